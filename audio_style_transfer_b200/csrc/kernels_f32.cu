// kernels_f32.cu -- CUDA-core fp32 kernels: the <=1e-4 accuracy tier and every non-GEMM stage.
// sm_100a only.  Reference lines are cited per kernel (paths relative to the upstream repo).
#include "common.cuh"
#include <algorithm>
#include <cooperative_groups.h>
#include <cstdlib>
#include "ptx.cuh"

// =============================================================================================
// helpers
// =============================================================================================
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// OR-reduce across each aligned group of 8 lanes
__device__ __forceinline__ uint32_t or8(uint32_t v) {
    v |= __shfl_xor_sync(0xffffffffu, v, 1);
    v |= __shfl_xor_sync(0xffffffffu, v, 2);
    v |= __shfl_xor_sync(0xffffffffu, v, 4);
    return v;
}

// =============================================================================================
// start conv (model.py:80-93): h0[t,c] = b0[c] + sum_k W0[k,0,c] * x[t+k-1] / 128, zero padded.
// one warp per row, lane owns channels 4*lane..4*lane+3.  Also emits the sign mask of h0.
// =============================================================================================
__global__ void __launch_bounds__(256) startconv_fwd_kernel(const float* __restrict__ x, const float* __restrict__ W0,
                                                            const float* __restrict__ b0, float* __restrict__ H0,
                                                            uint32_t* __restrict__ hmask, int T, int R) {
    const int lane = threadIdx.x & 31;
    const int warps_per_blk = blockDim.x >> 5;
    const float4 w0 = reinterpret_cast<const float4*>(W0)[lane];
    const float4 w1 = reinterpret_cast<const float4*>(W0 + AST_C)[lane];
    const float4 w2 = reinterpret_cast<const float4*>(W0 + 2 * AST_C)[lane];
    const float4 bb = reinterpret_cast<const float4*>(b0)[lane];
    for (int row = blockIdx.x * warps_per_blk + (threadIdx.x >> 5); row < R; row += gridDim.x * warps_per_blk) {
        const int t = row % T;
        const float xm = (t > 0) ? x[row - 1] * (1.0f / 128.0f) : 0.f;
        const float x0 = x[row] * (1.0f / 128.0f);
        const float xp = (t + 1 < T) ? x[row + 1] * (1.0f / 128.0f) : 0.f;
        float4 h;
        h.x = fmaf(w2.x, xp, fmaf(w1.x, x0, fmaf(w0.x, xm, bb.x)));
        h.y = fmaf(w2.y, xp, fmaf(w1.y, x0, fmaf(w0.y, xm, bb.y)));
        h.z = fmaf(w2.z, xp, fmaf(w1.z, x0, fmaf(w0.z, xm, bb.z)));
        h.w = fmaf(w2.w, xp, fmaf(w1.w, x0, fmaf(w0.w, xm, bb.w)));
        reinterpret_cast<float4*>(H0 + (size_t)row * AST_C)[lane] = h;
        uint32_t nib = (h.x > 0.f ? 1u : 0u) | (h.y > 0.f ? 2u : 0u) | (h.z > 0.f ? 4u : 0u) | (h.w > 0.f ? 8u : 0u);
        uint32_t wv = or8(nib << (4 * (lane & 7)));
        if ((lane & 7) == 0) hmask[(size_t)row * AST_MW + (lane >> 3)] = wv;
    }
}

void launch_startconv_fwd(const float* x, const float* W0, const float* b0, float* H0, uint32_t* hmask0, int T, int R,
                          cudaStream_t s) {
    int blocks = (R + 7) / 8;
    if (blocks > 148 * 8) blocks = 148 * 8;
    startconv_fwd_kernel<<<blocks, 256, 0, s>>>(x, W0, b0, H0, hmask0, T, R);
    ast_launch_counter += 1;
}

// adjoint of the start conv (SURVEY.md 8(a) last line):
//   p_k[t] = sum_c W0[k,0,c] g[t,c] ;  dx[t] = (p_0[t+1] + p_1[t] + p_2[t-1]) / 128
__global__ void __launch_bounds__(256) startconv_bwd_dots(const float* __restrict__ g, const float* __restrict__ W0,
                                                          float* __restrict__ p, int R) {
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    const float4 w0 = reinterpret_cast<const float4*>(W0)[lane];
    const float4 w1 = reinterpret_cast<const float4*>(W0 + AST_C)[lane];
    const float4 w2 = reinterpret_cast<const float4*>(W0 + 2 * AST_C)[lane];
    for (int row = blockIdx.x * wpb + (threadIdx.x >> 5); row < R; row += gridDim.x * wpb) {
        const float4 v = reinterpret_cast<const float4*>(g + (size_t)row * AST_C)[lane];
        float a0 = v.x * w0.x + v.y * w0.y + v.z * w0.z + v.w * w0.w;
        float a1 = v.x * w1.x + v.y * w1.y + v.z * w1.z + v.w * w1.w;
        float a2 = v.x * w2.x + v.y * w2.y + v.z * w2.z + v.w * w2.w;
        a0 = warp_sum(a0); a1 = warp_sum(a1); a2 = warp_sum(a2);
        if (lane == 0) { p[row] = a0; p[R + row] = a1; p[2 * (size_t)R + row] = a2; }
    }
}
__global__ void startconv_bwd_combine(const float* __restrict__ p, float* __restrict__ dx, int T, int R) {
    int row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= R) return;
    int t = row % T;
    float v = p[R + row];
    if (t + 1 < T) v += p[row + 1];
    if (t > 0) v += p[2 * (size_t)R + row - 1];
    dx[row] = v * (1.0f / 128.0f);
}
void launch_startconv_bwd_combine(const float* scratch3R, float* dx, int T, int R, cudaStream_t s) {
    startconv_bwd_combine<<<(R + 255) / 256, 256, 0, s>>>(scratch3R, dx, T, R);
    ast_launch_counter += 1;
}
void launch_startconv_bwd(const float* g, const float* W0, float* scratch3R, float* dx, int T, int R, cudaStream_t s) {
    int blocks = (R + 7) / 8;
    if (blocks > 148 * 8) blocks = 148 * 8;
    startconv_bwd_dots<<<blocks, 256, 0, s>>>(g, W0, scratch3R, R);
    ast_launch_counter += 1;
    startconv_bwd_combine<<<(R + 255) / 256, 256, 0, s>>>(scratch3R, dx, T, R);
    ast_launch_counter += 1;
}

// =============================================================================================
// layer kernel (fp32, CUDA cores): two chained GEMMs over a 64-row tile, 256 threads, 4x8 per thread.
//   forward  (model.py:99-114, masked.py:110-160 closed form):
//       z = b + sum_k relu(h)[t+(k-1)d] Wd[k] ; r = relu(z) ; h' = h + r Wr + br
//   backward:
//       g += (sum_k dz[t-(k-1)d] Wd[k]^T) * [h>0] + dF ; dz' = (g Wr'^T) * [z'>0]
// Rows outside [0,T) of the pair contribute zero: the reference's per-layer 'SAME' zero padding.
// =============================================================================================
#define LF_BM 64
#define LF_KC 32
#define LF_AS_LD 36
#define LF_RT_LD 132
#define LF_SMEM ((LF_BM * LF_AS_LD + LF_KC * AST_C + LF_BM * LF_RT_LD) * 4)

__device__ __forceinline__ void lf_load_w(float* Ws, const float* __restrict__ W, int tid) {
    // W points at a [32][128] fp32 chunk
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int idx = tid + i * 256;
        reinterpret_cast<float4*>(Ws)[idx] = reinterpret_cast<const float4*>(W)[idx];
    }
}
__device__ __forceinline__ void lf_mma(float (&acc)[4][8], const float* As, int lda, const float* Ws, int ty, int tx) {
#pragma unroll 8
    for (int kk = 0; kk < LF_KC; ++kk) {
        float a[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = As[(ty * 4 + i) * lda + kk];
        const float4 b0 = *reinterpret_cast<const float4*>(Ws + kk * AST_C + tx * 4);
        const float4 b1 = *reinterpret_cast<const float4*>(Ws + kk * AST_C + 64 + tx * 4);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            acc[i][0] = fmaf(a[i], b0.x, acc[i][0]); acc[i][1] = fmaf(a[i], b0.y, acc[i][1]);
            acc[i][2] = fmaf(a[i], b0.z, acc[i][2]); acc[i][3] = fmaf(a[i], b0.w, acc[i][3]);
            acc[i][4] = fmaf(a[i], b1.x, acc[i][4]); acc[i][5] = fmaf(a[i], b1.y, acc[i][5]);
            acc[i][6] = fmaf(a[i], b1.z, acc[i][6]); acc[i][7] = fmaf(a[i], b1.w, acc[i][7]);
        }
    }
}
// sign mask of a thread tile row -> 4 words of the row; lanes tx&7==0 write
__device__ __forceinline__ void lf_store_mask(const float (&v)[8], uint32_t* __restrict__ mrow, int tx) {
    uint32_t lo = (v[0] > 0.f ? 1u : 0u) | (v[1] > 0.f ? 2u : 0u) | (v[2] > 0.f ? 4u : 0u) | (v[3] > 0.f ? 8u : 0u);
    uint32_t hi = (v[4] > 0.f ? 1u : 0u) | (v[5] > 0.f ? 2u : 0u) | (v[6] > 0.f ? 4u : 0u) | (v[7] > 0.f ? 8u : 0u);
    lo = or8(lo << (4 * (tx & 7)));
    hi = or8(hi << (4 * (tx & 7)));
    if ((tx & 7) == 0) { mrow[tx >> 3] = lo; mrow[2 + (tx >> 3)] = hi; }
}
__device__ __forceinline__ void lf_apply_mask(float (&v)[8], const uint32_t* __restrict__ mrow, int tx) {
    const uint32_t lo = mrow[tx >> 3] >> (4 * (tx & 7));
    const uint32_t hi = mrow[2 + (tx >> 3)] >> (4 * (tx & 7));
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        if (!((lo >> j) & 1u)) v[j] = 0.f;
        if (!((hi >> j) & 1u)) v[4 + j] = 0.f;
    }
}

template <bool BWD>
__global__ void __launch_bounds__(256) layer_f32_kernel(const LayerArgsF32 a) {
    extern __shared__ __align__(16) float lf_smem[];
    float* As = lf_smem;
    float* Ws = As + LF_BM * LF_AS_LD;
    float* Rt = Ws + LF_KC * AST_C;
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int row0 = blockIdx.x * LF_BM;
    const int t0 = row0 % a.T;
    const int pbase = row0 - t0;

    float acc[4][8];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

    if (a.A1 != nullptr) {
        for (int k = 0; k < 3; ++k) {
            const int sh = (k - 1) * a.shift;
            if (t0 + sh + LF_BM <= 0 || t0 + sh >= a.T) continue;      // whole tap tile is padding (block-uniform)
            for (int kc = 0; kc < AST_C / LF_KC; ++kc) {
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const int idx = tid + i * 256, r = idx >> 3, q = idx & 7;
                    const int t = t0 + r + sh;
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (t >= 0 && t < a.T)
                        v = *reinterpret_cast<const float4*>(a.A1 + (size_t)(pbase + t) * AST_C + kc * LF_KC + q * 4);
                    if (!BWD) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
                    *reinterpret_cast<float4*>(As + r * LF_AS_LD + q * 4) = v;
                }
                lf_load_w(Ws, a.W1 + ((size_t)k * AST_C + kc * LF_KC) * AST_C, tid);
                __syncthreads();
                lf_mma(acc, As, LF_AS_LD, Ws, ty, tx);
                __syncthreads();
            }
        }
    }

    // ---- epilogue 1 -> Rt (A operand of GEMM2) ------------------------------------------------
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int row = row0 + ty * 4 + i;
        float v[8];
        if (!BWD) {
            const float4 b0 = *reinterpret_cast<const float4*>(a.bias1 + tx * 4);
            const float4 b1 = *reinterpret_cast<const float4*>(a.bias1 + 64 + tx * 4);
            v[0] = acc[i][0] + b0.x; v[1] = acc[i][1] + b0.y; v[2] = acc[i][2] + b0.z; v[3] = acc[i][3] + b0.w;
            v[4] = acc[i][4] + b1.x; v[5] = acc[i][5] + b1.y; v[6] = acc[i][6] + b1.z; v[7] = acc[i][7] + b1.w;
            lf_store_mask(v, a.zmask_out + (size_t)row * AST_MW, tx);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j], 0.f);
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = acc[i][j];
            if (a.A1 != nullptr) lf_apply_mask(v, a.hmask_in + (size_t)row * AST_MW, tx);
            float* grow = a.g + (size_t)row * AST_C;
            if (!a.g_init) {
                const float4 g0 = *reinterpret_cast<const float4*>(grow + tx * 4);
                const float4 g1 = *reinterpret_cast<const float4*>(grow + 64 + tx * 4);
                v[0] += g0.x; v[1] += g0.y; v[2] += g0.z; v[3] += g0.w;
                v[4] += g1.x; v[5] += g1.y; v[6] += g1.z; v[7] += g1.w;
            }
            if (a.dF != nullptr) {
                const float4 f0 = *reinterpret_cast<const float4*>(a.dF + (size_t)row * AST_C + tx * 4);
                const float4 f1 = *reinterpret_cast<const float4*>(a.dF + (size_t)row * AST_C + 64 + tx * 4);
                v[0] += f0.x; v[1] += f0.y; v[2] += f0.z; v[3] += f0.w;
                v[4] += f1.x; v[5] += f1.y; v[6] += f1.z; v[7] += f1.w;
            }
            *reinterpret_cast<float4*>(grow + tx * 4) = make_float4(v[0], v[1], v[2], v[3]);
            *reinterpret_cast<float4*>(grow + 64 + tx * 4) = make_float4(v[4], v[5], v[6], v[7]);
        }
        *reinterpret_cast<float4*>(Rt + (ty * 4 + i) * LF_RT_LD + tx * 4) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(Rt + (ty * 4 + i) * LF_RT_LD + 64 + tx * 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
    if (a.W2 == nullptr) return;
    __syncthreads();

    // ---- GEMM2: [64 x 128] . W2[128 x 128] ----------------------------------------------------
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    for (int kc = 0; kc < AST_C / LF_KC; ++kc) {
        lf_load_w(Ws, a.W2 + (size_t)kc * LF_KC * AST_C, tid);
        __syncthreads();
        lf_mma(acc, Rt + kc * LF_KC, LF_RT_LD, Ws, ty, tx);
        __syncthreads();
    }

    // ---- epilogue 2 ---------------------------------------------------------------------------
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int row = row0 + ty * 4 + i;
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = acc[i][j];
        if (!BWD) {
            const float4 b0 = *reinterpret_cast<const float4*>(a.bias2 + tx * 4);
            const float4 b1 = *reinterpret_cast<const float4*>(a.bias2 + 64 + tx * 4);
            const float4 h0 = *reinterpret_cast<const float4*>(a.Hin + (size_t)row * AST_C + tx * 4);
            const float4 h1 = *reinterpret_cast<const float4*>(a.Hin + (size_t)row * AST_C + 64 + tx * 4);
            // h' = h + (r Wr + br): same association as the reference (model.py:109-114)
            v[0] = h0.x + (v[0] + b0.x); v[1] = h0.y + (v[1] + b0.y); v[2] = h0.z + (v[2] + b0.z); v[3] = h0.w + (v[3] + b0.w);
            v[4] = h1.x + (v[4] + b1.x); v[5] = h1.y + (v[5] + b1.y); v[6] = h1.z + (v[6] + b1.z); v[7] = h1.w + (v[7] + b1.w);
            *reinterpret_cast<float4*>(a.Hout + (size_t)row * AST_C + tx * 4) = make_float4(v[0], v[1], v[2], v[3]);
            *reinterpret_cast<float4*>(a.Hout + (size_t)row * AST_C + 64 + tx * 4) = make_float4(v[4], v[5], v[6], v[7]);
            lf_store_mask(v, a.hmask_out + (size_t)row * AST_MW, tx);
        } else {
            lf_apply_mask(v, a.zmask_in + (size_t)row * AST_MW, tx);
            *reinterpret_cast<float4*>(a.dz_out + (size_t)row * AST_C + tx * 4) = make_float4(v[0], v[1], v[2], v[3]);
            *reinterpret_cast<float4*>(a.dz_out + (size_t)row * AST_C + 64 + tx * 4) = make_float4(v[4], v[5], v[6], v[7]);
        }
    }
}

int setup_heads_kernel();
int setup_f32_kernels() {
    cudaError_t e = cudaFuncSetAttribute(layer_f32_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, LF_SMEM);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(layer_f32_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, LF_SMEM);
    if (e != cudaSuccess) return (int)e;
    return setup_heads_kernel();
}
void launch_layer_f32(const LayerArgsF32& a, bool bwd, cudaStream_t s) {
    const int blocks = a.R / LF_BM;
    if (bwd) layer_f32_kernel<true><<<blocks, 256, LF_SMEM, s>>>(a);
    else     layer_f32_kernel<false><<<blocks, 256, LF_SMEM, s>>>(a);
    ast_launch_counter += 1;
}

// =============================================================================================
// Gram partials (methods.py:62-73).  Deterministic: every CTA writes its own partial, a second
// kernel sums the partials in a fixed order (no float atomics).
// =============================================================================================
// "ours": G[c,l,m] = sum_t F_l[t,c] F_m[t,c].  grid = (nchunk, 4 channel quads, B); part[b][chunk][c][l][m].
// A CTA owns 32 channels of one row chunk.  R rows of all L extracts are staged per round into shared memory
// with cp.async (3 stages, one barrier per round), so every element of F is requested from L2/HBM exactly once;
// block = 32 channels x ng(ng+1)/2 groups, each thread an 8x8 (l,m) register block of the UPPER triangle of
// 8-layer groups (the Gram is symmetric; the mirror is written at the end).  Layers L..8ng-1 are zero rows in
// shared memory, rows past the chunk end are zero-filled by cp.async (src-size 0).  Rows are accumulated in
// ascending order -> run-to-run reproducible.
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
    const int sz = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

#define GO_STAGES 3
template <int R>
__global__ void __launch_bounds__(320, 2) gram_ours_kernel(const float* const* __restrict__ F, int L, int T, int own_lo,
                                                        int own_hi, float* __restrict__ part, int nchunk,
                                                        int chunk_rows, int ng) {
    extern __shared__ __align__(16) float go_s[];          // [stage][8 ng][R][32]
    const int stage_f = ng * 8 * R * 32;
    const int cl = threadIdx.x, grp = threadIdx.y;
    const int nthr = blockDim.x * blockDim.y, tid = grp * 32 + cl;
    int lg = 0, mg = grp;                                   // triangular index -> (lg <= mg)
    for (int w = ng; mg >= w; --w) { mg -= w; ++lg; }
    mg += lg;
    const int c0 = blockIdx.y * 32, b = blockIdx.z;
    const int tb = blockIdx.x * chunk_rows;
    const int t_lo = max(tb, own_lo), t_hi = min(min(tb + chunk_rows, T), own_hi);
    for (int i = tid; i < GO_STAGES * stage_f; i += nthr)
        if ((i % stage_f) / (R * 32) >= L) go_s[i] = 0.f;
    const int nround = t_hi > t_lo ? (t_hi - t_lo + R - 1) / R : 0;
    auto issue = [&](int round) {
        if (round < nround) {
            float* dst = go_s + (round % GO_STAGES) * stage_f;
            const int t0 = t_lo + round * R;
            for (int i = tid; i < L * R * 8; i += nthr) {
                const int q = i & 7, r = (i >> 3) % R, l = i / (8 * R);
                const bool ok = t0 + r < t_hi;
                const float* src = F[l] + ((size_t)b * T + (ok ? t0 + r : t_lo)) * AST_C + c0 + q * 4;
                cp_async16(smem_u32(dst + (l * R + r) * 32 + q * 4), src, ok);
            }
        }
        cp_async_commit();
    };
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    issue(0);
    issue(1);
    for (int round = 0; round < nround; ++round) {
        cp_async_wait<1>();                 // this thread's copies of `round` have landed
        __syncthreads();                    // everyone's have; and everyone is done reading stage (round+2)%3
        issue(round + 2);
        const float* sl = go_s + (round % GO_STAGES) * stage_f + (lg * 8 * R) * 32 + cl;
        const float* sm = go_s + (round % GO_STAGES) * stage_f + (mg * 8 * R) * 32 + cl;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            float vl[8], vm[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) { vl[i] = sl[(i * R + r) * 32]; vm[i] = sm[(i * R + r) * 32]; }
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(vl[i], vm[j], acc[i][j]);
        }
    }
    cp_async_wait<0>();
    float* out = part + (((size_t)b * nchunk + blockIdx.x) * AST_C + c0 + cl) * L * L;
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int l = lg * 8 + i, m = mg * 8 + j;
            if (l < L && m < L) {
                out[l * L + m] = acc[i][j];
                if (lg != mg) out[m * L + l] = acc[i][j];
            }
        }
}
// Few style layers (L <= GS_MAXL: one stack, the reference's --stack 0/1/2): the blocked kernel above computes
// 3 x 64 products per (t, c) of which 55 are needed at L = 10 and is issue-bound (48 us for 84 MB at cfg0).  Here a
// thread owns ONE channel and ALL L(L+1)/2 pairs in registers (L loads + L(L+1)/2 FMAs per row) and 8 row groups
// share the channel; same cp.async staging (every byte of F requested once), 4 stages of 16 rows in flight.  The 8 row
// groups are summed through shared memory in a fixed order at the end -> run-to-run reproducible.
#define GS_MAXL 12
#define GS_R 16
#define GS_STAGES 4
#define GS_GROUPS 8
template <int L>
__global__ void __launch_bounds__(32 * GS_GROUPS) gram_ours_small_kernel(const float* const* __restrict__ F, int T, int own_lo,
                                                                         int own_hi, float* __restrict__ part, int nchunk,
                                                                         int chunk_rows) {
    extern __shared__ __align__(16) float go_s[];          // [stage][L][GS_R][32]; reused as [group][pair][32] at the end
    constexpr int NP = L * (L + 1) / 2;
    constexpr int stage_f = L * GS_R * 32;
    const int cl = threadIdx.x, grp = threadIdx.y;
    const int tid = grp * 32 + cl;
    constexpr int nthr = 32 * GS_GROUPS;
    const int c0 = blockIdx.y * 32, b = blockIdx.z;
    const int tb = blockIdx.x * chunk_rows;
    const int t_lo = max(tb, own_lo), t_hi = min(min(tb + chunk_rows, T), own_hi);
    const int nround = t_hi > t_lo ? (t_hi - t_lo + GS_R - 1) / GS_R : 0;
    auto issue = [&](int round) {
        if (round < nround) {
            float* dst = go_s + (round % GS_STAGES) * stage_f;
            const int t0 = t_lo + round * GS_R;
            for (int i = tid; i < L * GS_R * 8; i += nthr) {
                const int q = i & 7, r = (i >> 3) % GS_R, l = i / (8 * GS_R);
                const bool ok = t0 + r < t_hi;
                const float* src = F[l] + ((size_t)b * T + (ok ? t0 + r : t_lo)) * AST_C + c0 + q * 4;
                cp_async16(smem_u32(dst + (l * GS_R + r) * 32 + q * 4), src, ok);
            }
        }
        cp_async_commit();
    };
    float acc[NP];
#pragma unroll
    for (int i = 0; i < NP; ++i) acc[i] = 0.f;
#pragma unroll
    for (int st = 0; st < GS_STAGES - 1; ++st) issue(st);
    for (int round = 0; round < nround; ++round) {
        cp_async_wait<GS_STAGES - 2>();     // this thread's copies of `round` have landed
        __syncthreads();                    // everyone's have; and everyone is done reading the stage refilled next
        issue(round + GS_STAGES - 1);
        const float* sr = go_s + (round % GS_STAGES) * stage_f + cl;
#pragma unroll
        for (int rr = 0; rr < GS_R / GS_GROUPS; ++rr) {
            const int r = grp + rr * GS_GROUPS;
            float v[L];
#pragma unroll
            for (int l = 0; l < L; ++l) v[l] = sr[(l * GS_R + r) * 32];
            int p = 0;
#pragma unroll
            for (int l = 0; l < L; ++l)
#pragma unroll
                for (int m = l; m < L; ++m, ++p) acc[p] = fmaf(v[l], v[m], acc[p]);
        }
    }
    cp_async_wait<0>();
    __syncthreads();
    float* red = go_s;                                     // [GS_GROUPS][NP][32] <= the staging area
#pragma unroll
    for (int p = 0; p < NP; ++p) red[(grp * NP + p) * 32 + cl] = acc[p];
    __syncthreads();
    float* out = part + (((size_t)b * nchunk + blockIdx.x) * AST_C + c0 + cl) * L * L;
    for (int p = grp; p < NP; p += GS_GROUPS) {
        float a = red[p * 32 + cl];
#pragma unroll
        for (int g = 1; g < GS_GROUPS; ++g) a += red[(g * NP + p) * 32 + cl];
        int l = 0, rem = p;                                 // pair index -> (l <= m)
        while (rem >= L - l) { rem -= L - l; ++l; }
        const int m = l + rem;
        out[l * L + m] = a;
        out[m * L + l] = a;
    }
}
template <int L>
static void launch_gos(const float* const* d_F, int T, int B, int own_lo, int own_hi, float* part, int nchunk,
                       int chunk_rows, cudaStream_t s) {
    constexpr size_t smem = (size_t)GS_STAGES * L * GS_R * 32 * sizeof(float);
    static_assert((size_t)GS_GROUPS * (L * (L + 1) / 2) * 32 * sizeof(float) <= smem || L <= 2, "reduction scratch must fit the staging area");
    static bool once = false;
    if (!once) {
        cudaFuncSetAttribute(gram_ours_small_kernel<L>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max(smem, (size_t)GS_GROUPS * 3 * 32 * 4));
        once = true;
    }
    dim3 grid(nchunk, AST_C / 32, B), block(32, GS_GROUPS);
    gram_ours_small_kernel<L><<<grid, block, std::max(smem, (size_t)GS_GROUPS * 3 * 32 * 4), s>>>(d_F, T, own_lo, own_hi, part, nchunk, chunk_rows);
}
int gram_ours_max_small_L() { return GS_MAXL; }
void launch_gram_ours(const float* const* d_F, int L, int T, int B, int own_lo, int own_hi, float* part, int nchunk,
                      int chunk_rows, cudaStream_t s) {
    if (L <= GS_MAXL) {
#define GOS_CASE(n) case n: launch_gos<n>(d_F, T, B, own_lo, own_hi, part, nchunk, chunk_rows, s); break;
        switch (L) {
            GOS_CASE(1) GOS_CASE(2) GOS_CASE(3) GOS_CASE(4) GOS_CASE(5) GOS_CASE(6) GOS_CASE(7) GOS_CASE(8)
            GOS_CASE(9) GOS_CASE(10) GOS_CASE(11) GOS_CASE(12)
            default: break;
        }
#undef GOS_CASE
        ast_launch_counter += 1;
        return;
    }
    const int ng = (L + 7) / 8;
    dim3 grid(nchunk, AST_C / 32, B), block(32, ng * (ng + 1) / 2);
    if (ng <= 2) {
        const size_t smem = (size_t)GO_STAGES * ng * 8 * 8 * 32 * sizeof(float);      // <= 48 KB
        gram_ours_kernel<8><<<grid, block, smem, s>>>(d_F, L, T, own_lo, own_hi, part, nchunk, chunk_rows, ng);
    } else {
        // 8 rows per round (96 KB at ng = 4, opt-in) and two CTAs per SM: ncu had this kernel at ONE 10-warp CTA per SM
        // (100 registers x 320 threads, 64 KB carve-out) and 4 rows per barrier: 235 us for 252 MB at L = 30
        const size_t smem = (size_t)GO_STAGES * ng * 8 * 8 * 32 * sizeof(float);
        static bool once = false;
        if (!once) {
            cudaFuncSetAttribute(gram_ours_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 4 * 8 * 8 * 32 * 4);
            cudaFuncSetAttribute(gram_ours_kernel<8>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
            once = true;
        }
        gram_ours_kernel<8><<<grid, block, smem, s>>>(d_F, L, T, own_lo, own_hi, part, nchunk, chunk_rows, ng);
    }
    ast_launch_counter += 1;
}

// gatys: G_l = F_l^T F_l (128x128).  256 threads, 8x8 per thread (rows ty*4.., 64+ty*4.. ; cols likewise).
// grid = (nchunk, L, B).  part[(b*L+l)][chunk][128][128].
__global__ void __launch_bounds__(256) gram_gatys_kernel(const float* const* __restrict__ F, int L, int T, int own_lo,
                                                         int own_hi, float* __restrict__ part, int nchunk,
                                                         int chunk_rows) {
    __shared__ __align__(16) float Fs[32][AST_C];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int l = blockIdx.y, b = blockIdx.z;
    const float* Fl = F[l] + (size_t)b * T * AST_C;
    const int tb = blockIdx.x * chunk_rows;
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    for (int ts = tb; ts < min(tb + chunk_rows, T); ts += 32) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int idx = tid + i * 256, r = idx >> 5, q = idx & 31;
            const int t = ts + r;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (t < T && t >= own_lo && t < own_hi) v = *reinterpret_cast<const float4*>(Fl + (size_t)t * AST_C + q * 4);
            *reinterpret_cast<float4*>(&Fs[r][q * 4]) = v;
        }
        __syncthreads();
#pragma unroll 4
        for (int r = 0; r < 32; ++r) {
            const float4 a0 = *reinterpret_cast<const float4*>(&Fs[r][ty * 4]);
            const float4 a1 = *reinterpret_cast<const float4*>(&Fs[r][64 + ty * 4]);
            const float4 b0 = *reinterpret_cast<const float4*>(&Fs[r][tx * 4]);
            const float4 b1 = *reinterpret_cast<const float4*>(&Fs[r][64 + tx * 4]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
    float* out = part + (((size_t)b * L + l) * nchunk + blockIdx.x) * AST_C * AST_C;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int row = (i < 4) ? ty * 4 + i : 64 + ty * 4 + (i - 4);
        *reinterpret_cast<float4*>(out + row * AST_C + tx * 4) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        *reinterpret_cast<float4*>(out + row * AST_C + 64 + tx * 4) = make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]);
    }
}
void launch_gram_gatys(const float* const* d_F, int L, int T, int B, int own_lo, int own_hi, float* part, int nchunk,
                       int chunk_rows, cudaStream_t s) {
    dim3 grid(nchunk, L, B);
    gram_gatys_kernel<<<grid, 256, 0, s>>>(d_F, L, T, own_lo, own_hi, part, nchunk, chunk_rows);
    ast_launch_counter += 1;
}

// G[o][e] = sum_{ch} part[o][ch][e], ch ascending (fixed order), double accumulate.
__global__ void gram_reduce_kernel(const float* __restrict__ part, float* __restrict__ G, int nchunk, int E) {
    const int o = blockIdx.y;
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < E; e += gridDim.x * blockDim.x) {
        const float* p = part + (size_t)o * nchunk * E + e;
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;          // fixed association: run-to-run reproducible
        int ch = 0;
        for (; ch + 4 <= nchunk; ch += 4) {
            s0 += (double)p[(size_t)ch * E]; s1 += (double)p[(size_t)(ch + 1) * E];
            s2 += (double)p[(size_t)(ch + 2) * E]; s3 += (double)p[(size_t)(ch + 3) * E];
        }
        for (; ch < nchunk; ++ch) s0 += (double)p[(size_t)ch * E];
        G[(size_t)o * E + e] = (float)((s0 + s1) + (s2 + s3));
    }
}
// many chunks (channel-wise Gram, 128 of them): block = 32 elements x 8 chunk slices; slice k sums chunks
// k, k+8, ... in two interleaved chains, then the 8 slice sums are combined in fixed order through smem.
__global__ void __launch_bounds__(256) gram_reduce_sliced_kernel(const float* __restrict__ part, float* __restrict__ G,
                                                                 int nchunk, int E) {
    __shared__ double red[8][32];
    const int o = blockIdx.y, k = threadIdx.y;
    const int e = blockIdx.x * 32 + threadIdx.x;
    double s0 = 0.0, s1 = 0.0;
    if (e < E) {
        const float* p = part + (size_t)o * nchunk * E + e;
        int ch = k;
        for (; ch + 8 < nchunk; ch += 16) { s0 += (double)p[(size_t)ch * E]; s1 += (double)p[(size_t)(ch + 8) * E]; }
        if (ch < nchunk) s0 += (double)p[(size_t)ch * E];
    }
    red[k][threadIdx.x] = s0 + s1;
    __syncthreads();
    if (k == 0 && e < E) {
        double t = 0.0;
#pragma unroll
        for (int i = 0; i < 8; ++i) t += red[i][threadIdx.x];
        G[(size_t)o * E + e] = (float)t;
    }
}
void launch_gram_reduce(const float* part, float* G, int gatys, int L, int B, int nchunk, cudaStream_t s) {
    const int nouter = gatys ? B * L : B;
    const int E = gatys ? AST_C * AST_C : AST_C * L * L;
    if (nchunk >= 32) {
        dim3 grid((E + 31) / 32, nouter), block(32, 8);
        gram_reduce_sliced_kernel<<<grid, block, 0, s>>>(part, G, nchunk, E);
    } else {
        dim3 grid((E + 255) / 256, nouter);
        gram_reduce_kernel<<<grid, 256, 0, s>>>(part, G, nchunk, E);
    }
    ast_launch_counter += 1;
}

// =============================================================================================
// content head (methods.py:58,116-117): squared error of Ec - Pc over the owned rows.
// grid = (nblk, B); part[b][blk] (double), then a fixed-order final sum -> sse[b].
// =============================================================================================
__global__ void __launch_bounds__(256) content_sse_kernel(const float* const* __restrict__ Hc, int n_cont, int cnt_ch,
                                                          const float* __restrict__ Pc, int T, int own_lo, int own_hi,
                                                          double* __restrict__ part) {
    __shared__ double red[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int b = blockIdx.y;
    const int ncols = n_cont * cnt_ch;
    double s = 0.0;
    const bool vec = (cnt_ch & 3) == 0;               // 16-byte loads (Pc rows stay 16-byte aligned: ncols % 4 == 0)
    for (int t = own_lo + blockIdx.x * 8 + warp; t < own_hi; t += gridDim.x * 8) {
        const size_t row = (size_t)b * T + t;
        for (int slot = 0; slot < n_cont; ++slot) {
            const float* h = Hc[slot] + row * AST_C;
            const float* p = Pc + row * ncols + (size_t)slot * cnt_ch;
            if (vec) {
                for (int j = lane * 4; j < cnt_ch; j += 128) {
                    const float4 a = *reinterpret_cast<const float4*>(h + j), c = *reinterpret_cast<const float4*>(p + j);
                    const float d0 = a.x - c.x, d1 = a.y - c.y, d2 = a.z - c.z, d3 = a.w - c.w;
                    s += ((double)d0 * (double)d0 + (double)d1 * (double)d1) + ((double)d2 * (double)d2 + (double)d3 * (double)d3);
                }
            } else {
                for (int j = lane; j < cnt_ch; j += 32) {
                    const float d = h[j] - p[j];
                    s += (double)d * (double)d;
                }
            }
        }
    }
    s = warp_sum_d(s);
    if (lane == 0) red[warp] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double tot = 0.0;
        for (int i = 0; i < 8; ++i) tot += red[i];
        part[(size_t)b * gridDim.x + blockIdx.x] = tot;
    }
}
__global__ void content_sse_final(const double* __restrict__ part, int nblk, float* __restrict__ sse) {
    double tot = 0.0;                           // one warp; lane-strided then a fixed shuffle tree: deterministic
    for (int i = threadIdx.x; i < nblk; i += 32) tot += part[(size_t)blockIdx.x * nblk + i];
    tot = warp_sum_d(tot);
    if (threadIdx.x == 0) sse[blockIdx.x] = (float)tot;
}
void launch_content_sse(const float* const* d_Hc, int n_cont, int cnt_ch, const float* Pc, int T, int B, int own_lo,
                        int own_hi, float* part, int nblk, float* sse, cudaStream_t s) {
    dim3 grid(nblk, B);
    content_sse_kernel<<<grid, 256, 0, s>>>(d_Hc, n_cont, cnt_ch, Pc, T, own_lo, own_hi, reinterpret_cast<double*>(part));
    ast_launch_counter += 1;
    if (!sse) return;
    content_sse_final<<<B, 32, 0, s>>>(reinterpret_cast<const double*>(part), nblk, sse);
    ast_launch_counter += 1;
}

// =============================================================================================
// heads: Frobenius normalise (methods.py:74), style MSE (methods.py:118-119) and its gradient
// through the normalisation:  dG = (dGn - Gn <dGn,Gn>) * rsqrt(max(|G|^2, 1e-12)).
// grid = (nmat, B), one CTA per matrix.
// =============================================================================================
__device__ __forceinline__ double block_sum_d(double v, double* red) {
    v = warp_sum_d(v);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    double t = 0.0;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += red[i];
    return t;
}
// sl = sum of the per-matrix squared differences, sc = content squared error
__device__ __forceinline__ void write_losses(const HeadArgs& a, int b, double sl, double sc) {
    const double Ls = (double)a.ls_scale * sl;
    const double Lc = (double)a.lc_scale * sc;
    // methods.py:125.  The reference also evaluates the regulariser when gamma = 0 (logging only); here it is 0 then.
    const double reg = a.gamma != 0.f ? (double)a.reg_scale * (double)a.reg_raw[b] : 0.0;
    a.losses[b * 4 + 0] = (float)(Lc + (double)a.lambd * Ls + (double)a.gamma * reg);
    a.losses[b * 4 + 1] = (float)Lc;
    a.losses[b * 4 + 2] = (float)Ls;
    a.losses[b * 4 + 3] = (float)reg;
}
__global__ void __launch_bounds__(1024) heads_kernel(const HeadArgs a) {
    extern __shared__ float dgs[];            // [n][n + 1]: dG, read back transposed for Ssym = dG + dG^T
    __shared__ double red[32];
    __shared__ double red2[2][32];
    __shared__ int s_last;
    const int mat = blockIdx.x, b = blockIdx.y;
    const int n = a.n, msize = n * n;
    const float* G = a.G + ((size_t)b * a.nmat + mat) * msize;
    const bool used = mat < a.nmat_used;
    const float* Ps = used ? a.Ps + ((size_t)b * a.nmat_used + mat) * msize : nullptr;
    // The kernel is a chain of block reductions, i.e. latency: G and Ps are loaded ONCE into registers (<= 16 elements per
    // thread: msize <= 16 blockDim for both head types), and the two sums of the second pass share one reduction.
    // The style loss keeps the direct sum of (gn - Ps)^2: the algebraic one-pass form cancels when gn is close to Ps.
    // gram_part != NULL (single-GPU step, Gatys layout [b][l][chunk][n*n]): the split-K partials are summed here, in
    // gram_reduce_kernel's order, instead of by a kernel of their own.
    float gr[16], pr[16];
    double ss = 0.0;
    if (a.gram_part) {
        // chunk-outer so that 16 independent loads are in flight per iteration (a per-element chunk loop serialises 16
        // round trips); ascending chunk order, fp32 adds of the (few) fp32 chunk sums
        const float* p = a.gram_part + ((size_t)b * a.nmat + mat) * a.nchunk * msize;
#pragma unroll
        for (int k = 0; k < 16; ++k) gr[k] = 0.f;
        for (int ch = 0; ch < a.nchunk; ++ch, p += msize) {
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                const int e = threadIdx.x + k * blockDim.x;
                if (e < msize) gr[k] += p[e];
            }
        }
    } else {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const int e = threadIdx.x + k * blockDim.x;
            gr[k] = e < msize ? G[e] : 0.f;
        }
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int e = threadIdx.x + k * blockDim.x;
        pr[k] = (used && e < msize) ? Ps[e] : 0.f;
        ss += (double)gr[k] * (double)gr[k];
    }
    ss = block_sum_d(ss, red);
    const double inv = 1.0 / sqrt(fmax(ss, 1e-12));
    double ls = 0.0, dot = 0.0;
    if (used) {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const double gn = (double)gr[k] * inv;
            const double d = gn - (double)pr[k];
            ls += d * d;
            dot += (double)a.coef_dGn * d * gn;
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (int)(blockDim.x >> 5);
    {
        ls = warp_sum_d(ls); dot = warp_sum_d(dot);
        if (lane == 0) { red2[0][warp] = ls; red2[1][warp] = dot; }
        __syncthreads();
        ls = 0.0; dot = 0.0;
        for (int i = 0; i < nw; ++i) { ls += red2[0][i]; dot += red2[1][i]; }      // same order in every thread
    }
    // ---- losses by the last CTA of this pair to get here (fixed-order sums: the value does not depend on which) ----
    if (a.done) {
        if (threadIdx.x == 0) {
            a.ls_part[(size_t)b * a.nmat + mat] = (float)ls;
            __threadfence();
            const unsigned prev = atomicAdd(a.done + b, 1u);
            s_last = (prev == (unsigned)a.nmat - 1u);
        }
        __syncthreads();
        if (s_last) {                                           // block-uniform
            __threadfence();
            double sl = 0.0, sc = 0.0;
            for (int m = threadIdx.x; m < a.nmat_used; m += blockDim.x) sl += (double)__ldcg(a.ls_part + (size_t)b * a.nmat + m);
            if (a.sse_part)
                for (int i = threadIdx.x; i < a.sse_nblk; i += blockDim.x) sc += __ldcg(a.sse_part + (size_t)b * a.sse_nblk + i);
            sl = block_sum_d(sl, red);
            sc = a.sse_part ? block_sum_d(sc, red) : (double)a.sse[b];
            if (threadIdx.x == 0) {
                write_losses(a, b, sl, sc);
                a.done[b] = 0u;                                 // ready for the next launch (stream order)
            }
        }
    } else if (threadIdx.x == 0) {
        a.ls_part[(size_t)b * a.nmat + mat] = (float)ls;
    }
    const bool clamped = ss < 1e-12;
    float* Gn = a.Gn ? a.Gn + ((size_t)b * a.nmat + mat) * msize : nullptr;
    float* Ss = a.Ssym ? a.Ssym + ((size_t)b * a.nmat + mat) * msize : nullptr;
    __nv_bfloat16* Sb = a.Ssym_bf16 ? a.Ssym_bf16 + ((size_t)b * a.nmat + mat) * msize : nullptr;
    __nv_bfloat16* Sbl = a.Ssym_bf16_lo ? a.Ssym_bf16_lo + ((size_t)b * a.nmat + mat) * msize : nullptr;
    const int ld = n + 1;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int e = threadIdx.x + k * blockDim.x;
        if (e >= msize) break;
        const int i = e / n, j = e - i * n;
        const double gn = (double)gr[k] * inv;
        if (Gn) Gn[e] = (float)gn;
        if (Ss) {
            double dg = 0.0;
            if (used) {
                const double dgn = (double)a.coef_dGn * (gn - (double)pr[k]);
                dg = clamped ? dgn * inv : (dgn - gn * dot) * inv;
            }
            dgs[i * ld + j] = (float)dg;
        }
    }
    if (!Ss) return;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int e = threadIdx.x + k * blockDim.x;
        if (e >= msize) break;
        const int i = e / n, j = e - i * n;
        const float v = dgs[i * ld + j] + dgs[j * ld + i];
        Ss[e] = v;
        if (Sb) Sb[e] = __float2bfloat16_rn(v);
        if (Sbl) Sbl[e] = __float2bfloat16_rn(v - __bfloat162float(__float2bfloat16_rn(v)));
    }
}
// The same head for the Gatys matrices (n = 128) with a CLUSTER of 4 CTAs per matrix: 30 one-CTA launches of the kernel
// above leave 118 SMs idle and chain ~8 dependent memory round trips per CTA (27 us under ncu).  Here a CTA owns 32 rows
// (8 elements per thread, every load of a phase in flight at once), the three sums go through DSMEM + one cluster
// barrier each, and the transposed term of Ssym = dG + dG^T is read from the peer CTA's shared memory.
#define HG_CL 4
#define HG_THREADS 512
#define HG_E 8                                  // elements per thread: 32 rows x 128 / 512
__device__ __forceinline__ void hg_cluster_sum(double* v, int n, double (*wred)[2], double (*slot)[HG_CL][2], int& par) {
    namespace cgx = cooperative_groups;
    cgx::cluster_group cl = cgx::this_cluster();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int j = 0; j < n; ++j) v[j] = warp_sum_d(v[j]);
    if (lane == 0) for (int j = 0; j < n; ++j) wred[warp][j] = v[j];
    __syncthreads();
    if ((int)threadIdx.x < n) {
        double t = 0.0;
        for (int w = 0; w < HG_THREADS / 32; ++w) t += wred[w][threadIdx.x];
        const unsigned me = cl.block_rank();
        for (unsigned r = 0; r < HG_CL; ++r) *cl.map_shared_rank(&slot[par][me][threadIdx.x], r) = t;
    }
    cl.sync();
    for (int j = 0; j < n; ++j) {
        double t = 0.0;
        for (int r = 0; r < HG_CL; ++r) t += slot[par][r][j];
        v[j] = t;
    }
    par ^= 1;
}
__global__ void __cluster_dims__(HG_CL, 1, 1) __launch_bounds__(HG_THREADS) heads_gatys_cluster_kernel(const HeadArgs a) {
    namespace cgx = cooperative_groups;
    cgx::cluster_group cl = cgx::this_cluster();
    __shared__ float dgs[32][AST_C + 1];       // this CTA's 32 rows of dG
    __shared__ double wred[HG_THREADS / 32][2];
    __shared__ double slot[2][HG_CL][2];
    __shared__ double red[32];
    __shared__ int s_last;
    const int rank = (int)cl.block_rank();
    const int mat = blockIdx.x / HG_CL, b = blockIdx.y;
    constexpr int n = AST_C, msize = AST_C * AST_C;
    const bool used = mat < a.nmat_used;
    const float* Ps = used ? a.Ps + ((size_t)b * a.nmat_used + mat) * msize : nullptr;
    const int e0 = rank * (msize / HG_CL) + threadIdx.x;                 // element e0 + k * HG_THREADS
    int par = 0;
    float gr[HG_E], pr[HG_E];
#pragma unroll
    for (int k = 0; k < HG_E; ++k) pr[k] = used ? Ps[e0 + k * HG_THREADS] : 0.f;
    if (a.gram_part) {                                                   // split-K partials, ascending chunk order, fp32 adds
        const float* p = a.gram_part + ((size_t)b * a.nmat + mat) * a.nchunk * msize;
#pragma unroll
        for (int k = 0; k < HG_E; ++k) gr[k] = 0.f;
        for (int c0 = 0; c0 < a.nchunk; c0 += 4) {
            float t[4][HG_E];
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int k = 0; k < HG_E; ++k) t[c][k] = c0 + c < a.nchunk ? p[(size_t)(c0 + c) * msize + e0 + k * HG_THREADS] : 0.f;
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int k = 0; k < HG_E; ++k) gr[k] += t[c][k];
        }
    } else {
        const float* G = a.G + ((size_t)b * a.nmat + mat) * msize;
#pragma unroll
        for (int k = 0; k < HG_E; ++k) gr[k] = G[e0 + k * HG_THREADS];
    }
    double v[2] = {0.0, 0.0};
#pragma unroll
    for (int k = 0; k < HG_E; ++k) v[0] += (double)gr[k] * (double)gr[k];
    hg_cluster_sum(v, 1, wred, slot, par);
    const double ss = v[0];
    const double inv = 1.0 / sqrt(fmax(ss, 1e-12));
    v[0] = 0.0; v[1] = 0.0;                                              // style squared error, <dGn, Gn>
    if (used) {
#pragma unroll
        for (int k = 0; k < HG_E; ++k) {
            const double gn = (double)gr[k] * inv;
            const double d = gn - (double)pr[k];
            v[0] += d * d;
            v[1] += (double)a.coef_dGn * d * gn;
        }
    }
    hg_cluster_sum(v, 2, wred, slot, par);
    const double ls = v[0], dot = v[1];
    // ---- losses: one CTA per matrix reports; the last matrix of the pair to report writes them (fixed-order sums) ----
    if (rank == 0) {
        if (a.done) {
            if (threadIdx.x == 0) {
                a.ls_part[(size_t)b * a.nmat + mat] = (float)ls;
                __threadfence();
                const unsigned prev = atomicAdd(a.done + b, 1u);
                s_last = (prev == (unsigned)a.nmat - 1u);
            }
            __syncthreads();
            if (s_last) {
                __threadfence();
                double sl = 0.0, sc = 0.0;
                for (int m = threadIdx.x; m < a.nmat_used; m += HG_THREADS) sl += (double)__ldcg(a.ls_part + (size_t)b * a.nmat + m);
                if (a.sse_part)
                    for (int i = threadIdx.x; i < a.sse_nblk; i += HG_THREADS) sc += __ldcg(a.sse_part + (size_t)b * a.sse_nblk + i);
                sl = block_sum_d(sl, red);
                sc = a.sse_part ? block_sum_d(sc, red) : (double)a.sse[b];
                if (threadIdx.x == 0) {
                    write_losses(a, b, sl, sc);
                    a.done[b] = 0u;
                }
            }
        } else if (threadIdx.x == 0) {
            a.ls_part[(size_t)b * a.nmat + mat] = (float)ls;
        }
    }
    const bool clamped = ss < 1e-12;
    float* Gn = a.Gn ? a.Gn + ((size_t)b * a.nmat + mat) * msize : nullptr;
    float* Ss = a.Ssym ? a.Ssym + ((size_t)b * a.nmat + mat) * msize : nullptr;
    __nv_bfloat16* Sb = a.Ssym_bf16 ? a.Ssym_bf16 + ((size_t)b * a.nmat + mat) * msize : nullptr;
    __nv_bfloat16* Sbl = a.Ssym_bf16_lo ? a.Ssym_bf16_lo + ((size_t)b * a.nmat + mat) * msize : nullptr;
#pragma unroll
    for (int k = 0; k < HG_E; ++k) {
        const int e = e0 + k * HG_THREADS;
        const int i = e / n, j = e - i * n;
        const double gn = (double)gr[k] * inv;
        if (Gn) Gn[e] = (float)gn;
        double dg = 0.0;
        if (used) {
            const double dgn = (double)a.coef_dGn * (gn - (double)pr[k]);
            dg = clamped ? dgn * inv : (dgn - gn * dot) * inv;
        }
        dgs[i - rank * 32][j] = (float)dg;
    }
    cl.sync();                                   // every CTA's rows of dG are in its shared memory
    if (Ss) {
#pragma unroll
        for (int k = 0; k < HG_E; ++k) {
            const int e = e0 + k * HG_THREADS;
            const int i = e / n, j = e - i * n;
            const float* peer = cl.map_shared_rank(&dgs[j & 31][i], (unsigned)(j >> 5));      // dG[j][i]
            const float val = dgs[i - rank * 32][j] + *peer;
            Ss[e] = val;
            if (Sb) Sb[e] = __float2bfloat16_rn(val);
            if (Sbl) Sbl[e] = __float2bfloat16_rn(val - __bfloat162float(__float2bfloat16_rn(val)));
        }
    }
    cl.sync();                                   // nobody exits while a peer may still read its shared memory
}
// losses[b] = {Lc + lambd Ls + gamma reg, Lc, Ls, reg}  (methods.py:116-125)
__global__ void losses_final_kernel(const HeadArgs a) {
    const int b = blockIdx.x;
    double s = 0.0;                             // one warp, fixed order
    for (int m = threadIdx.x; m < a.nmat_used; m += 32) s += (double)a.ls_part[(size_t)b * a.nmat + m];
    s = warp_sum_d(s);
    if (threadIdx.x != 0) return;
    write_losses(a, b, s, (double)a.sse[b]);
}
int setup_heads_kernel() {
    return (int)cudaFuncSetAttribute(heads_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     AST_C * (AST_C + 1) * (int)sizeof(float));
}
void launch_heads(const HeadArgs& a, cudaStream_t s) {
    dim3 grid(a.nmat, a.B);
    const size_t smem = a.Ssym ? (size_t)a.n * (a.n + 1) * sizeof(float) : 0;
#ifdef AST_EXPERIMENTS
    static const bool one_cta = getenv("AST_B200_HEADS_1CTA") != nullptr;    // A/B switch for profiling
#else
    const bool one_cta = false;
#endif
    if (a.n == AST_C && !one_cta) heads_gatys_cluster_kernel<<<dim3(HG_CL * a.nmat, a.B), HG_THREADS, 0, s>>>(a);
    else heads_kernel<<<grid, a.n * a.n >= 4096 ? 1024 : 256, smem, s>>>(a);
    ast_launch_counter += 1;
    if (a.losses && !a.done) { losses_final_kernel<<<a.B, 32, 0, s>>>(a); ast_launch_counter += 1; }
}

// =============================================================================================
// head gradients
// =============================================================================================
// ours: dF_l[t,c] = sum_m Ssym[c,l,m] F_m[t,c].  block = (32 channels, L layers), grid = (T/64, 4 channel
// quads, B).  Each thread keeps its row Ssym[c,l,:] in registers and walks 64 rows, ROWS at a time: the rows of
// all L extracts are shared through a double-buffered smem tile, and the next round's global loads are issued
// into registers BEFORE the current round's FMAs, so memory latency overlaps the arithmetic (one barrier per
// round).  L is a template parameter: the inner product is exactly L LDS + L FMA per output (a runtime-L loop
// with an `m < L` guard issued 32 predicated pairs and was issue-bound at 3x the work for L = 10).
// Measured dead ends on B200: no staging at all (every thread reads F_m through L1) 2.4x slower; 32-row rounds
// without the register prefetch 1.8x slower than 8-row rounds (all resident CTAs sit in their load phase at once).
template <int L, int ROWS>
__global__ void __launch_bounds__(32 * L) style_grad_ours_kernel(const float* const* __restrict__ F,
                                                                 float* const* __restrict__ dF,
                                                                 const float* __restrict__ Ssym, int T) {
    __shared__ float Fs[2][L][ROWS][32];
    const int cl = threadIdx.x, l = threadIdx.y;
    const int c = blockIdx.y * 32 + cl, b = blockIdx.z;
    float S[L];
#pragma unroll
    for (int m = 0; m < L; ++m) S[m] = Ssym[(((size_t)b * AST_C + c) * L + l) * L + m];
    const int tb = blockIdx.x * 64;
    const float* Fl = F[l] + ((size_t)b * T + tb) * AST_C + c;
    float* dFl = dF[l] + ((size_t)b * T + tb) * AST_C + c;
    float nxt[ROWS];
#pragma unroll
    for (int r = 0; r < ROWS; ++r) nxt[r] = __ldg(Fl + (size_t)r * AST_C);
#pragma unroll
    for (int r = 0; r < ROWS; ++r) Fs[0][l][r][cl] = nxt[r];
    __syncthreads();
    constexpr int NIT = 64 / ROWS;
#pragma unroll 1
    for (int it = 0; it < NIT; ++it) {
        const int cur = it & 1;
        if (it + 1 < NIT) {
#pragma unroll
            for (int r = 0; r < ROWS; ++r) nxt[r] = __ldg(Fl + (size_t)((it + 1) * ROWS + r) * AST_C);
        }
#pragma unroll
        for (int r = 0; r < ROWS; ++r) {
            float acc = 0.f;
#pragma unroll
            for (int m = 0; m < L; ++m) acc = fmaf(S[m], Fs[cur][m][r][cl], acc);
            dFl[(size_t)(it * ROWS + r) * AST_C] = acc;
        }
        if (it + 1 < NIT) {
#pragma unroll
            for (int r = 0; r < ROWS; ++r) Fs[cur ^ 1][l][r][cl] = nxt[r];
        }
        __syncthreads();
    }
}
template <int L>
static void launch_sgo(const float* const* d_F, float* const* d_dF, const float* Ssym, int T, int B, cudaStream_t s) {
    constexpr int ROWS = L <= 24 ? 8 : 4;                      // 2 * L * ROWS * 128 B <= 48 KB of static smem
    dim3 grid(T / 64, AST_C / 32, B), block(32, L);
    style_grad_ours_kernel<L, ROWS><<<grid, block, 0, s>>>(d_F, d_dF, Ssym, T);
}
// One stack (L <= GS_MAXL): the kernel above issues one LDS per FMA (a thread = one (channel, layer) output) and sits at
// 47 us for 168 MB at cfg0.  Here a thread owns one channel and ALL L outputs of a row: the symmetric L x L coefficient
// block lives in L(L+1)/2 registers, a row costs L LDS + L^2 FMAs, outputs go straight from registers to dF (32 channels =
// one 128-byte line per (layer, row)).  Staging as in gram_ours_small_kernel.
#define SGS_ROWS 256            // rows per CTA
template <int L>
__global__ void __launch_bounds__(32 * GS_GROUPS) style_grad_ours_small_kernel(const float* const* __restrict__ F,
                                                                               float* const* __restrict__ dF,
                                                                               const float* __restrict__ Ssym, int T) {
    extern __shared__ __align__(16) float go_s[];          // [stage][L][GS_R][32]
    constexpr int NP = L * (L + 1) / 2;
    constexpr int stage_f = L * GS_R * 32;
    constexpr int nthr = 32 * GS_GROUPS;
    const int cl = threadIdx.x, grp = threadIdx.y, tid = grp * 32 + cl;
    const int c0 = blockIdx.y * 32, b = blockIdx.z;
    const int t_lo = blockIdx.x * SGS_ROWS, t_hi = min(t_lo + SGS_ROWS, T);
    const int nround = (t_hi - t_lo + GS_R - 1) / GS_R;
    auto issue = [&](int round) {
        if (round < nround) {
            float* dst = go_s + (round % GS_STAGES) * stage_f;
            const int t0 = t_lo + round * GS_R;
            for (int i = tid; i < L * GS_R * 8; i += nthr) {
                const int q = i & 7, r = (i >> 3) % GS_R, l = i / (8 * GS_R);
                const bool ok = t0 + r < t_hi;
                const float* src = F[l] + ((size_t)b * T + (ok ? t0 + r : t_lo)) * AST_C + c0 + q * 4;
                cp_async16(smem_u32(dst + (l * GS_R + r) * 32 + q * 4), src, ok);
            }
        }
        cp_async_commit();
    };
#pragma unroll
    for (int st = 0; st < GS_STAGES - 1; ++st) issue(st);
    float S[NP];                                            // upper triangle of the symmetric block of this channel
    {
        const float* Sc = Ssym + ((size_t)b * AST_C + c0 + cl) * L * L;
        int p = 0;
#pragma unroll
        for (int l = 0; l < L; ++l)
#pragma unroll
            for (int m = l; m < L; ++m, ++p) S[p] = Sc[l * L + m];
    }
    for (int round = 0; round < nround; ++round) {
        cp_async_wait<GS_STAGES - 2>();
        __syncthreads();
        issue(round + GS_STAGES - 1);
        const float* sr = go_s + (round % GS_STAGES) * stage_f + cl;
#pragma unroll
        for (int rr = 0; rr < GS_R / GS_GROUPS; ++rr) {
            const int r = grp + rr * GS_GROUPS;
            const int t = t_lo + round * GS_R + r;
            float v[L], o[L];
#pragma unroll
            for (int l = 0; l < L; ++l) { v[l] = sr[(l * GS_R + r) * 32]; o[l] = 0.f; }
            int p = 0;
#pragma unroll
            for (int l = 0; l < L; ++l)
#pragma unroll
                for (int m = l; m < L; ++m, ++p) {
                    o[l] = fmaf(S[p], v[m], o[l]);
                    if (m != l) o[m] = fmaf(S[p], v[l], o[m]);
                }
            if (t < t_hi) {
#pragma unroll
                for (int l = 0; l < L; ++l) dF[l][((size_t)b * T + t) * AST_C + c0 + cl] = o[l];
            }
        }
    }
    cp_async_wait<0>();
}
template <int L>
static void launch_sgos(const float* const* d_F, float* const* d_dF, const float* Ssym, int T, int B, cudaStream_t s) {
    constexpr size_t smem = (size_t)GS_STAGES * L * GS_R * 32 * sizeof(float);
    static bool once = false;
    if (!once) {
        cudaFuncSetAttribute(style_grad_ours_small_kernel<L>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        once = true;
    }
    dim3 grid((T + SGS_ROWS - 1) / SGS_ROWS, AST_C / 32, B), block(32, GS_GROUPS);
    style_grad_ours_small_kernel<L><<<grid, block, smem, s>>>(d_F, d_dF, Ssym, T);
}
// Several stacks (L > GS_MAXL, e.g. the CLI default: all 30 layers): the per-(channel, layer) kernel below issues one
// LDS per FMA (900 of each per (t, c) at L = 30).  Here a thread owns one channel, SGT_LO = 6 outputs and SGT_RT = 4
// rows: per input layer m it loads 6 coefficients + 4 values for 24 FMAs (2.4 FMA per LDS).  The coefficient block of
// the CTA's 32 channels sits in shared memory as [l][m][32] (conflict-free, 115 KB at L = 30), rows are staged by
// cp.async, 8 per round.  Same summation order over m as the other two kernels.
#define SGT_LO 6
#define SGT_RT 4
#define SGT_R 8
template <int LT>          // LT > 0: L known at compile time (immediate shared-memory offsets, unrolled m loop); 0: runtime L
__global__ void __launch_bounds__(32 * 6 * (SGT_R / SGT_RT)) style_grad_ours_tiled_kernel(const float* const* __restrict__ F,
                                                                                         float* const* __restrict__ dF,
                                                                                         const float* __restrict__ Ssym,
                                                                                         int Lrt, int T, int nstage, int rows_per_cta) {
    const int L = LT ? LT : Lrt;
    extern __shared__ __align__(16) float go_s[];          // coef [L][L][33] | stages [nstage][L][SGT_R][32]
    const int cl = threadIdx.x, og = threadIdx.y, rg = threadIdx.z;
    const int nthr = blockDim.x * blockDim.y * blockDim.z;
    const int tid = (rg * blockDim.y + og) * 32 + cl;
    const int c0 = blockIdx.y * 32, b = blockIdx.z;
    float* coef = go_s;                                     // row pitch 33: the transposing fill and the reads are both conflict-free
    float* stg = go_s + ((L * L * 33 + 3) & ~3);
    const int stage_f = L * SGT_R * 32;
    const int t_lo = blockIdx.x * rows_per_cta, t_hi = min(t_lo + rows_per_cta, T);
    const int nround = t_hi > t_lo ? (t_hi - t_lo + SGT_R - 1) / SGT_R : 0;
    auto issue = [&](int round) {
        if (round < nround) {
            float* dst = stg + (round % nstage) * stage_f;
            const int t0 = t_lo + round * SGT_R;
            for (int i = tid; i < L * SGT_R * 8; i += nthr) {
                const int q = i & 7, r = (i >> 3) % SGT_R, l = i / (8 * SGT_R);
                const bool ok = t0 + r < t_hi;
                const float* src = F[l] + ((size_t)b * T + (ok ? t0 + r : t_lo)) * AST_C + c0 + q * 4;
                cp_async16(smem_u32(dst + (l * SGT_R + r) * 32 + q * 4), src, ok);
            }
        }
        cp_async_commit();
    };
    for (int st = 0; st < nstage - 1; ++st) issue(st);
    {   // coef[l][m][ch] = Ssym[b][c0 + ch][l][m]: the 32 channels' blocks are one contiguous run of 32 L^2 floats; read it
        // coalesced (the first version gathered 32 lines per warp load and spent a quarter of the kernel here)
        const float* src = Ssym + ((size_t)b * AST_C + c0) * L * L;
        const int LL = L * L;
        for (int i = tid; i < LL * 32; i += nthr) {
            const int ch = i / LL, lm = i - ch * LL;
            coef[lm * 33 + ch] = src[i];
        }
    }
    const int l0 = og * SGT_LO;
    for (int round = 0; round < nround; ++round) {
        if (nstage == 3) cp_async_wait<1>(); else cp_async_wait<0>();
        __syncthreads();                                    // (also publishes coef before its first use)
        issue(round + nstage - 1);
        const float* sr = stg + (round % nstage) * stage_f + (rg * SGT_RT) * 32 + cl;
        float acc[SGT_LO][SGT_RT];
#pragma unroll
        for (int i = 0; i < SGT_LO; ++i)
#pragma unroll
            for (int r = 0; r < SGT_RT; ++r) acc[i][r] = 0.f;
#pragma unroll (LT ? 6 : 2)
        for (int m = 0; m < L; ++m) {
            float v[SGT_RT], cf[SGT_LO];
#pragma unroll
            for (int r = 0; r < SGT_RT; ++r) v[r] = sr[(m * SGT_R + r) * 32];
#pragma unroll
            for (int i = 0; i < SGT_LO; ++i) cf[i] = l0 + i < L ? coef[((l0 + i) * L + m) * 33 + cl] : 0.f;
#pragma unroll
            for (int i = 0; i < SGT_LO; ++i)
#pragma unroll
                for (int r = 0; r < SGT_RT; ++r) acc[i][r] = fmaf(cf[i], v[r], acc[i][r]);
        }
#pragma unroll
        for (int r = 0; r < SGT_RT; ++r) {
            const int t = t_lo + round * SGT_R + rg * SGT_RT + r;
            if (t < t_hi) {
#pragma unroll
                for (int i = 0; i < SGT_LO; ++i)
                    if (l0 + i < L) dF[l0 + i][((size_t)b * T + t) * AST_C + c0 + cl] = acc[i][r];
            }
        }
        if (nstage == 2) __syncthreads();                   // two stages: the one just read is refilled by the next issue
    }
    cp_async_wait<0>();
}
static bool launch_sgot(const float* const* d_F, float* const* d_dF, const float* Ssym, int L, int T, int B, cudaStream_t s) {
    const int nog = (L + SGT_LO - 1) / SGT_LO;
    if (nog > 6) return false;
    int nstage = 3;
    auto bytes = [&](int ns) { return ((size_t)((L * L * 33 + 3) & ~3) + (size_t)ns * L * SGT_R * 32) * sizeof(float); };
    if (bytes(nstage) > 227 * 1024) nstage = 2;
    if (bytes(nstage) > 227 * 1024) return false;
    static size_t set_for = 0;
    if (bytes(nstage) > set_for) {
        if (cudaFuncSetAttribute(style_grad_ours_tiled_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes(nstage)) != cudaSuccess ||
            cudaFuncSetAttribute(style_grad_ours_tiled_kernel<30>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes(nstage)) != cudaSuccess)
            return false;
        set_for = bytes(nstage);
    }
    // one CTA per SM (207 KB of shared memory) and equal work per CTA: exactly one wave
    const int nx = std::max(1, 148 / (4 * B));
    const int rows = std::max(SGT_R, ((T + nx - 1) / nx + SGT_R - 1) / SGT_R * SGT_R);
    dim3 grid((T + rows - 1) / rows, AST_C / 32, B), block(32, nog, SGT_R / SGT_RT);
    if (L == 30) style_grad_ours_tiled_kernel<30><<<grid, block, bytes(nstage), s>>>(d_F, d_dF, Ssym, L, T, nstage, rows);   // all three stacks
    else         style_grad_ours_tiled_kernel<0><<<grid, block, bytes(nstage), s>>>(d_F, d_dF, Ssym, L, T, nstage, rows);
    return true;
}
void launch_style_grad_ours(const float* const* d_F, float* const* d_dF, const float* Ssym, int L, int T, int B,
                            cudaStream_t s) {
#ifdef AST_EXPERIMENTS
    static const bool old_path = getenv("AST_B200_SGO_OLD") != nullptr;      // A/B switch for profiling: per-(channel, layer) kernel
#else
    const bool old_path = false;
#endif
    if (L > GS_MAXL && !old_path && launch_sgot(d_F, d_dF, Ssym, L, T, B, s)) { ast_launch_counter += 1; return; }
    if (L <= GS_MAXL) {
#define SGOS_CASE(n) case n: launch_sgos<n>(d_F, d_dF, Ssym, T, B, s); break;
        switch (L) {
            SGOS_CASE(1) SGOS_CASE(2) SGOS_CASE(3) SGOS_CASE(4) SGOS_CASE(5) SGOS_CASE(6) SGOS_CASE(7) SGOS_CASE(8)
            SGOS_CASE(9) SGOS_CASE(10) SGOS_CASE(11) SGOS_CASE(12)
            default: break;
        }
#undef SGOS_CASE
        ast_launch_counter += 1;
        return;
    }
#define SGO_CASE(n) case n: launch_sgo<n>(d_F, d_dF, Ssym, T, B, s); break;
    switch (L) {
        SGO_CASE(1) SGO_CASE(2) SGO_CASE(3) SGO_CASE(4) SGO_CASE(5) SGO_CASE(6) SGO_CASE(7) SGO_CASE(8)
        SGO_CASE(9) SGO_CASE(10) SGO_CASE(11) SGO_CASE(12) SGO_CASE(13) SGO_CASE(14) SGO_CASE(15) SGO_CASE(16)
        SGO_CASE(17) SGO_CASE(18) SGO_CASE(19) SGO_CASE(20) SGO_CASE(21) SGO_CASE(22) SGO_CASE(23) SGO_CASE(24)
        SGO_CASE(25) SGO_CASE(26) SGO_CASE(27) SGO_CASE(28) SGO_CASE(29) SGO_CASE(30) SGO_CASE(31) SGO_CASE(32)
        default: break;                                        // ast_create rejects L outside 1..32
    }
#undef SGO_CASE
    ast_launch_counter += 1;
}

// gatys: dF_l = F_l (dG_l + dG_l^T): a [rows x 128].[128 x 128] GEMM per style layer.  grid = (B*T/64, L).
__global__ void __launch_bounds__(256) style_grad_gatys_kernel(const float* const* __restrict__ F,
                                                               float* const* __restrict__ dF,
                                                               const float* __restrict__ Ssym, int L, int T) {
    __shared__ __align__(16) float As[LF_BM * LF_AS_LD];
    __shared__ __align__(16) float Ws[LF_KC * AST_C];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int l = blockIdx.y;
    const int row0 = blockIdx.x * LF_BM;
    const int b = row0 / T;
    const float* A = F[l];
    const float* W = Ssym + ((size_t)b * L + l) * AST_C * AST_C;
    float acc[4][8];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    for (int kc = 0; kc < AST_C / LF_KC; ++kc) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int idx = tid + i * 256, r = idx >> 3, q = idx & 7;
            *reinterpret_cast<float4*>(As + r * LF_AS_LD + q * 4) =
                *reinterpret_cast<const float4*>(A + (size_t)(row0 + r) * AST_C + kc * LF_KC + q * 4);
        }
        lf_load_w(Ws, W + (size_t)kc * LF_KC * AST_C, tid);
        __syncthreads();
        lf_mma(acc, As, LF_AS_LD, Ws, ty, tx);
        __syncthreads();
    }
    float* out = dF[l];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const size_t row = (size_t)row0 + ty * 4 + i;
        *reinterpret_cast<float4*>(out + row * AST_C + tx * 4) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        *reinterpret_cast<float4*>(out + row * AST_C + 64 + tx * 4) = make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]);
    }
}
void launch_style_grad_gatys(const float* const* d_F, float* const* d_dF, const float* Ssym, int L, int T, int B,
                             cudaStream_t s) {
    dim3 grid(B * T / LF_BM, L);
    style_grad_gatys_kernel<<<grid, 256, 0, s>>>(d_F, d_dF, Ssym, L, T);
    ast_launch_counter += 1;
}

// content: dF[t,j] (+)= coef (H[t,j] - Pc[t, slot*cnt + j]) for j < cnt (methods.py:58,116-117)
__global__ void content_grad_kernel(const float* __restrict__ H, const float* __restrict__ Pc, float* __restrict__ dF,
                                    int slot, int n_cont, int cnt_ch, float coef, int accumulate, int R) {
    const size_t i4 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;      // one thread = 4 consecutive channels of a row
    if (i4 >= (size_t)R * (AST_C / 4)) return;
    const size_t row = i4 / (AST_C / 4);
    const int j = (int)(i4 % (AST_C / 4)) * 4;
    const float4 h = reinterpret_cast<const float4*>(H)[i4];
    const float* p = Pc + row * (size_t)(n_cont * cnt_ch) + (size_t)slot * cnt_ch + j;
    float4 v;
    v.x = j + 0 < cnt_ch ? coef * (h.x - p[0]) : 0.f;
    v.y = j + 1 < cnt_ch ? coef * (h.y - p[1]) : 0.f;
    v.z = j + 2 < cnt_ch ? coef * (h.z - p[2]) : 0.f;
    v.w = j + 3 < cnt_ch ? coef * (h.w - p[3]) : 0.f;
    float4* out = reinterpret_cast<float4*>(dF) + i4;
    if (accumulate) { const float4 o = *out; v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w; }
    *out = v;
}
void launch_content_grad(const float* H, const float* Pc, float* dF, int slot, int n_cont, int cnt_ch, float coef,
                         int accumulate, int R, cudaStream_t s) {
    const size_t n4 = (size_t)R * (AST_C / 4);
    content_grad_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, s>>>(H, Pc, dF, slot, n_cont, cnt_ch, coef, accumulate, R);
    ast_launch_counter += 1;
}
// Ec[t, slot*cnt + j] = extracts[id_slot][t, j]  (methods.py:58)
__global__ void content_gather_kernel(const float* const* __restrict__ Hc, int n_cont, int cnt_ch,
                                      float* __restrict__ Ec, int R) {
    const int ncols = n_cont * cnt_ch;
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)R * ncols) return;
    const size_t row = i / ncols;
    const int col = (int)(i % ncols);
    const int slot = col / cnt_ch, j = col % cnt_ch;
    Ec[i] = Hc[slot][row * AST_C + j];
}
void launch_content_gather(const float* const* d_Hc, int n_cont, int cnt_ch, float* Ec, int R, cudaStream_t s) {
    const size_t n = (size_t)R * n_cont * cnt_ch;
    content_gather_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(d_Hc, n_cont, cnt_ch, Ec, R);
    ast_launch_counter += 1;
}
