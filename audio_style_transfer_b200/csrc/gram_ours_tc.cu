// gram_ours_tc.cu -- channel-wise ("ours") Gram for many style layers (13 <= L <= 32: the CLI default of all 30 layers,
// BASELINE configs[4]) on tensor cores.  Reference: methods.py:69,73  G[c,l,m] = sum_t S[l,t,c] S[m,t,c].
//
// Per channel c this is a 32 x 32 x T product V_c V_c^T with V_c = [layer][t].  The CUDA-core kernel it replaces in the
// tensor-core tier (gram_ours_kernel, kernels_f32.cu) is issue-bound: 640 FMAs per (t, c), 155 us for 252 MB at L = 30,
// T = 16384 (FMA bound 37 us, HBM bound 39 us).  The shape -- 128 independent 32 x 32 accumulators, K streamed from HBM once
// -- does not fit tcgen05's 128-row tiles and shared-memory descriptors, so this is warp-level `mma.sync.m16n8k8` with
// single-pass TF32 operands (round to nearest, 2^-11 per operand -- four times finer than the bf16 extract copies the
// Gatys Gram of the same tier multiplies; the unbiased product errors average over the 16384 terms of a sum: measured
// Gn rel-L2 vs fp64 in tests/test_gpu_parity.py), fp32 accumulate.  Round 1 split the operands in two terms (3 MMAs per
// product): ncu showed the kernel throttling on exactly that tensor work (math pipe throttle the top stall, 48 % pipe
// activity, profiles/r2_ncu_stack_and_head_kernels.md).
//
// A warp owns 2 channels.  With K = t, the A fragment (rows = layer, cols = t) and the B fragment (k = t, n = layer) of
// V V^T are the same loads: per k-step of 8 rows a lane reads V[g + 8a][tig + 4b] (g = lane / 4, tig = lane % 4,
// a < 4, b < 2) = 8 values, which feed the 6 upper-triangular 16 x 8 tiles.  For those loads to be conflict-free the
// staging buffer is [channel][layer][t] with a row pitch of 12 floats (banks 12 g + tig are all distinct) and a plane
// pitch of 32 * 12 + 1; HBM rows are [t][channel], so the transposition goes through registers: coalesced 16-byte loads
// of 4 channels, prefetched TWO rounds ahead in two register sets (one round ahead measured 113 us: a round then costs
// latency + transfer), 4 scattered (conflict-free) 4-byte shared stores each.
#include "common.cuh"
#include "ptx.cuh"

#include <algorithm>
#include <cstdlib>

#define GT_THREADS 512                      // 16 warps x 2 channels = 32 channels per CTA
#define GT_R 8                              // rows (t) per round = 1 k-step
#define GT_PT 12                            // row pitch (floats): 12 g + tig distinct mod 32 for g < 8, tig < 4
#define GT_PL (32 * GT_PT + 1)              // channel-plane pitch: 4 q * GT_PL = 4 q (mod 32) for the transposing stores
#define GT_STAGE (32 * GT_PL)               // floats per stage
#define GT_NLD ((32 * GT_R * 8) / GT_THREADS)                          // 16-byte loads per thread per round at L = 32

__device__ __forceinline__ uint32_t tf32_of(float x) {
    uint32_t u;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
    return u;
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// grid = (nchunk, 4 channel quads, B); part[b][chunk][c][l][m] like the CUDA-core kernels.
__global__ void __launch_bounds__(GT_THREADS, 1) gram_ours_tc_kernel(const __grid_constant__ OursPtrs P, int L, int T, int own_lo,
                                                                    int own_hi, float* __restrict__ part, int nchunk, int chunk_rows) {
    extern __shared__ __align__(16) float gt_s[];          // [2 stages][32 channels][32 layers][GT_PT] (+1 per plane)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, tig = lane & 3;
    const int c0 = blockIdx.y * 32, b = blockIdx.z;
    const int tb = blockIdx.x * chunk_rows;
    const int t_lo = max(tb, own_lo), t_hi = min(min(tb + chunk_rows, T), own_hi);
    const int nround = t_hi > t_lo ? (t_hi - t_lo + GT_R - 1) / GT_R : 0;
    const int nld = (L * GT_R * 8 + GT_THREADS - 1) / GT_THREADS;          // <= GT_NLD
    for (int i = tid; i < 2 * GT_STAGE; i += GT_THREADS) gt_s[i] = 0.f;    // layers L..31 stay zero
    __syncthreads();

    auto fetch = [&](float4 (&pf)[GT_NLD], int round) {                    // global -> registers (coalesced: 8 lanes = one 128-byte row segment)
        const int t0 = t_lo + round * GT_R;
#pragma unroll
        for (int it = 0; it < GT_NLD; ++it) {
            const int i = tid + it * GT_THREADS;
            const int q = i & 7, t = (i >> 3) & (GT_R - 1), l = i / (8 * GT_R);
            const bool ok = it < nld && l < L && t0 + t < t_hi;
            pf[it] = ok ? *reinterpret_cast<const float4*>(P.F[l < 32 ? l : 0] + ((size_t)b * T + t0 + t) * AST_C + c0 + q * 4)
                        : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    };
    auto stash = [&](const float4 (&pf)[GT_NLD], int stage) {              // registers -> [channel][layer][t]
        float* dst = gt_s + stage * GT_STAGE;
#pragma unroll
        for (int it = 0; it < GT_NLD; ++it) {
            const int i = tid + it * GT_THREADS;
            const int q = i & 7, t = (i >> 3) & (GT_R - 1), l = i / (8 * GT_R);
            if (it < nld && l < L) {
                float* d = dst + (q * 4) * GT_PL + l * GT_PT + t;
                d[0] = pf[it].x; d[GT_PL] = pf[it].y; d[2 * GT_PL] = pf[it].z; d[3 * GT_PL] = pf[it].w;
            }
        }
    };

    float acc[2][6][4];
#pragma unroll
    for (int ch = 0; ch < 2; ++ch)
#pragma unroll
        for (int tl = 0; tl < 6; ++tl)
#pragma unroll
            for (int k = 0; k < 4; ++k) acc[ch][tl][k] = 0.f;

    auto compute = [&](int stage) {
        const float* st = gt_s + stage * GT_STAGE;
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
            const float* pl = st + (warp * 2 + ch) * GT_PL + g * GT_PT + tig;
#pragma unroll
            for (int ks = 0; ks < GT_R / 8; ++ks) {
                uint32_t hi[4][2];                                         // V[g + 8a][8 ks + tig + 4 bb]
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int bb = 0; bb < 2; ++bb) hi[a][bb] = tf32_of(pl[(8 * a) * GT_PT + 8 * ks + 4 * bb]);
                // tiles (row block of 16 layers, column block of 8 layers), upper triangle: (0,0..3), (1,2..3)
#pragma unroll
                for (int tl = 0; tl < 6; ++tl) {
                    const int mt = tl < 4 ? 0 : 1, j = tl < 4 ? tl : tl - 2;
                    const int a0 = 2 * mt, a1 = 2 * mt + 1;
                    mma_tf32(acc[ch][tl], hi[a0][0], hi[a1][0], hi[a0][1], hi[a1][1], hi[j][0], hi[j][1]);
                }
            }
        }
    };
    // On entry to round r: stage r & 1 holds round r, register set (r + 1) & 1 holds round r + 1 (requested a round ago), the
    // other set holds round r + 2 (just requested).  A round stashes r + 1, re-requests r + 3 into the freed set, computes r.
    float4 s0[GT_NLD], s1[GT_NLD];
    auto round_body = [&](int r, float4 (&x)[GT_NLD]) {
        if (r + 1 < nround) stash(x, (r + 1) & 1);
        if (r + 3 < nround) fetch(x, r + 3);
        compute(r & 1);
        __syncthreads();
    };
    if (nround > 0) {
        fetch(s0, 0);
        stash(s0, 0);
        if (nround > 1) fetch(s1, 1);
        if (nround > 2) fetch(s0, 2);
    }
    __syncthreads();
    for (int r = 0; r < nround; r += 2) {
        round_body(r, s1);
        if (r + 1 < nround) round_body(r + 1, s0);
    }
    // C fragment: c0 (row g, col 2 tig), c1 (g, 2 tig + 1), c2 (g + 8, 2 tig), c3 (g + 8, 2 tig + 1).  Entries with l <= m are written
    // together with their mirror, so the partial is exactly symmetric.
#pragma unroll
    for (int ch = 0; ch < 2; ++ch) {
        float* out = part + (((size_t)b * nchunk + blockIdx.x) * AST_C + c0 + warp * 2 + ch) * L * L;
#pragma unroll
        for (int tl = 0; tl < 6; ++tl) {
            const int mt = tl < 4 ? 0 : 1, j = tl < 4 ? tl : tl - 2;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int l = 16 * mt + g + (k >> 1) * 8, m = 8 * j + 2 * tig + (k & 1);
                if (l <= m && m < L) {
                    out[l * L + m] = acc[ch][tl][k];
                    out[m * L + l] = acc[ch][tl][k];
                }
            }
        }
    }
}

int gram_ours_tc_supported(int L) { return L > gram_ours_max_small_L() && L <= 32; }

void launch_gram_ours_tc(const OursPtrs& ptrs, int L, int T, int B, int own_lo, int own_hi, float* part, int nchunk,
                         int chunk_rows, cudaStream_t s) {
    const size_t smem = (size_t)2 * GT_STAGE * sizeof(float);
    static bool once = false;
    if (!once) {
        cudaFuncSetAttribute(gram_ours_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        once = true;
    }
    dim3 grid(nchunk, AST_C / 32, B);
    gram_ours_tc_kernel<<<grid, GT_THREADS, smem, s>>>(ptrs, L, T, own_lo, own_hi, part, nchunk, chunk_rows);
    ast_launch_counter += 1;
}

// =====================================================================================================================
// Channel-wise style gradient on tensor cores (13 <= L <= 32), reference SURVEY.md 8(a):
//   dF_l[t, c] = sum_m Ssym[c, l, m] F_m[t, c]        (Ssym = dG + dG^T, symmetric)
// Per channel: OUT[t][l] = V^T[t][m] S[m][l]: M = t (16 rows per round), K = m (4 k-steps), N = l (4 tiles of 8).
// The coefficient block of a warp's 2 channels lives in registers for the whole CTA as TF32 B fragments (rounded once:
// 2^-11 relative, the same class as the bf16 Ssym of the Gatys path); so is the A operand (round 1 split it in two terms).
// Staging [channel][layer][t] with row pitch 24 (banks 24 tig + g all distinct) is filled by 4-byte cp.async (HBM rows are
// [t][channel]; no register staging, so the next round is in flight while this one is computed and written out).
// Outputs go back through the consumed stage as [l][t][c] (pitches 548 / 33: conflict-free both ways) and leave as
// 128-byte rows (8-byte stores straight from the accumulator registers -- 32 rows per warp instruction, partial sectors --
// were measured: 302 us against 207 us; so was the transposed product OUT^T = S V with 8-row rounds, four 33 KB stages,
// a separate output double buffer and ONE barrier per round: 252 us -- the kernel is not waiting for its loads).  The CUDA-core kernel it replaces (style_grad_ours_tiled_kernel) is 258 us for 504 MB at L = 30.
// =====================================================================================================================
#define ST_THREADS 512
#define ST_R 16
#define ST_PT 24
#define ST_PL (32 * ST_PT + 1)
#define ST_STAGE (32 * ST_PL)                 // floats per input stage (98.4 KB)
#define ST_OPL 548                            // output staging: index = l * 548 + t * 33 + c
#define ST_OPT 33

__device__ __forceinline__ void cp_async_commit_st() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src, bool valid) {
    const int sz = valid ? 4 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}

__global__ void __launch_bounds__(ST_THREADS, 1) style_grad_ours_tc_kernel(const __grid_constant__ OursPtrs P,
                                                                          const float* __restrict__ Ssym, int L, int T,
                                                                          int rows_per_cta) {
    extern __shared__ __align__(16) float st_s[];          // [2][ST_STAGE]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, tig = lane & 3;
    const int c0 = blockIdx.y * 32, b = blockIdx.z;
    const int t_lo = blockIdx.x * rows_per_cta, t_hi = min(t_lo + rows_per_cta, T);
    const int nround = t_hi > t_lo ? (t_hi - t_lo + ST_R - 1) / ST_R : 0;
    for (int i = tid; i < 2 * ST_STAGE; i += ST_THREADS) st_s[i] = 0.f;      // layers L..31 stay zero
    // B fragments: b0 = S[m = 8 ks + tig][l = 8 j + g], b1 = S[m + 4][l]
    uint32_t bf[2][4][4][2];
#pragma unroll
    for (int ch = 0; ch < 2; ++ch) {
        const float* Sc = Ssym + ((size_t)b * AST_C + c0 + warp * 2 + ch) * L * L;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int m = 8 * ks + tig + 4 * h, l = 8 * j + g;
                    bf[ch][ks][j][h] = tf32_of(m < L && l < L ? Sc[l * L + m] : 0.f);
                }
    }
    __syncthreads();
    auto issue = [&](int round) {                                          // 4-byte cp.async, transposing [t][c] -> [c][l][t]
        if (round < nround) {
            const uint32_t dst = smem_u32(st_s + (round & 1) * ST_STAGE);
            const int t0 = t_lo + round * ST_R;
            // ST_THREADS == 32 * ST_R: a thread keeps its (t, c) and walks the layers; the layer's base pointer is uniform
            const int c = tid & 31, t = tid >> 5;
            const bool ok = t0 + t < t_hi;
            const size_t off = ((size_t)b * T + (ok ? t0 + t : t_lo)) * AST_C + c0 + c;
            const uint32_t d0 = dst + (uint32_t)(c * ST_PL + t) * 4u;
            for (int l = 0; l < L; ++l) cp_async4(d0 + (uint32_t)(l * ST_PT) * 4u, P.F[l] + off, ok);
        }
        cp_async_commit_st();
    };
    issue(0);
    issue(1);
    for (int round = 0; round < nround; ++round) {
        asm volatile("cp.async.wait_group 1;" ::: "memory");              // this thread's copies of `round` have landed
        __syncthreads();                                                  // everyone's have
        float* st = st_s + (round & 1) * ST_STAGE;
        float acc[2][4][4];
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int k = 0; k < 4; ++k) acc[ch][j][k] = 0.f;
            const float* pl = st + (warp * 2 + ch) * ST_PL + tig * ST_PT + g;
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                // A(row t, col m) = V[m][t]: a0 (t = g, m = 8 ks + tig), a1 (t = g + 8, m), a2 (t = g, m + 4), a3 (t = g + 8, m + 4)
                uint32_t ah[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) ah[k] = tf32_of(pl[(8 * ks + 4 * (k >> 1)) * ST_PT + 8 * (k & 1)]);
#pragma unroll
                for (int j = 0; j < 4; ++j) mma_tf32(acc[ch][j], ah[0], ah[1], ah[2], ah[3], bf[ch][ks][j][0], bf[ch][ks][j][1]);
            }
        }
        __syncthreads();                                                  // everyone is done reading this stage: reuse it for the outputs
        // C fragment: c0 (t = g, l = 8 j + 2 tig), c1 (t = g, l + 1), c2 (t = g + 8, l), c3 (t = g + 8, l + 1)
#pragma unroll
        for (int ch = 0; ch < 2; ++ch)
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int l = 8 * j + 2 * tig + (k & 1), t = g + 8 * (k >> 1);
                    st[l * ST_OPL + t * ST_OPT + warp * 2 + ch] = acc[ch][j][k];
                }
        __syncthreads();
        const int t0 = t_lo + round * ST_R;
        {                                                                 // 32 lanes = one 128-byte row of dF_l
            const int c = tid & 31, t = tid >> 5;
            if (t0 + t < t_hi) {
                const size_t off = ((size_t)b * T + t0 + t) * AST_C + c0 + c;
                const float* so = st + t * ST_OPT + c;
#pragma unroll 6
                for (int l = 0; l < L; ++l) P.dF[l][off] = so[l * ST_OPL];
            }
        }
        __syncthreads();                                                  // outputs are out: the stage may be refilled
        // the refill must leave layers L..31 of the INPUT layout zero: the output staging has just overwritten parts of them
        for (int i = tid; i < (32 - L) * 32 * ST_R; i += ST_THREADS) {
            const int t = i & (ST_R - 1), c = (i / ST_R) & 31, l = L + i / (32 * ST_R);
            st[c * ST_PL + l * ST_PT + t] = 0.f;
        }
        issue(round + 2);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

int style_grad_ours_tc_supported(int L) { return gram_ours_tc_supported(L); }

static_assert(ST_THREADS == 32 * ST_R, "a thread of the style-gradient kernel owns one (t, c) of a round");
void launch_style_grad_ours_tc(const OursPtrs& ptrs, const float* Ssym, int L, int T, int B, cudaStream_t s) {
    const size_t smem = (size_t)2 * ST_STAGE * sizeof(float);
    static bool once = false;
    if (!once) {
        cudaFuncSetAttribute(style_grad_ours_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        once = true;
    }
    const int nx = std::max(1, 148 / (4 * B));                            // one CTA per SM, one wave
    const int rows = std::max(ST_R, ((T + nx - 1) / nx + ST_R - 1) / ST_R * ST_R);
    dim3 grid((T + rows - 1) / rows, AST_C / 32, B);
    style_grad_ours_tc_kernel<<<grid, ST_THREADS, smem, s>>>(ptrs, Ssym, L, T, rows);
    ast_launch_counter += 1;
}
