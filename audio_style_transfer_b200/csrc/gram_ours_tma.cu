// gram_ours_tma.cu -- channel-wise ("ours") Gram and style gradient (1 <= L <= 32 style layers: one stack = the reference's
// --stack 0/1/2, BASELINE configs[0]; all 30 layers = the CLI default, configs[4]) fed by TMA from TRANSPOSED extract
// copies.  Reference: methods.py:62-73
//   G[c,l,m] = sum_t S[l,t,c] S[m,t,c]        dF_l[t,c] = sum_m Ssym[c,l,m] F_m[t,c]
//
// Both are per-channel products over V_c = [layer][t].  The extracts leave the encoder as [t][channel] rows, so round 1/2
// (gram_ours_tc.cu, kept for the per-layer kernel path) transposed every element on its way into shared memory: through
// registers and four scattered 4-byte shared stores (Gram), or one 4-byte cp.async per element (gradient).  ncu, round 2:
// the Gram kernel at 2.8 TB/s, the gradient kernel latency-bound at 25 % occupancy, 2.2 TB/s.  Here the whole-stack forward
// kernel writes a second copy of every style extract as Ht[row / 32][slot][channel][row % 32] in fp16 (stack_fwd.cuh
// finish(): a warp's 32 TMEM lanes are 32 consecutive rows, so one column of a warp is one 64-byte run), and both kernels
// stream it with one TMA box per stage: {32 t, LP layers, 32 channels} fp16 = LP x 2 KB, SWIZZLE_64B, layers L..LP-1
// zero-filled by the out-of-bounds rule.  The 32-row blocking keeps DRAM pages busy on both sides: a stage is L runs of
// 2 KB (32 channels x 64 B) 8 KB apart, a tile-layer of the forward writes 4 runs of 8 KB.  (First version:
// Ht[slot][channel][row] in fp32 -- a stage was then 960 runs of 64 B, 64 KB apart, and both kernels stalled at 1.7 TB/s.)
//
// Operand precision.  fp16 keeps 11 significant bits -- exactly what the single-pass TF32 operands of the kernels this
// file replaces kept (cvt.rna.tf32) -- at half the bytes of the fp32 copies the second version streamed; the conversion
// saturates at +-65504 (cvt.rn.satfinite), values below 6e-5 go subnormal (absolute error <= 3e-8: nothing next to a Gram
// sum).  Accumulation is fp32 (mma.sync.m16n8k16.f16.f32).  The coefficient block of the gradient is scaled per channel by
// a power of two that brings its largest entry next to 1 before it is rounded to fp16, and the outputs are scaled back.
//
// No thread touches an operand before it is an mma.sync fragment:
//   Gram:      A (rows = layer, cols = t) and B (k = t, n = layer) are the same 32-bit loads (two consecutive t);
//   gradient:  OUT[t][l] = V^T S: A = V^T by ldmatrix.trans (an 8 x 8 block of [layer][t] arrives as [t][layer] pairs), the
//              coefficient block S as register-resident B fragments.
// Both are conflict-free under the 64-byte swizzle (16-byte chunk index ^= bits 7..8 of the offset): see the address
// comments.  Thread 0 keeps the ring full (full / empty mbarriers); 16 warps own 2 channels each.
#include "common.cuh"
#include "ptx.cuh"

#include <cuda.h>
#include <cuda_fp16.h>
#include <algorithm>

#define HT_TK 32                              // t per stage = rows per block of Ht
#define HT_CH 32                              // channels per CTA
// LP = layers per stage: 32 (17..32 style layers) or 16 (one stack, the reference's --stack 0/1/2: <= 16); a stage is
// [channel 32][layer LP][t 32] fp16 = LP * 2048 bytes
#define HT_STAGE(LP) (HT_CH * (LP) * HT_TK * 2)
#define HT_THREADS 512                        // 16 warps x 2 channels; no producer warp: a 17th warp would cap the kernel at 96 registers

__device__ __forceinline__ void ht_mma(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ht_bar_all() { asm volatile("bar.sync 1, %0;" ::"n"(HT_THREADS) : "memory"); }
__device__ __forceinline__ uint32_t ht_lds(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void ht_ldmatrix_x4_trans(uint32_t (&d)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]) : "r"(addr));
}
__device__ __forceinline__ uint32_t ht_pack_f16(float lo, float hi) {
    uint32_t u;
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(u) : "f"(hi), "f"(lo));      // low half <- lo
    return u;
}

// Thread 0 requests round r into stage r % NS (the caller has made sure the stage is free).
template <int LP, int NS>
__device__ __forceinline__ void ht_request(const CUtensorMap* tm, uint32_t stage0, uint32_t bar_full, int r, int row0, int c0) {
    const int s = r % NS;
    mbar_expect_tx(bar_full + 8u * s, HT_STAGE(LP));
    tma_load_4d(stage0 + (uint32_t)s * HT_STAGE(LP), tm, bar_full + 8u * s, 0, 0, c0, (row0 >> 5) + r);
}

// ---------------------------------------------------------------------------------------------------------------------
// Gram.  grid = (nchunk, 4 channel quads, B); part[b][chunk][c][l][m] like the other channel-wise Gram kernels.
// chunk_rows, own_lo, own_hi and T are multiples of 32: every round is a full stage.
// ---------------------------------------------------------------------------------------------------------------------
#define HG_NS(LP) ((LP) == 32 ? 3 : 5)
#define HG_SMEM(LP) (HG_NS(LP) * HT_STAGE(LP) + 1024 + 64)

template <int LP>
__global__ void __launch_bounds__(HT_THREADS, 1) gram_ours_tma_kernel(const __grid_constant__ CUtensorMap tmHt, int L, int T,
                                                                     int own_lo, int own_hi, float* __restrict__ part,
                                                                     int nchunk, int chunk_rows) {
    constexpr int NS = HG_NS(LP), NA = LP / 8, NT = LP == 32 ? 6 : 2;     // stages; 8-layer blocks; upper-triangular 16 x 8 tiles
    extern __shared__ uint8_t hg_raw[];
    const uint32_t base = (smem_u32(hg_raw) + 1023u) & ~1023u;
    const uint32_t bar_full = base + NS * HT_STAGE(LP), bar_empty = bar_full + 8u * NS;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int c0 = blockIdx.y * HT_CH, b = blockIdx.z;
    const int tb = blockIdx.x * chunk_rows;
    const int t_lo = max(tb, own_lo), t_hi = min(min(tb + chunk_rows, T), own_hi);
    const int nround = t_hi > t_lo ? (t_hi - t_lo) / HT_TK : 0;
    if (tid == 0) {
        for (int s = 0; s < NS; ++s) { mbar_init(bar_full + 8u * s, 1); mbar_init(bar_empty + 8u * s, HT_THREADS / 32); }
        mbar_init_fence();
        tma_prefetch_desc(&tmHt);
    }
    __syncthreads();
    const int row0 = b * T + t_lo;
    if (tid == 0)
        for (int r = 0; r < NS - 1 && r < nround; ++r) ht_request<LP, NS>(&tmHt, base, bar_full, r, row0, c0);
    const int g = lane >> 2, tig = lane & 3;
    float acc[2][NT][4];
#pragma unroll
    for (int ch = 0; ch < 2; ++ch)
#pragma unroll
        for (int tl = 0; tl < NT; ++tl)
#pragma unroll
            for (int k = 0; k < 4; ++k) acc[ch][tl][k] = 0.f;
    // element (c, l, t) of a stage: offset c * LP * 64 + l * 64 + t * 2, 16-byte chunk (t >> 3) ^= (l >> 1) & 3.
    // A lane reads l = 8 a + g, t = 16 ks + 2 tig + 8 h (two consecutive t = one 32-bit word): chunk (2 ks + h) ^ (g >> 1) & 3,
    // bank (g & 1) * 16 + chunk * 4 + tig: 32 distinct banks over (g, tig).
    const uint32_t lane_off = (uint32_t)(g * 64 + tig * 4), gx = (uint32_t)((g >> 1) & 3);
    for (int r = 0; r < nround; ++r) {
        const int s = r % NS;
        // round r + NS - 1 goes into the stage round r - 1 used: thread 0 waits until all 16 warps have released it
        if (tid == 0 && r + NS - 1 < nround) {
            if (r >= 1) mbar_wait(bar_empty + 8u * ((r - 1) % NS), (uint32_t)((r - 1) / NS) & 1u);
            ht_request<LP, NS>(&tmHt, base, bar_full, r + NS - 1, row0, c0);
        }
        __syncwarp();
        mbar_wait(bar_full + 8u * s, (uint32_t)(r / NS) & 1u);
        const uint32_t st = base + (uint32_t)s * HT_STAGE(LP);
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
            const uint32_t pl = st + (uint32_t)(warp * 2 + ch) * (uint32_t)(LP * 64) + lane_off;
#pragma unroll
            for (int ks = 0; ks < HT_TK / 16; ++ks) {
                uint32_t v[NA][2];                                         // V[g + 8 a][16 ks + 2 tig + 8 h .. + 1]
#pragma unroll
                for (int a = 0; a < NA; ++a)
#pragma unroll
                    for (int h = 0; h < 2; ++h)
                        v[a][h] = ht_lds(pl + (uint32_t)(a * 512) + ((((uint32_t)(2 * ks + h)) ^ gx) << 4));
                // tiles (row block of 16 layers, column block of 8 layers), upper triangle: (0,0..3), (1,2..3); LP = 16: (0,0..1)
                // A: a0 (row g, k 2 tig..), a1 (row g + 8), a2 (row g, k + 8), a3 (row g + 8, k + 8); B: b0 (k 2 tig.., n g), b1 (k + 8)
#pragma unroll
                for (int tl = 0; tl < NT; ++tl) {
                    const int mt = tl < 4 ? 0 : 1, j = tl < 4 ? tl : tl - 2;
                    ht_mma(acc[ch][tl], v[2 * mt][0], v[2 * mt + 1][0], v[2 * mt][1], v[2 * mt + 1][1], v[j][0], v[j][1]);
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_empty + 8u * s);
    }
    // C fragment: c0 (row g, col 2 tig), c1 (g, 2 tig + 1), c2 (g + 8, 2 tig), c3 (g + 8, 2 tig + 1).  Entries with l <= m are
    // written together with their mirror, so the partial is exactly symmetric.
#pragma unroll
    for (int ch = 0; ch < 2; ++ch) {
        float* out = part + (((size_t)b * nchunk + blockIdx.x) * AST_C + c0 + warp * 2 + ch) * L * L;
#pragma unroll
        for (int tl = 0; tl < NT; ++tl) {
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

// ---------------------------------------------------------------------------------------------------------------------
// Style gradient.  grid = (row blocks, 4 channel quads, B).  Per channel, stage and 16-row half mt of the stage:
// OUT[t][l] (16 x LP) = V^T (16 x LP) S.
//   A = V^T: ldmatrix.x4.trans of the blocks (m 0..7 | 8..15) x (t 0..7 | 8..15) of a 16-layer k-step; lane i addresses row
//     m = i & 7 (+ 8 for i >= 16) of chunk 2 mt + (i >> 3 & 1): within 8 lanes the 16-byte slot (m & 1) * 4 + (chunk ^ (m >> 1) & 3)
//     takes all 8 values -- conflict-free;
//   B = S: b0 = S[m = 16 ks + 2 tig .. + 1][l = 8 j + g], b1 = S[m + 8 ..][l], fp16 pairs in registers for the whole CTA;
//   output staging so[l * 548 + t * 33 + c]: bank 4 l + t + c = 8 tig (+4) + g (+8): distinct over a warp.
// Outputs leave as 128-byte rows of dF_l (32 lanes = 32 channels).
// ---------------------------------------------------------------------------------------------------------------------
#define HS_NS(LP) ((LP) == 32 ? 2 : 4)
#define HS_OPL 548
#define HS_OPT 33
#define HS_OUT_BYTES(LP) ((LP) * HS_OPL * 4)
#define HS_SMEM(LP) (HS_NS(LP) * HT_STAGE(LP) + HS_OUT_BYTES(LP) + 1024 + 64)

template <int LP>
__global__ void __launch_bounds__(HT_THREADS, 1) style_grad_ours_tma_kernel(const __grid_constant__ CUtensorMap tmHt,
                                                                           const __grid_constant__ OursPtrs P,
                                                                           const float* __restrict__ Ssym, int L, int T,
                                                                           int rows_per_cta) {
    constexpr int NS = HS_NS(LP), NK = LP / 16, NJ = LP / 8;               // stages; 16-layer k-steps; 8-layer column tiles
    extern __shared__ uint8_t hs_raw[];
    const uint32_t base = (smem_u32(hs_raw) + 1023u) & ~1023u;
    const uint32_t so_u32 = base + NS * HT_STAGE(LP);
    const uint32_t bar_full = so_u32 + HS_OUT_BYTES(LP);
    float* so = reinterpret_cast<float*>(hs_raw + (so_u32 - smem_u32(hs_raw)));
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int c0 = blockIdx.y * HT_CH, b = blockIdx.z;
    const int t_lo = blockIdx.x * rows_per_cta, t_hi = min(t_lo + rows_per_cta, T);
    const int nround = t_hi > t_lo ? (t_hi - t_lo) / HT_TK : 0;
    if (tid == 0) {
        for (int s = 0; s < NS; ++s) mbar_init(bar_full + 8u * s, 1);
        mbar_init_fence();
        tma_prefetch_desc(&tmHt);
    }
    __syncthreads();
    const int row0 = b * T + t_lo;
    if (tid == 0)
        for (int r = 0; r < NS && r < nround; ++r) ht_request<LP, NS>(&tmHt, base, bar_full, r, row0, c0);
    const int g = lane >> 2, tig = lane & 3;
    // coefficient block -> fp16 B fragments, scaled per channel by 2^-e with 2^e > max |S_c| (exact; undone on the outputs)
    uint32_t bf[2][NK][NJ][2];
    float unscale[2];
#pragma unroll
    for (int ch = 0; ch < 2; ++ch) {
        const float* Sc = Ssym + ((size_t)b * AST_C + c0 + warp * 2 + ch) * L * L;
        float sv[NK][NJ][2][2];
        float mx = 0.f;
#pragma unroll
        for (int ks = 0; ks < NK; ++ks)
#pragma unroll
            for (int j = 0; j < NJ; ++j)
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int m = 16 * ks + 2 * tig + 8 * h + e, l = 8 * j + g;
                        const float v = (m < L && l < L) ? Sc[l * L + m] : 0.f;
                        sv[ks][j][h][e] = v;
                        mx = fmaxf(mx, fabsf(v));
                    }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        int ex = 0;
        if (mx > 0.f && mx < 3.0e38f) frexpf(mx, &ex);                    // mx = f * 2^ex, 0.5 <= f < 1
        ex = max(-100, min(100, ex));
        const float sc = ldexpf(1.0f, -ex);
        unscale[ch] = ldexpf(1.0f, ex);
#pragma unroll
        for (int ks = 0; ks < NK; ++ks)
#pragma unroll
            for (int j = 0; j < NJ; ++j)
#pragma unroll
                for (int h = 0; h < 2; ++h) bf[ch][ks][j][h] = ht_pack_f16(sv[ks][j][h][0] * sc, sv[ks][j][h][1] * sc);
    }
    // ldmatrix row address of this lane: matrix q = lane >> 3: layers (q >> 1) * 8 + (lane & 7), t chunk (q & 1) (+ 2 mt)
    const int lm = ((lane >> 4) & 1) * 8 + (lane & 7);
    const uint32_t lm_off = (uint32_t)(lm * 64), lm_x = (uint32_t)((lm >> 1) & 3), lm_c = (uint32_t)((lane >> 3) & 1);
    // C fragment: c0 (t = g, l = 8 j + 2 tig), c1 (t, l + 1), c2 (t + 8, l), c3 (t + 8, l + 1)
    float* so_w = so + (2 * tig) * HS_OPL + g * HS_OPT + warp * 2;
    const int oc = tid & 31, ot = tid >> 5;                                 // output pass: this thread's channel and row of a half stage
    const float* so_r = so + ot * HS_OPT + oc;
    for (int r = 0; r < nround; ++r) {
        const int s = r % NS;
        mbar_wait(bar_full + 8u * s, (uint32_t)(r / NS) & 1u);
        const uint32_t st = base + (uint32_t)s * HT_STAGE(LP);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
            float acc[2][NJ][4];
#pragma unroll
            for (int ch = 0; ch < 2; ++ch) {
#pragma unroll
                for (int j = 0; j < NJ; ++j)
#pragma unroll
                    for (int k = 0; k < 4; ++k) acc[ch][j][k] = 0.f;
                const uint32_t pl = st + (uint32_t)(warp * 2 + ch) * (uint32_t)(LP * 64) + lm_off;
#pragma unroll
                for (int ks = 0; ks < NK; ++ks) {
                    uint32_t a[4];                                         // a0 (t g, m 2 tig..), a1 (t g + 8), a2 (m + 8), a3 (t g + 8, m + 8)
                    ht_ldmatrix_x4_trans(a, pl + (uint32_t)(ks * 1024) + ((((uint32_t)(2 * mt) + lm_c) ^ lm_x) << 4));
#pragma unroll
                    for (int j = 0; j < NJ; ++j) ht_mma(acc[ch][j], a[0], a[1], a[2], a[3], bf[ch][ks][j][0], bf[ch][ks][j][1]);
                }
            }
            ht_bar_all();                             // the previous outputs have been read; mt = 1: every warp is done with the stage
            if (mt == 1 && tid == 0 && r + NS < nround) ht_request<LP, NS>(&tmHt, base, bar_full, r + NS, row0, c0);   // refilled while the outputs leave
#pragma unroll
            for (int ch = 0; ch < 2; ++ch)
#pragma unroll
                for (int j = 0; j < NJ; ++j) {
                    float* d = so_w + (8 * j) * HS_OPL + ch;
                    d[0] = acc[ch][j][0] * unscale[ch];
                    d[HS_OPL] = acc[ch][j][1] * unscale[ch];
                    d[8 * HS_OPT] = acc[ch][j][2] * unscale[ch];
                    d[HS_OPL + 8 * HS_OPT] = acc[ch][j][3] * unscale[ch];
                }
            ht_bar_all();
            {                                                             // 32 lanes = one 128-byte row of dF_l
                const size_t off = ((size_t)b * T + t_lo + r * HT_TK + mt * 16 + ot) * AST_C + c0 + oc;
#pragma unroll 6
                for (int l = 0; l < L; ++l) P.dF[l][off] = so_r[l * HS_OPL];
            }
        }
    }
}

// ---- host side ------------------------------------------------------------------------------------------------------
typedef CUresult (*ht_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static inline int ours_tma_lp(int L) { return L <= 16 ? 16 : 32; }
int ours_tma_supported(int L) { return L >= 1 && L <= 32; }
size_t ours_tma_bytes(int L, int R) { return (size_t)L * R * AST_C * sizeof(__half); }

// Ht = [R/32][L][128 channels][32 rows] fp16.  Tensor-map dimensions in the order (row % 32, layer, channel, row / 32): the
// box lands in shared memory as [channel][layer][t].
int ours_tma_make_map(OursTma* o, void* Ht, int L, int R) {
    static ht_encode_fn enc = nullptr;
    if (!enc) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPointByVersion("cuTensorMapEncodeTiled", &fn, 12000, cudaEnableDefault, &qres) != cudaSuccess ||
            qres != cudaDriverEntryPointSuccess || !fn)
            return -1;
        enc = reinterpret_cast<ht_encode_fn>(fn);
    }
    static_assert(sizeof(o->map) >= sizeof(CUtensorMap), "OursTma::map holds a CUtensorMap");
    const cuuint64_t run = HT_TK * sizeof(__half);                         // 64 bytes: 32 rows of one (layer, channel)
    const cuuint64_t dims[4] = {HT_TK, (cuuint64_t)L, (cuuint64_t)AST_C, (cuuint64_t)(R / HT_TK)};
    const cuuint64_t str[3] = {(cuuint64_t)AST_C * run, run, (cuuint64_t)L * AST_C * run};
    const cuuint32_t box[4] = {HT_TK, (cuuint32_t)ours_tma_lp(L), HT_CH, 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    if (enc(reinterpret_cast<CUtensorMap*>(o->map), CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, Ht, dims, str, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return -1;
    if (cudaFuncSetAttribute(gram_ours_tma_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, HG_SMEM(32)) != cudaSuccess ||
        cudaFuncSetAttribute(gram_ours_tma_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, HG_SMEM(16)) != cudaSuccess ||
        cudaFuncSetAttribute(style_grad_ours_tma_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, HS_SMEM(32)) != cudaSuccess ||
        cudaFuncSetAttribute(style_grad_ours_tma_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, HS_SMEM(16)) != cudaSuccess)
        return -1;
    return 0;
}

int ours_tma_chunk_rows(int T, int target) {
    const int rows = (T + target - 1) / target;
    return std::max(HT_TK, (rows + HT_TK - 1) / HT_TK * HT_TK);
}

void launch_gram_ours_tma(const OursTma& o, int L, int T, int B, int own_lo, int own_hi, float* part, int nchunk,
                          int chunk_rows, cudaStream_t s) {
    dim3 grid(nchunk, AST_C / HT_CH, B);
    const CUtensorMap& tm = *reinterpret_cast<const CUtensorMap*>(o.map);
    if (ours_tma_lp(L) == 32) gram_ours_tma_kernel<32><<<grid, HT_THREADS, HG_SMEM(32), s>>>(tm, L, T, own_lo, own_hi, part, nchunk, chunk_rows);
    else                      gram_ours_tma_kernel<16><<<grid, HT_THREADS, HG_SMEM(16), s>>>(tm, L, T, own_lo, own_hi, part, nchunk, chunk_rows);
    ast_launch_counter += 1;
}

void launch_style_grad_ours_tma(const OursTma& o, const OursPtrs& ptrs, const float* Ssym, int L, int T, int B, cudaStream_t s) {
    const int nx = std::max(1, 148 / (4 * B));                            // one CTA per SM, one wave
    const int rows = std::max(HT_TK, ((T + nx - 1) / nx + HT_TK - 1) / HT_TK * HT_TK);
    dim3 grid((T + rows - 1) / rows, AST_C / HT_CH, B);
    const CUtensorMap& tm = *reinterpret_cast<const CUtensorMap*>(o.map);
    if (ours_tma_lp(L) == 32) style_grad_ours_tma_kernel<32><<<grid, HT_THREADS, HS_SMEM(32), s>>>(tm, ptrs, Ssym, L, T, rows);
    else                      style_grad_ours_tma_kernel<16><<<grid, HT_THREADS, HS_SMEM(16), s>>>(tm, ptrs, Ssym, L, T, rows);
    ast_launch_counter += 1;
}
