// gram_ours_tma.cu -- channel-wise ("ours") Gram and style gradient (1 <= L <= 32 style layers: one stack = the reference's
// --stack 0/1/2, BASELINE configs[0]; all 30 layers = the CLI default, configs[4]) fed by TMA from TRANSPOSED extract
// copies.  Reference: methods.py:62-73
//   G[c,l,m] = sum_t S[l,t,c] S[m,t,c]        dF_l[t,c] = sum_m Ssym[c,l,m] F_m[t,c]
//
// Both are per-channel products over V_c = [layer][t].  The extracts leave the encoder as [t][channel] rows, so round 1/2
// (gram_ours_tc.cu, kept for the per-layer kernel path) transposed every element on its way into shared memory: through
// registers and four scattered 4-byte shared stores (Gram), or one 4-byte cp.async per element (gradient).  ncu, round 2:
// the Gram kernel at 2.8 TB/s, the gradient kernel latency-bound at 25 % occupancy, 2.2 TB/s.  Here the whole-stack forward
// kernel writes a second copy of every style extract as Ht[row / 16][slot][channel][row % 16] (stack_fwd.cuh finish(): a
// warp's 32 TMEM lanes are 32 consecutive rows, so one column of a warp is two 64-byte runs), and both kernels stream it
// with one TMA box per stage: {16 t, 32 layers, 32 channels} fp32 = 64 KB, SWIZZLE_64B, layers L..31 zero-filled by the
// out-of-bounds rule.  The 16-row blocking keeps DRAM pages busy on both sides: a stage is 30 runs of 2 KB (32 channels
// x 64 B) 8 KB apart, a tile-layer of the forward writes 8 runs of 8 KB.  (First version: Ht[slot][channel][row] -- a stage
// was then 960 runs of 64 B, 64 KB apart, and both kernels stalled at 1.7 TB/s.)  No thread touches an operand before it is an mma.sync fragment load:
//   Gram:      A (rows = layer, cols = t) and B (k = t, n = layer) are the same 8 loads per k-step, 6 upper-triangular tiles;
//   gradient:  OUT[t][l] = V^T S, the coefficient block S as register-resident TF32 B fragments.
// Fragment loads are conflict-free under the 64-byte swizzle (16-byte chunk index ^= bits 7..8 of the offset): see the
// address comments.  Thread 0 keeps the ring full (full / empty mbarriers); 16 warps own 2 channels each.
// Operands are single-pass TF32 (cvt.rna), fp32 accumulate -- the same arithmetic class as gram_ours_tc.cu.
#include "common.cuh"
#include "ptx.cuh"

#include <cuda.h>
#include <algorithm>

#define HT_TK 16                              // t per stage
#define HT_CH 32                              // channels per CTA
// LP = layers per stage: 32 (13..32 style layers) or 16 (one stack, the reference's --stack 0/1/2: <= 16); a stage is
// [channel 32][layer LP][t 16] fp32 = LP * 2048 bytes
#define HT_STAGE(LP) (HT_CH * (LP) * HT_TK * 4)
#define HT_CONSUMERS 512                      // 16 warps x 2 channels
#define HT_THREADS HT_CONSUMERS               // no producer warp: a 17th warp would cap the kernel at 96 registers per thread

__device__ __forceinline__ uint32_t ht_tf32(float x) {
    uint32_t u;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
    return u;
}
__device__ __forceinline__ void ht_mma(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ht_bar_consumers() { asm volatile("bar.sync 1, %0;" ::"n"(HT_CONSUMERS) : "memory"); }
__device__ __forceinline__ float ht_lds(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}

// Thread 0 requests round r into stage r % NS (the caller has made sure the stage is free).
template <int LP, int NS>
__device__ __forceinline__ void ht_request(const CUtensorMap* tm, uint32_t stage0, uint32_t bar_full, int r, int row0, int c0) {
    const int s = r % NS;
    mbar_expect_tx(bar_full + 8u * s, HT_STAGE(LP));
    tma_load_4d(stage0 + (uint32_t)s * HT_STAGE(LP), tm, bar_full + 8u * s, 0, 0, c0, (row0 >> 4) + r);
}

// ---------------------------------------------------------------------------------------------------------------------
// Gram.  grid = (nchunk, 4 channel quads, B); part[b][chunk][c][l][m] like the other channel-wise Gram kernels.
// chunk_rows, own_lo, own_hi and T are multiples of 16: every round is a full stage.
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
        for (int s = 0; s < NS; ++s) { mbar_init(bar_full + 8u * s, 1); mbar_init(bar_empty + 8u * s, HT_CONSUMERS / 32); }
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
    // element (c, l, t) of a stage: offset c * LP * 64 + l * 64 + t * 4, 16-byte chunk (t >> 2) ^= (l >> 1) & 3.
    // A lane reads l = 8 a + g, t = 8 ks + tig + 4 bb: chunk (2 ks + bb) ^ (g >> 1) & 3, bank (g & 1) * 16 + chunk * 4 + tig:
    // 32 distinct banks over (g, tig).
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
            for (int ks = 0; ks < HT_TK / 8; ++ks) {
                uint32_t v[NA][2];                                         // V[g + 8 a][8 ks + tig + 4 bb]
#pragma unroll
                for (int a = 0; a < NA; ++a)
#pragma unroll
                    for (int bb = 0; bb < 2; ++bb)
                        v[a][bb] = ht_tf32(ht_lds(pl + (uint32_t)(a * 512) + ((((uint32_t)(2 * ks + bb)) ^ gx) << 4)));
                // tiles (row block of 16 layers, column block of 8 layers), upper triangle: (0,0..3), (1,2..3); LP = 16: (0,0..1)
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
// Style gradient.  grid = (row blocks, 4 channel quads, B).  Per channel and stage: OUT[t][l] (16 x 32) = V^T (16 x 32) S.
// MMA row g of the stage is t(g) = (g >> 2) * 8 + (g & 3), row g + 8 is t(g) + 4; MMA column n of tile j is layer
// 8 j + perm(n), perm = 0 2 1 3 4 6 5 7.  Both permutations only exist to make the shared-memory accesses conflict-free:
//   A loads: m = 8 ks + tig (+ 4), chunk ((g >> 2) * 2 + k) ^ ((tig >> 1) + 2 h), bank (tig & 1) * 16 + chunk * 4 + (g & 3);
//   output staging so[l * 548 + t * 33 + c]: bank 4 l + t + c, with l in {0,1,4,5} (+2) + 8 j and t in {0..3, 8..11} (+4).
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
    constexpr int NS = HS_NS(LP), NK = LP / 8;                             // stages; 8-layer blocks (k-steps and column tiles)
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
    const int pg = (g & 4) | ((g & 1) << 1) | ((g >> 1) & 1);              // perm(g): 0 2 1 3 4 6 5 7
    // B fragments: b0 = S[m = 8 ks + tig][l = 8 j + perm(g)], b1 = S[m + 4][l]
    uint32_t bf[2][NK][NK][2];
#pragma unroll
    for (int ch = 0; ch < 2; ++ch) {
        const float* Sc = Ssym + ((size_t)b * AST_C + c0 + warp * 2 + ch) * L * L;
#pragma unroll
        for (int ks = 0; ks < NK; ++ks)
#pragma unroll
            for (int j = 0; j < NK; ++j)
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int m = 8 * ks + tig + 4 * h, l = 8 * j + pg;
                    bf[ch][ks][j][h] = ht_tf32(m < L && l < L ? Sc[l * L + m] : 0.f);
                }
    }
    const int tg = (g >> 2) * 8 + (g & 3);                                 // t of MMA row g (row g + 8: tg + 4)
    // A(row t, col m) = V[m][t]: a0 (tg, m = 8 ks + tig), a1 (tg + 4, m), a2 (tg, m + 4), a3 (tg + 4, m + 4)
    const uint32_t a_lane = (uint32_t)(tig * 64 + (g & 3) * 4);
    const uint32_t ch0 = (uint32_t)((g >> 2) * 2), xl = (uint32_t)(tig >> 1);
    // C fragment: c0 (tg, l0), c1 (tg, l0 + 2), c2 (tg + 4, l0), c3 (tg + 4, l0 + 2), l0 = 8 j + (tig & 1) + 4 (tig >> 1)
    const int l0 = (tig & 1) + 4 * (tig >> 1);
    float* so_w = so + l0 * HS_OPL + tg * HS_OPT + warp * 2;
    const int oc = tid & 31, ot = tid >> 5;                                 // output pass: this thread's channel and row of a stage
    const float* so_r = so + ot * HS_OPT + oc;
    for (int r = 0; r < nround; ++r) {
        const int s = r % NS;
        mbar_wait(bar_full + 8u * s, (uint32_t)(r / NS) & 1u);
        const uint32_t st = base + (uint32_t)s * HT_STAGE(LP);
        float acc[2][NK][4];
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
#pragma unroll
            for (int j = 0; j < NK; ++j)
#pragma unroll
                for (int k = 0; k < 4; ++k) acc[ch][j][k] = 0.f;
            const uint32_t pl = st + (uint32_t)(warp * 2 + ch) * (uint32_t)(LP * 64) + a_lane;
#pragma unroll
            for (int ks = 0; ks < NK; ++ks) {
                uint32_t a[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {                              // k & 1: row + 8 (t + 4), k >> 1: m + 4
                    const uint32_t chunk = (ch0 + (uint32_t)(k & 1)) ^ ((xl + 2u * (uint32_t)(k >> 1)) & 3u);
                    a[k] = ht_tf32(ht_lds(pl + (uint32_t)((8 * ks + 4 * (k >> 1)) * 64) + (chunk << 4)));
                }
#pragma unroll
                for (int j = 0; j < NK; ++j) ht_mma(acc[ch][j], a[0], a[1], a[2], a[3], bf[ch][ks][j][0], bf[ch][ks][j][1]);
            }
        }
        ht_bar_consumers();                       // every warp is done with the stage, and with reading the previous round's outputs
        if (tid == 0 && r + NS < nround) ht_request<LP, NS>(&tmHt, base, bar_full, r + NS, row0, c0);   // refilled while the outputs leave
#pragma unroll
        for (int ch = 0; ch < 2; ++ch)
#pragma unroll
            for (int j = 0; j < NK; ++j) {
                float* d = so_w + (8 * j) * HS_OPL + ch;
                d[0] = acc[ch][j][0];
                d[2 * HS_OPL] = acc[ch][j][1];
                d[4 * HS_OPT] = acc[ch][j][2];
                d[2 * HS_OPL + 4 * HS_OPT] = acc[ch][j][3];
            }
        ht_bar_consumers();
        {                                                                 // 32 lanes = one 128-byte row of dF_l
            const size_t off = ((size_t)b * T + t_lo + r * HT_TK + ot) * AST_C + c0 + oc;
#pragma unroll 6
            for (int l = 0; l < L; ++l) P.dF[l][off] = so_r[l * HS_OPL];
        }
    }
}

// ---- host side ------------------------------------------------------------------------------------------------------
typedef CUresult (*ht_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// Ht = [R/16][L][128 channels][16 rows] fp32.  Tensor-map dimensions in the order (row % 16, layer, channel, row / 16): the
// box lands in shared memory as [channel][layer][t].
static inline int ours_tma_lp(int L) { return L <= 16 ? 16 : 32; }
int ours_tma_supported(int L) { return L >= 1 && L <= 32; }

int ours_tma_make_map(OursTma* o, float* Ht, int L, int R) {
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
    const cuuint64_t dims[4] = {HT_TK, (cuuint64_t)L, (cuuint64_t)AST_C, (cuuint64_t)(R / HT_TK)};
    const cuuint64_t str[3] = {(cuuint64_t)AST_C * HT_TK * sizeof(float), (cuuint64_t)HT_TK * sizeof(float),
                               (cuuint64_t)L * AST_C * HT_TK * sizeof(float)};
    const cuuint32_t box[4] = {HT_TK, (cuuint32_t)ours_tma_lp(L), HT_CH, 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    if (enc(reinterpret_cast<CUtensorMap*>(o->map), CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, Ht, dims, str, box, estr,
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
