// kernels_tc.cu -- the dilated-conv stack on 5th-gen tensor cores (sm_100a only): tcgen05.mma with TMEM
// accumulators, every operand AND every activation tile moved by TMA (128B swizzle), warp-specialised
// (warp 0 = TMA producer, warp 1 = MMA issuer, warps 2-17 = epilogue: TMEM lane quadrant x 32-column group).
// One CTA = one 128-row time tile of one layer, two chained GEMMs:
//   forward  (model.py:99-114):  z = conv3_d(relu(h)) + bd ; h' = h + relu(z) Wr + br
//   backward :  g += conv3_d^T(dz_l) * [h_l>0] + dF ; dz_{l-1} = (g Wr_{l-1}^T) * [z_{l-1}>0]
// The K=3 dilated conv is an implicit GEMM: three TMA loads of the same activation matrix at row
// offsets t0-d, t0, t0+d; rows outside [0,T) of the pair are zero-filled by TMA, which IS the
// reference's per-layer 'SAME' zero padding (masked.py:139).  A tap whose whole tile is padding is
// skipped (no load, no MMA).
// Forward operands are 2-term bf16 splits (x = hi + lo, products hi*hi + lo*hi + hi*lo): plain bf16
// (or fp16 / tf32) forward activations flip too many ReLU signs through 30 layers for the <=1e-2
// gradient tier (oracle/precision_study.py, DESIGN.md "accuracy tiers").
// Epilogues never touch global memory row-per-thread: fp32 tiles (residual stream, running gradient,
// head gradient) arrive by TMA into swizzled smem, are updated in place by the thread that owns the
// row (= its TMEM lane), and leave by TMA store; only the 16-byte-per-row sign masks use plain
// (fully coalesced) global accesses.
#include "tc.cuh"
#include "ptx.cuh"

#include <cuda.h>
#include <cuda_fp16.h>
#include <cudaTypedefs.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

#define TILE16 16384                            // [128 rows][128 B]: one bf16 K-half or one fp32 32-column group
static constexpr uint32_t IDESC = umma_idesc_bf16(128, 128);

#ifdef AST_EXPERIMENTS
// optional per-CTA timeline (profiling aid, scripts/timeline_stack.py; -DAST_EXPERIMENTS builds only): slot i of CTA c <- clock64()
__device__ unsigned long long* g_timeline = nullptr;
#define TL(i) do { if (g_timeline) g_timeline[(size_t)blockIdx.x * 64 + (i)] = clock64(); } while (0)
void tc_set_timeline(unsigned long long* d_buf) { cudaMemcpyToSymbol(g_timeline, &d_buf, sizeof(d_buf)); }
void tc_set_tl_layer(int l);
// fault injection for tests/test_gpu_loud.py: tile `g_dbg_skip_publish - 1` never releases its hand-off flags
__device__ int g_dbg_skip_publish = 0;
void tc_set_dbg_skip_publish(int v) { cudaMemcpyToSymbol(g_dbg_skip_publish, &v, sizeof(v)); }
// tests: pretend the GPU has only this many SMs when planning segments / groups (0 = the real count), so that a clip
// that also fits ONE launch runs as several groups and must reproduce the one-launch result bit for bit
static int g_dbg_group_tiles = 0;
void tc_set_dbg_group_tiles(int v) { g_dbg_group_tiles = v; }
#define TC_GROUP_CAP(s) (g_dbg_group_tiles > 0 ? g_dbg_group_tiles : (s)->n_sm)
#else
#define TC_GROUP_CAP(s) ((s)->n_sm)
#define TL(i) do { } while (0)
#endif

__device__ __forceinline__ bool tap_live(int t0, int k, int shift, int T) {
    const int t = t0 + (k - 1) * shift;
    return t + 128 > 0 && t < T;
}

// =============================================================================================
// forward layer
// =============================================================================================
struct alignas(64) FwdArgs {
    CUtensorMap tmA_hi, tmA_lo;      // relu(h) hi / lo, bf16 [2][B][T][128], box {64,128,1,1} (loads and stores)
    CUtensorMap tmH;                 // fp32 residual stream [NL+1][B][T][128], box {32,128,1,1} (loads and stores)
    CUtensorMap tmW1, tmW2;          // weight operands, rows of 128 k-contiguous bf16, box {64,128}
    CUtensorMap tmF;                 // bf16 copy of the extract for the Gram head [n_style][B][T][128], box {64,128,1,1}
    int w1_row, w2_row;              // first row of this layer's block in tmW1 / tmW2
    int f_slot;                      // style slot of extracts[l] (row of tmF), or -1
    int T, shift, l, a_in, a_out;    // rows per pair; dilation; layer; ping-pong indices of the bf16 operand buffers
    const float* bias1; const float* bias2;
    uint32_t* zmask_out; uint32_t* hmask_out;
};
// smem: slot 0 | slot 1 | slot 2 (64 KB each: A_hi, A_lo, W_hi, W_lo x [128][64] bf16).  Slot 2 is a pipeline
// stage only during GEMM1; afterwards it is RT: relu(z) hi/lo (A operand of the 1x1 GEMM), then the staging
// area of relu(h') hi/lo.  After GEMM1 the A half of slots 0/1 receives the fp32 H tile (two 32-column groups
// each), updated in place to H' and stored from there; their W_hi quarter, dead after GEMM2, stages bf16(h').
#define FWD_STAGE 65536
#define FWD_SMEM (3 * FWD_STAGE + 1024 + 128)
#define FWD_THREADS 576                          // warp 0 TMA, warp 1 MMA, 16 epilogue warps (quadrant x column group)
// barrier block
#define FB_FULL(i)  (8u * (i))
#define FB_EMPTY(i) (24u + 8u * (i))
#define FB_ACC1 48u
#define FB_RT(j) (56u + 8u * (j))
#define FB_ACC2 72u
#define FB_TMEM 80u

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
// relu'd (or raw) 32 floats -> 4 x 16-byte chunks of bf16 hi (and lo = v - hi) in a swizzled [128][64] tile pair
__device__ __forceinline__ void stage_hi_lo(uint8_t* tile_hi, uint8_t* tile_lo, int r, int chunk0, const float (&v)[32]) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t off = sw128_off(r, chunk0 + i);
        float hi[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) hi[e] = bf16_round(v[8 * i + e]);
        *reinterpret_cast<uint4*>(tile_hi + off) = make_uint4(pack_bf16(hi[0], hi[1]), pack_bf16(hi[2], hi[3]),
                                                             pack_bf16(hi[4], hi[5]), pack_bf16(hi[6], hi[7]));
        *reinterpret_cast<uint4*>(tile_lo + off) =
            make_uint4(pack_bf16(v[8 * i] - hi[0], v[8 * i + 1] - hi[1]), pack_bf16(v[8 * i + 2] - hi[2], v[8 * i + 3] - hi[3]),
                       pack_bf16(v[8 * i + 4] - hi[4], v[8 * i + 5] - hi[5]), pack_bf16(v[8 * i + 6] - hi[6], v[8 * i + 7] - hi[7]));
    }
}
__device__ __forceinline__ void stage_bf16(uint8_t* tile, int r, int chunk0, const float (&v)[32]) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
        *reinterpret_cast<uint4*>(tile + sw128_off(r, chunk0 + i)) =
            make_uint4(pack_bf16(v[8 * i], v[8 * i + 1]), pack_bf16(v[8 * i + 2], v[8 * i + 3]),
                       pack_bf16(v[8 * i + 4], v[8 * i + 5]), pack_bf16(v[8 * i + 6], v[8 * i + 7]));
}

__global__ void __launch_bounds__(FWD_THREADS, 1) tc_fwd_kernel(const __grid_constant__ FwdArgs p) {
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t base = (smem_u32(tc_smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = tc_smem_raw + (base - smem_u32(tc_smem_raw));          // generic pointer to `base`
    const uint32_t s_rt = base + 2 * FWD_STAGE, bar = base + 3 * FWD_STAGE;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row0 = blockIdx.x * 128;
    const int b = row0 / p.T, t0 = row0 - b * p.T;
    const int nlive = (int)tap_live(t0, 0, p.shift, p.T) + 1 + (int)tap_live(t0, 2, p.shift, p.T);
    const int n1 = 2 * nlive;                                              // GEMM1 K-slices of 64
    if (threadIdx.x == 0) TL(0);

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.tmA_hi); tma_prefetch_desc(&p.tmA_lo); tma_prefetch_desc(&p.tmW1); tma_prefetch_desc(&p.tmW2);
        tma_prefetch_desc(&p.tmH);
        if (p.f_slot >= 0) tma_prefetch_desc(&p.tmF);
        for (int i = 0; i < 3; ++i) { mbar_init(bar + FB_FULL(i), 1); mbar_init(bar + FB_EMPTY(i), 1); }
        mbar_init(bar + FB_ACC1, 1); mbar_init(bar + FB_RT(0), 8); mbar_init(bar + FB_RT(1), 8); mbar_init(bar + FB_ACC2, 1);
        mbar_init_fence();
    }
    if (warp == 1) tmem_alloc<256>(bar + FB_TMEM);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gen + (bar + FB_TMEM - base));
    const uint32_t acc1 = tmem, acc2 = tmem + 128;
    pdl_launch_dependents();
    // Everything above overlapped the previous kernel's tail.  The producer thread goes further before it waits:
    // weight tiles are constants, so the W half of the first three stages is requested ahead of griddepcontrol.wait.
    if (threadIdx.x != 0) pdl_wait();

    if (warp == 0) {
        // ================================ TMA producer ========================================
        if (lane == 0) {
            int fills[3] = {0, 0, 0};
            int tap_of[6];
            {
                int n = 0;
                for (int k = 0; k < 3; ++k)
                    if (tap_live(t0, k, p.shift, p.T)) { tap_of[n++] = k; tap_of[n++] = k; }
            }
            auto load_w = [&](int it) {                      // arms the stage and requests its two weight tiles
                const int slot = it % 3, k = tap_of[it], j = it & 1;
                const uint32_t st = base + slot * FWD_STAGE, fb = bar + FB_FULL(slot);
                mbar_expect_tx(fb, 4 * TILE16);
                tma_load_2d(st + 2 * TILE16, &p.tmW1, fb, j * 64, p.w1_row + (k * 2 + 0) * 128);
                tma_load_2d(st + 3 * TILE16, &p.tmW1, fb, j * 64, p.w1_row + (k * 2 + 1) * 128);
            };
            const int npre = n1 < 3 ? n1 : 3;
            for (int it = 0; it < npre; ++it) load_w(it);
            pdl_wait();                                      // activations of the previous layer are now complete and visible
            TL(1);
            for (int it = 0; it < n1; ++it) {
                const int slot = it % 3, k = tap_of[it], j = it & 1;
                const int t = t0 + (k - 1) * p.shift;
                if (it >= npre) {
                    mbar_wait(bar + FB_EMPTY(slot), (fills[slot] & 1) ^ 1);
                    load_w(it);
                }
                ++fills[slot];
                const uint32_t st = base + slot * FWD_STAGE, fb = bar + FB_FULL(slot);
                tma_load_4d(st, &p.tmA_hi, fb, j * 64, t, b, p.a_in);
                tma_load_4d(st + TILE16, &p.tmA_lo, fb, j * 64, t, b, p.a_in);
                TL(2 + it);
            }
            // 1x1 weights (K-half j) + two column groups of the fp32 residual tile, into slots 0 and 1
            for (int j = 0; j < 2; ++j) {
                mbar_wait(bar + FB_EMPTY(j), (fills[j] & 1) ^ 1);
                ++fills[j];
                const uint32_t st = base + j * FWD_STAGE, fb = bar + FB_FULL(j);
                mbar_expect_tx(fb, 4 * TILE16);
                tma_load_2d(st + 2 * TILE16, &p.tmW2, fb, j * 64, p.w2_row);
                tma_load_2d(st + 3 * TILE16, &p.tmW2, fb, j * 64, p.w2_row + 128);
                tma_load_4d(st, &p.tmH, fb, (2 * j) * 32, t0, b, p.l);
                tma_load_4d(st + TILE16, &p.tmH, fb, (2 * j + 1) * 32, t0, b, p.l);
                TL(8 + j);
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ==========================================
        if (lane == 0) {
            int uses[3] = {0, 0, 0};
            uint32_t accum = 0;
            for (int it = 0; it < n1; ++it) {
                const int slot = it % 3;
                mbar_wait(bar + FB_FULL(slot), uses[slot] & 1);
                ++uses[slot];
                tc_fence_after();
                TL(10 + it);
                const uint32_t st = base + slot * FWD_STAGE;
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    const uint64_t a_hi = umma_desc_kmajor_sw128(st + kk * 32);
                    const uint64_t a_lo = umma_desc_kmajor_sw128(st + TILE16 + kk * 32);
                    const uint64_t w_hi = umma_desc_kmajor_sw128(st + 2 * TILE16 + kk * 32);
                    const uint64_t w_lo = umma_desc_kmajor_sw128(st + 3 * TILE16 + kk * 32);
                    umma_f16(acc1, a_hi, w_hi, IDESC, accum);
                    umma_f16(acc1, a_lo, w_hi, IDESC, 1u);
                    umma_f16(acc1, a_hi, w_lo, IDESC, 1u);
                    accum = 1u;
                }
                umma_commit(bar + FB_EMPTY(slot));
            }
            umma_commit(bar + FB_ACC1);
            TL(18);
            for (int j = 0; j < 2; ++j) {
                mbar_wait(bar + FB_RT(j), 0);
                mbar_wait(bar + FB_FULL(j), uses[j] & 1);
                tc_fence_after();
                TL(16 + j * 3);
                const uint32_t st = base + j * FWD_STAGE;
                const uint32_t ra = s_rt + j * 2 * TILE16;                  // [K-half j][hi, lo]
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    const uint64_t a_hi = umma_desc_kmajor_sw128(ra + kk * 32);
                    const uint64_t a_lo = umma_desc_kmajor_sw128(ra + TILE16 + kk * 32);
                    const uint64_t w_hi = umma_desc_kmajor_sw128(st + 2 * TILE16 + kk * 32);
                    const uint64_t w_lo = umma_desc_kmajor_sw128(st + 3 * TILE16 + kk * 32);
                    umma_f16(acc2, a_hi, w_hi, IDESC, (j > 0 || kk > 0) ? 1u : 0u);
                    umma_f16(acc2, a_lo, w_hi, IDESC, 1u);
                    umma_f16(acc2, a_hi, w_lo, IDESC, 1u);
                }
            }
            umma_commit(bar + FB_ACC2);
            TL(20);
        }
        __syncwarp();
    } else {
        // ================================ epilogue warps ======================================
        // warp -> (TMEM lane quadrant q = warp % 4, 32-column group cc): 4 warps per scheduler hide each other's latency
        const int q = warp & 3, cc = (warp - 2) >> 2, half = cc >> 1;
        const int r = q * 32 + lane;                              // row within the tile = TMEM lane
        const size_t row = (size_t)row0 + r;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        uint8_t* rt = gen + 2 * FWD_STAGE + half * 2 * TILE16;    // [K-half][hi, lo]
        float v[32];

        // ---- epilogue 1: z = acc1 + bd -> sign mask, relu(z) hi/lo -> A operand of the 1x1 GEMM ----
        mbar_wait(bar + FB_ACC1, 0);
        tc_fence_after();
        if (threadIdx.x == 64) TL(24);
        tmem_ld32(acc1 + lane_off + cc * 32, v);
        {
            uint32_t word = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float4 bb = __ldg(reinterpret_cast<const float4*>(p.bias1 + cc * 32) + i);
                v[4 * i] += bb.x; v[4 * i + 1] += bb.y; v[4 * i + 2] += bb.z; v[4 * i + 3] += bb.w;
            }
#pragma unroll
            for (int i = 0; i < 32; ++i) { word |= (v[i] > 0.f ? 1u : 0u) << i; v[i] = fmaxf(v[i], 0.f); }
            stage_hi_lo(rt, rt + TILE16, r, (cc & 1) * 4, v);
            fence_proxy_async_smem();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar + FB_RT(half));
            p.zmask_out[row * AST_MW + cc] = word;
        }
        if (threadIdx.x == 64) TL(25);

        // ---- epilogue 2: h' = h + (acc2 + br) in place in smem; bf16(h'); relu(h') hi/lo -> RT; mask ----
        mbar_wait(bar + FB_ACC2, 0);
        tc_fence_after();
        if (threadIdx.x == 64) TL(26);
        // residual tile landed: last fill of slot `half` (fill index = its number of GEMM1 slices)
        mbar_wait(bar + FB_FULL(half), (uint32_t)((n1 - half + 2) / 3) & 1u);
        if (threadIdx.x == 64) TL(27);
        tmem_ld32(acc2 + lane_off + cc * 32, v);
        if (threadIdx.x == 64) TL(40);
        {
            uint8_t* hcg = gen + half * FWD_STAGE + (cc & 1) * TILE16;           // fp32 column group cc
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float4* hp = reinterpret_cast<float4*>(hcg + sw128_off(r, i));
                const float4 h = *hp;
                const float4 bb = __ldg(reinterpret_cast<const float4*>(p.bias2 + cc * 32) + i);
                // h' = h + (r Wr + br): same association as the reference (model.py:109-114)
                v[4 * i] = h.x + (v[4 * i] + bb.x); v[4 * i + 1] = h.y + (v[4 * i + 1] + bb.y);
                v[4 * i + 2] = h.z + (v[4 * i + 2] + bb.z); v[4 * i + 3] = h.w + (v[4 * i + 3] + bb.w);
                *hp = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
            }
            if (threadIdx.x == 64) TL(41);
            if (p.f_slot >= 0) stage_bf16(gen + half * FWD_STAGE + 2 * TILE16, r, (cc & 1) * 4, v);
            uint32_t word = 0;
#pragma unroll
            for (int i = 0; i < 32; ++i) { word |= (v[i] > 0.f ? 1u : 0u) << i; v[i] = fmaxf(v[i], 0.f); }
            if (threadIdx.x == 64) TL(42);
            stage_hi_lo(rt, rt + TILE16, r, (cc & 1) * 4, v);
            if (threadIdx.x == 64) TL(43);
            p.hmask_out[row * AST_MW + cc] = word;
        }
        if (threadIdx.x == 64) TL(44);
        fence_proxy_async_smem();
        if (threadIdx.x == 64) TL(28);
        // the 8 warps of this 64-column half are done: one thread stores the half's tiles
        named_bar_sync(1 + half, 256);
        if (q == 0 && (cc & 1) == 0 && lane == 0) {
            const uint32_t st = base + half * FWD_STAGE;
            tma_store_4d(&p.tmH, st, (2 * half) * 32, t0, b, p.l + 1);
            tma_store_4d(&p.tmH, st + TILE16, (2 * half + 1) * 32, t0, b, p.l + 1);
            tma_store_4d(&p.tmA_hi, s_rt + half * 2 * TILE16, half * 64, t0, b, p.a_out);
            tma_store_4d(&p.tmA_lo, s_rt + half * 2 * TILE16 + TILE16, half * 64, t0, b, p.a_out);
            if (p.f_slot >= 0) tma_store_4d(&p.tmF, st + 2 * TILE16, half * 64, t0, b, p.f_slot);
            tma_store_commit();
            if (half == 0) TL(29);
            tma_store_wait_read();
            if (half == 0) TL(30);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) TL(31);
    if (warp == 1) tmem_dealloc<256>(tmem);
}

#include "stack_fwd.cuh"
#include "stack_bwd.cuh"
#include "stack_bwd_split.cuh"
#ifdef AST_EXPERIMENTS
void tc_set_tl_layer(int l) { cudaMemcpyToSymbol(g_tl_layer, &l, sizeof(l)); }
#endif

// =============================================================================================
// backward boundary l (between layer l-1 and layer l), plain bf16 operands
// =============================================================================================
struct alignas(64) BwdArgs {
    CUtensorMap tmDz;                // bf16 dz ping-pong [2][B][T][128], box {64,128,1,1} (loads and stores)
    CUtensorMap tmG;                 // fp32 running gradient [1][B][T][128], box {32,128,1,1} (loads and stores)
    CUtensorMap tmDF;                // fp32 head gradients [n_head][B][T][128], box {32,128,1,1}
    CUtensorMap tmW1, tmW2;
    CUtensorMap tmF, tmS;            // fused Gatys style gradient: bf16 extract F [L][B][T][128] and bf16 Ssym rows [B*L*128][128]
    int sg_slot, sg_L;               // style slot whose gradient F.Ssym is added at this boundary (-1: none; excludes df_idx >= 0)
    int w1_row, w2_row;
    int T, shift;                    // tap k reads dz row t + (k-1)*shift, shift = -d
    int do_gemm1, do_gemm2, g_init, df_idx;      // df_idx < 0: no head gradient at this boundary
    int dz_in, dz_out;
    const uint32_t* hmask_in; const uint32_t* zmask_in;
};
// smem: slots 0,1,2 x (A 16 KB | W 16 KB) | G 64 KB (fp32, updated in place) | DF 64 KB (fp32 head gradient, OR the
// operands F (2 K-halves) and Ssym (2 K-halves) of the fused style-gradient GEMM; then dz staging)
// Slot 2 is a pipeline stage during GEMM1, then RT = bf16(g') (A operand of the 1x1 GEMM, 2 K-halves).
#define BWD_STAGE 32768
#define BWD_G (3 * BWD_STAGE)
#define BWD_DF (BWD_G + 65536)
#define BWD_SMEM (BWD_DF + 65536 + 1024 + 128)
#define BWD_THREADS 576
#define BB_IO 88u

__global__ void __launch_bounds__(BWD_THREADS, 1) tc_bwd_kernel(const __grid_constant__ BwdArgs p) {
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t base = (smem_u32(tc_smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = tc_smem_raw + (base - smem_u32(tc_smem_raw));
    const uint32_t s_rt = base + 2 * BWD_STAGE, s_g = base + BWD_G, s_df = base + BWD_DF, bar = s_df + 65536;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row0 = blockIdx.x * 128;
    const int b = row0 / p.T, t0 = row0 - b * p.T;
    const bool has_io = !p.g_init || p.df_idx >= 0 || p.sg_slot >= 0;
    const int n1 = p.do_gemm1 ? 2 * ((int)tap_live(t0, 0, p.shift, p.T) + 1 + (int)tap_live(t0, 2, p.shift, p.T)) : 0;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.tmDz); tma_prefetch_desc(&p.tmW1); tma_prefetch_desc(&p.tmW2); tma_prefetch_desc(&p.tmG);
        tma_prefetch_desc(&p.tmDF);
        if (p.sg_slot >= 0) { tma_prefetch_desc(&p.tmF); tma_prefetch_desc(&p.tmS); }
        for (int i = 0; i < 3; ++i) { mbar_init(bar + FB_FULL(i), 1); mbar_init(bar + FB_EMPTY(i), 1); }
        mbar_init(bar + FB_ACC1, 1); mbar_init(bar + FB_RT(0), 8); mbar_init(bar + FB_RT(1), 8); mbar_init(bar + FB_ACC2, 1);
        mbar_init(bar + BB_IO, 1);
        mbar_init_fence();
    }
    if (warp == 1) tmem_alloc<512>(bar + FB_TMEM);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gen + (bar + FB_TMEM - base));
    const uint32_t acc1 = tmem, acc2 = tmem + 128, acc3 = tmem + 256;
    const bool sg = p.sg_slot >= 0;
    pdl_launch_dependents();
    if (threadIdx.x != 0) pdl_wait();      // the producer thread first requests constant weight tiles (see forward kernel)

    if (warp == 0) {
        // ================================ TMA producer ========================================
        if (lane == 0) {
            int fills[3] = {0, 0, 0};
            int tap_of[6];
            {
                int n = 0;
                for (int k = 0; k < 3; ++k)
                    if (p.do_gemm1 && tap_live(t0, k, p.shift, p.T)) { tap_of[n++] = k; tap_of[n++] = k; }
            }
            auto load_w = [&](int it) {
                const int slot = it % 3;
                const uint32_t fb = bar + FB_FULL(slot);
                mbar_expect_tx(fb, 2 * TILE16);
                tma_load_2d(base + slot * BWD_STAGE + TILE16, &p.tmW1, fb, (it & 1) * 64, p.w1_row + tap_of[it] * 128);
            };
            const int npre = n1 < 3 ? n1 : 3;
            for (int it = 0; it < npre; ++it) load_w(it);
            pdl_wait();
            bool io_sent = !has_io;
            auto send_io = [&]() {
                if (io_sent) return;
                io_sent = true;
                mbar_expect_tx(bar + BB_IO, (uint32_t)((p.g_init ? 0 : 4) + (p.df_idx >= 0 ? 4 : 0) + (sg ? 4 : 0)) * TILE16);
                if (sg)
                    for (int j = 0; j < 2; ++j) {
                        tma_load_4d(s_df + j * TILE16, &p.tmF, bar + BB_IO, j * 64, t0, b, p.sg_slot);
                        tma_load_2d(s_df + (2 + j) * TILE16, &p.tmS, bar + BB_IO, j * 64, (b * p.sg_L + p.sg_slot) * 128);
                    }
                if (!p.g_init)
                    for (int cg = 0; cg < 4; ++cg) tma_load_4d(s_g + cg * TILE16, &p.tmG, bar + BB_IO, cg * 32, t0, b, 0);
                if (p.df_idx >= 0)
                    for (int cg = 0; cg < 4; ++cg) tma_load_4d(s_df + cg * TILE16, &p.tmDF, bar + BB_IO, cg * 32, t0, b, p.df_idx);
            };
            for (int it = 0; it < n1; ++it) {
                const int slot = it % 3;
                if (it == 3) send_io();                  // after the first three operand stages are in flight
                if (it >= npre) {
                    mbar_wait(bar + FB_EMPTY(slot), (fills[slot] & 1) ^ 1);
                    load_w(it);
                }
                ++fills[slot];
                tma_load_4d(base + slot * BWD_STAGE, &p.tmDz, bar + FB_FULL(slot), (it & 1) * 64,
                            t0 + (tap_of[it] - 1) * p.shift, b, p.dz_in);
            }
            send_io();
            if (p.do_gemm2) {                            // Wr^T K-halves into the W part of slots 0 and 1
                for (int j = 0; j < 2; ++j) {
                    mbar_wait(bar + FB_EMPTY(j), (fills[j] & 1) ^ 1);
                    ++fills[j];
                    const uint32_t fb = bar + FB_FULL(j);
                    mbar_expect_tx(fb, TILE16);
                    tma_load_2d(base + j * BWD_STAGE + TILE16, &p.tmW2, fb, j * 64, p.w2_row);
                }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ==========================================
        if (lane == 0) {
            int uses[3] = {0, 0, 0};
            if (p.do_gemm1) {
                uint32_t accum = 0;
                for (int it = 0; it < n1; ++it) {
                    const int slot = it % 3;
                    mbar_wait(bar + FB_FULL(slot), uses[slot] & 1);
                    ++uses[slot];
                    tc_fence_after();
                    const uint32_t st = base + slot * BWD_STAGE;
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) {
                        umma_f16(acc1, umma_desc_kmajor_sw128(st + kk * 32), umma_desc_kmajor_sw128(st + TILE16 + kk * 32), IDESC, accum);
                        accum = 1u;
                    }
                    umma_commit(bar + FB_EMPTY(slot));
                }
            }
            if (sg) {                                    // acc3 = F_tile . Ssym  (the Gatys style gradient of this extract)
                mbar_wait(bar + BB_IO, 0);
                tc_fence_after();
#pragma unroll
                for (int j = 0; j < 2; ++j)
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk)
                        umma_f16(acc3, umma_desc_kmajor_sw128(s_df + j * TILE16 + kk * 32),
                                 umma_desc_kmajor_sw128(s_df + (2 + j) * TILE16 + kk * 32), IDESC, (j > 0 || kk > 0) ? 1u : 0u);
            }
            if (p.do_gemm1 || sg) umma_commit(bar + FB_ACC1);
            if (p.do_gemm2) {
                for (int j = 0; j < 2; ++j) {
                    mbar_wait(bar + FB_RT(j), 0);
                    mbar_wait(bar + FB_FULL(j), uses[j] & 1);
                    tc_fence_after();
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk)
                        umma_f16(acc2, umma_desc_kmajor_sw128(s_rt + j * TILE16 + kk * 32),
                                 umma_desc_kmajor_sw128(base + j * BWD_STAGE + TILE16 + kk * 32), IDESC, (j > 0 || kk > 0) ? 1u : 0u);
                }
                umma_commit(bar + FB_ACC2);
            }
        }
        __syncwarp();
    } else {
        // ================================ epilogue warps ======================================
        const int q = warp & 3, cc = (warp - 2) >> 2, half = cc >> 1;
        const int r = q * 32 + lane;
        const size_t row = (size_t)row0 + r;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        float v[32];

        // ---- epilogue 1: g' = g + acc1 * [h>0] + dF, in place in smem -> TMA store; bf16(g') -> RT ----
        uint32_t mask = 0;
        if (p.do_gemm1) mask = p.hmask_in[row * AST_MW + cc];
        if (p.do_gemm1 || sg) {
            mbar_wait(bar + FB_ACC1, 0);
            tc_fence_after();
        }
        if (p.do_gemm1) {
            tmem_ld32(acc1 + lane_off + cc * 32, v);
#pragma unroll
            for (int i = 0; i < 32; ++i) if (!((mask >> i) & 1u)) v[i] = 0.f;
        } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = 0.f;
        }
        if (sg) {
            float w[32];
            tmem_ld32(acc3 + lane_off + cc * 32, w);
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] += w[i];
        }
        if (has_io) mbar_wait(bar + BB_IO, 0);
        {
            uint8_t* gcg = gen + BWD_G + cc * TILE16;
            const uint8_t* dcg = gen + BWD_DF + cc * TILE16;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint32_t off = sw128_off(r, i);
                if (!p.g_init) {
                    const float4 t = *reinterpret_cast<const float4*>(gcg + off);
                    v[4 * i] += t.x; v[4 * i + 1] += t.y; v[4 * i + 2] += t.z; v[4 * i + 3] += t.w;
                }
                if (p.df_idx >= 0) {
                    const float4 t = *reinterpret_cast<const float4*>(dcg + off);
                    v[4 * i] += t.x; v[4 * i + 1] += t.y; v[4 * i + 2] += t.z; v[4 * i + 3] += t.w;
                }
                *reinterpret_cast<float4*>(gcg + off) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
            }
            if (p.do_gemm2) stage_bf16(gen + 2 * BWD_STAGE + half * TILE16, r, (cc & 1) * 4, v);
        }
        fence_proxy_async_smem();
        if (p.do_gemm2) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar + FB_RT(half));
        }
        named_bar_sync(1 + cc, 128);                       // the 4 quadrant warps of this column group
        if (q == 0 && lane == 0) {
            tma_store_4d(&p.tmG, s_g + cc * TILE16, cc * 32, t0, b, 0);
            tma_store_commit();
        }
        if (p.do_gemm2) {
            // ---- epilogue 2: dz = acc2 * [z>0] -> bf16, staged in the (now free) DF region -------------
            mask = p.zmask_in[row * AST_MW + cc];
            mbar_wait(bar + FB_ACC2, 0);
            tc_fence_after();
            tmem_ld32(acc2 + lane_off + cc * 32, v);
#pragma unroll
            for (int i = 0; i < 32; ++i) if (!((mask >> i) & 1u)) v[i] = 0.f;
            // every warp of this half must have finished READING the DF tile before it is overwritten:
            // acc2 is complete only after all 16 warps arrived on RT, i.e. after their epilogue-1 reads
            stage_bf16(gen + BWD_DF + half * TILE16, r, (cc & 1) * 4, v);
            fence_proxy_async_smem();
            named_bar_sync(5 + half, 256);
            if (q == 0 && (cc & 1) == 0 && lane == 0) {
                tma_store_4d(&p.tmDz, s_df + half * TILE16, half * 64, t0, b, p.dz_out);
                tma_store_commit();
            }
        }
        if (q == 0 && lane == 0) tma_store_wait_read();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<512>(tmem);
}

// =============================================================================================
// operand preparation
// =============================================================================================
// hi/lo bf16 split of relu(H0): the A operands of layer 0 (later layers get theirs from the epilogue)
__global__ void tc_split_relu_kernel(const float* __restrict__ H, __nv_bfloat16* __restrict__ hi,
                                     __nv_bfloat16* __restrict__ lo, size_t n4) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    float4 v = reinterpret_cast<const float4*>(H)[i];
    v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
    const float hx = bf16_round(v.x), hy = bf16_round(v.y), hz = bf16_round(v.z), hw = bf16_round(v.w);
    reinterpret_cast<uint2*>(hi)[i] = make_uint2(pack_bf16(hx, hy), pack_bf16(hz, hw));
    reinterpret_cast<uint2*>(lo)[i] = make_uint2(pack_bf16(v.x - hx, v.y - hy), pack_bf16(v.z - hz, v.w - hw));
}
// Weight operand rows (each row = 128 K-contiguous bf16, "K-major B operand", row index = n):
//   fwd:  out[(m*2 + part)*128 + n][k] = split_part(W[m][k][n])      (m = layer*3+tap or layer; transposed)
//   bwd:  out[m*128 + n][k]            = bf16(W[m][n][k])            (natural order)
__global__ void tc_weights_fwd_kernel(const float* __restrict__ W, __nv_bfloat16* __restrict__ out, int nmat) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)nmat * AST_C * AST_C) return;
    const int m = (int)(i / (AST_C * AST_C)), n = (int)((i / AST_C) % AST_C), k = (int)(i % AST_C);
    const float w = W[((size_t)m * AST_C + k) * AST_C + n];
    const float h = bf16_round(w);
    out[(((size_t)m * 2 + 0) * AST_C + n) * AST_C + k] = __float2bfloat16_rn(h);
    out[(((size_t)m * 2 + 1) * AST_C + n) * AST_C + k] = __float2bfloat16_rn(w - h);
}
__global__ void tc_weights_bwd_kernel(const float* __restrict__ W, __nv_bfloat16* __restrict__ out, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = __float2bfloat16_rn(W[i]);
}
// lo part of the 2-term split of the backward operands (accurate tier): bf16(W - float(bf16(W)))
__global__ void tc_weights_bwd_lo_kernel(const float* __restrict__ W, __nv_bfloat16* __restrict__ out, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = __float2bfloat16_rn(W[i] - __bfloat162float(__float2bfloat16_rn(W[i])));
}

// =============================================================================================
// host side
// =============================================================================================
struct TcState {
    int T, B, R, NL;
    __nv_bfloat16 *a_hi, *a_lo, *dz;           // [2][R][128] each (ping-pong)
    __nv_bfloat16 *w1f, *w2f, *w1b, *w2b;
    // accurate tier (AST_PREC_TC3): lo parts of the backward operands, the head gradients' base pointer
    bool split;
    __nv_bfloat16 *dz_lo, *w1b_lo, *w2b_lo;
    CUtensorMap tm_dz_lo, tm_w1b_lo, tm_w2b_lo;
    const float* dF;
    __nv_bfloat16 *F, *ssym;                   // Gatys head: bf16 extracts [n_fslots][R][128], bf16 Ssym [B*n_fslots][128][128]
    __nv_bfloat16 *F_lo, *ssym_lo;             // accurate tier: their lo parts
    CUtensorMap tm_F_lo, tm_ssym_lo;
    int n_fslots;
    __half* Ht; int n_tslots;                  // transposed fp16 extract copies [R/32][n_tslots][128][32] of the channel-wise head, or NULL (owned by the caller)
    CUtensorMap tm_F, tm_ssym, tm_gpart, tm_dF_st;
    CUtensorMap tm_a_hi, tm_a_lo, tm_dz, tm_H, tm_g, tm_dF;      // box = 128 rows x 128 bytes, loads and stores
    CUtensorMap tm_w1f, tm_w2f, tm_w1b, tm_w2b;
    uint4 *ll_f, *ll_b;            // near hand-off lines of the whole-stack kernels (stack_fwd.cuh / stack_bwd.cuh)
    int cur;                       // dz buffer holding the most recent backward output
    unsigned* flags;               // [2][R/128][2] publish counters (per tile and K-half) of the forward / backward stack
                                   // kernels, then err[2], then {epoch, CTAs done} of the forward and of the backward kernel
    int n_sm, coop;                // SM count; cooperative launch supported
    int nseg, seg_lo[4], seg_hi[4], seg_halo[4];   // layer segments of the whole-stack kernels and their halo (tiles per side)
    // peer-mapped halos of a time shard (tc_set_peer): side 0 = the rank before this window, 1 = after
    struct {
        bool on;
        const float* H[2]; const float* g[2]; const float* g_alt[2]; unsigned* sig[2];
        int R[2]; int row_off[2];
        int own_lo, own_hi;            // owned range in window coordinates
        unsigned* sig_local;           // this rank's signal words [2 sides][2 directions]
        unsigned* d_step;              // device [2]: forward / backward evaluations started (all ranks count alike); bumped on the stream
    } peer;
};

static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;

// [nbuf][B][T][128] activation tensor; box = {cols, rows, 1, 1}; 128B swizzle (cols * elem = 128 B)
static bool make_act_map(CUtensorMap* m, void* ptr, bool fp32, int T, int B, int nbuf, int box_rows) {
    const uint64_t es = fp32 ? 4 : 2;
    const uint64_t dims[4] = {AST_C, (uint64_t)T, (uint64_t)B, (uint64_t)nbuf};
    const uint64_t str[3] = {AST_C * es, (uint64_t)T * AST_C * es, (uint64_t)B * T * AST_C * es};
    const uint32_t box[4] = {fp32 ? 32u : 64u, (uint32_t)box_rows, 1, 1};
    const uint32_t estr[4] = {1, 1, 1, 1};
    return g_encode(m, fp32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, ptr, dims, str, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
static bool make_w_map(CUtensorMap* m, void* ptr, size_t rows) {             // [rows][128] bf16, box [128][64]
    const uint64_t dims[2] = {AST_C, (uint64_t)rows};
    const uint64_t str[1] = {AST_C * 2};
    const uint32_t box[2] = {64, 128};
    const uint32_t estr[2] = {1, 1};
    return g_encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ptr, dims, str, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static bool make_rows_map(CUtensorMap* m, void* ptr, bool fp32, size_t rows, int box_rows) {   // [rows][128]
    const uint64_t es = fp32 ? 4 : 2;
    const uint64_t dims[2] = {AST_C, (uint64_t)rows};
    const uint64_t str[1] = {AST_C * es};
    const uint32_t box[2] = {fp32 ? 32u : 64u, (uint32_t)box_rows};
    const uint32_t estr[2] = {1, 1};
    return g_encode(m, fp32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ptr, dims, str, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

TcState* tc_create(int T, int B, int NL, float* H, float* g, float* dF, int n_head, int n_fslots,
                   float* gram_part, int64_t gram_part_mats, bool split, int64_t* bytes, const char** why) {
    if (!g_encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPointByVersion("cuTensorMapEncodeTiled", &fn, 12000, cudaEnableDefault, &qres) != cudaSuccess ||
            qres != cudaDriverEntryPointSuccess || !fn) {
            *why = "cuTensorMapEncodeTiled not available from the driver";
            return nullptr;
        }
        g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
    }
    TcState* s = new TcState();
    memset(s, 0, sizeof(*s));
    s->T = T; s->B = B; s->R = B * T; s->NL = NL; s->split = split; s->dF = dF;
    const size_t act = (size_t)2 * s->R * AST_C * sizeof(__nv_bfloat16);
    const size_t mat = (size_t)AST_C * AST_C * sizeof(__nv_bfloat16);
    bool ok = cudaMalloc(&s->a_hi, act) == cudaSuccess && cudaMalloc(&s->a_lo, act) == cudaSuccess &&
              cudaMalloc(&s->dz, act) == cudaSuccess && cudaMalloc(&s->w1f, mat * NL * 3 * 2) == cudaSuccess &&
              cudaMalloc(&s->w2f, mat * NL * 2) == cudaSuccess && cudaMalloc(&s->w1b, mat * NL * 3) == cudaSuccess &&
              cudaMalloc(&s->w2b, mat * NL) == cudaSuccess;
    ok = ok && cudaMalloc(&s->flags, ((size_t)4 * (s->R / 128) + 6) * sizeof(unsigned)) == cudaSuccess;
    if (!ok) { *why = "cudaMalloc failed"; tc_destroy(s); return nullptr; }
    cudaMemset(s->flags, 0, ((size_t)4 * (s->R / 128) + 6) * sizeof(unsigned));
    {   // LL lines: [2 ping-pong][tiles][2 sides][2 K-halves][4 cc]([2 parts])[4][32 rows] x 16 B; tag 0 never validates
        const size_t nt_ = (size_t)(s->R / 128);
        const size_t lf = nt_ * 2 * 2 * 2 * LL_ROWS * 8 * 4 * sizeof(uint4), lb = split ? lf : lf / 2;   // split: hi and lo lines
        if (cudaMalloc(&s->ll_f, lf) != cudaSuccess || cudaMalloc(&s->ll_b, lb) != cudaSuccess) { *why = "cudaMalloc failed"; tc_destroy(s); return nullptr; }
        cudaMemset(s->ll_f, 0, lf); cudaMemset(s->ll_b, 0, lb);
        *bytes += (int64_t)(lf + lb);
    }
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&s->n_sm, cudaDevAttrMultiProcessorCount, dev);
        cudaDeviceGetAttribute(&s->coop, cudaDevAttrCooperativeLaunch, dev);
    }
    *bytes += (int64_t)(3 * act + mat * NL * 12);
    // far taps of a group-edge tile read rows of these buffers that no launch of the group has written: keep them finite
    cudaMemset(s->a_hi, 0, act); cudaMemset(s->a_lo, 0, act); cudaMemset(s->dz, 0, act);
    tc_plan_segments(s, false);
    ok = make_act_map(&s->tm_a_hi, s->a_hi, false, T, B, 2, 128) && make_act_map(&s->tm_a_lo, s->a_lo, false, T, B, 2, 128) &&
         make_act_map(&s->tm_dz, s->dz, false, T, B, 2, 128) && make_act_map(&s->tm_H, H, true, T, B, NL + 1, 128) &&
         make_act_map(&s->tm_g, g, true, T, B, 1, 128) && make_act_map(&s->tm_dF, dF, true, T, B, n_head > 0 ? n_head : 1, 128) &&
         make_w_map(&s->tm_w1f, s->w1f, (size_t)NL * 3 * 2 * AST_C) && make_w_map(&s->tm_w2f, s->w2f, (size_t)NL * 2 * AST_C) &&
         make_w_map(&s->tm_w1b, s->w1b, (size_t)NL * 3 * AST_C) && make_w_map(&s->tm_w2b, s->w2b, (size_t)NL * AST_C);
    if (ok && split) {
        if (cudaMalloc(&s->dz_lo, act) != cudaSuccess || cudaMalloc(&s->w1b_lo, mat * NL * 3) != cudaSuccess ||
            cudaMalloc(&s->w2b_lo, mat * NL) != cudaSuccess) { *why = "cudaMalloc failed"; tc_destroy(s); return nullptr; }
        cudaMemset(s->dz_lo, 0, act);
        *bytes += (int64_t)(act + mat * NL * 4);
        ok = make_act_map(&s->tm_dz_lo, s->dz_lo, false, T, B, 2, 128) && make_w_map(&s->tm_w1b_lo, s->w1b_lo, (size_t)NL * 3 * AST_C) &&
             make_w_map(&s->tm_w2b_lo, s->w2b_lo, (size_t)NL * AST_C);
    }
    s->n_fslots = n_fslots;
    if (ok && n_fslots > 0) {
        const size_t fb = (size_t)n_fslots * s->R * AST_C * sizeof(__nv_bfloat16), sb = (size_t)B * n_fslots * mat;
        if (cudaMalloc(&s->F, fb) != cudaSuccess || cudaMalloc(&s->ssym, sb) != cudaSuccess) {
            *why = "cudaMalloc failed"; tc_destroy(s); return nullptr;
        }
        *bytes += (int64_t)(fb + sb);
        ok = make_act_map(&s->tm_F, s->F, false, T, B, n_fslots, 128) &&
             make_rows_map(&s->tm_ssym, s->ssym, false, (size_t)B * n_fslots * AST_C, 128) &&
             make_rows_map(&s->tm_gpart, gram_part, true, (size_t)gram_part_mats * AST_C, 32) &&
             make_act_map(&s->tm_dF_st, dF, true, T, B, n_head > 0 ? n_head : 1, 32) && setup_gram_tc_kernels() == 0;
        if (ok && split) {                       // the accurate tier's heads multiply hi / lo pairs of the extracts and of Ssym
            if (cudaMalloc(&s->F_lo, fb) != cudaSuccess || cudaMalloc(&s->ssym_lo, sb) != cudaSuccess) {
                *why = "cudaMalloc failed"; tc_destroy(s); return nullptr;
            }
            *bytes += (int64_t)(fb + sb);
            ok = make_act_map(&s->tm_F_lo, s->F_lo, false, T, B, n_fslots, 128) &&
                 make_rows_map(&s->tm_ssym_lo, s->ssym_lo, false, (size_t)B * n_fslots * AST_C, 128);
        }
    }
    if (!ok) { *why = "cuTensorMapEncodeTiled failed"; tc_destroy(s); return nullptr; }
    if (cudaFuncSetAttribute(tc_fwd_stack_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, F2_SMEM) != cudaSuccess ||
        cudaFuncSetAttribute(tc_bwd_stack_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, B2_SMEM) != cudaSuccess ||
        cudaFuncSetAttribute(tc_bwd_stack_split_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, S2_SMEM) != cudaSuccess ||
        cudaFuncSetAttribute(tc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM) != cudaSuccess ||
        cudaFuncSetAttribute(tc_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM) != cudaSuccess) {
        *why = "cudaFuncSetAttribute(max dynamic smem) failed";
        tc_destroy(s);
        return nullptr;
    }
    return s;
}

void tc_destroy(TcState* s) {
    if (!s) return;
    cudaFree(s->a_hi); cudaFree(s->a_lo); cudaFree(s->dz);
    cudaFree(s->w1f); cudaFree(s->w2f); cudaFree(s->w1b); cudaFree(s->w2b);
    cudaFree(s->F); cudaFree(s->ssym); cudaFree(s->F_lo); cudaFree(s->ssym_lo); cudaFree(s->flags); cudaFree(s->ll_f); cudaFree(s->ll_b);
    cudaFree(s->dz_lo); cudaFree(s->w1b_lo); cudaFree(s->w2b_lo);
    delete s;
}

int tc_set_weights(TcState* s, const float* Wd, const float* Wr, cudaStream_t st) {
    const size_t nd = (size_t)s->NL * 3 * AST_C * AST_C, nr = (size_t)s->NL * AST_C * AST_C;
    tc_weights_fwd_kernel<<<(unsigned)((nd + 255) / 256), 256, 0, st>>>(Wd, s->w1f, s->NL * 3);
    tc_weights_fwd_kernel<<<(unsigned)((nr + 255) / 256), 256, 0, st>>>(Wr, s->w2f, s->NL);
    tc_weights_bwd_kernel<<<(unsigned)((nd + 255) / 256), 256, 0, st>>>(Wd, s->w1b, nd);
    tc_weights_bwd_kernel<<<(unsigned)((nr + 255) / 256), 256, 0, st>>>(Wr, s->w2b, nr);
    if (s->split) {
        tc_weights_bwd_lo_kernel<<<(unsigned)((nd + 255) / 256), 256, 0, st>>>(Wd, s->w1b_lo, nd);
        tc_weights_bwd_lo_kernel<<<(unsigned)((nr + 255) / 256), 256, 0, st>>>(Wr, s->w2b_lo, nr);
    }
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

void tc_prepare_input(TcState* s, const float* H0, cudaStream_t st) {
    const size_t n4 = (size_t)s->R * AST_C / 4;
    tc_split_relu_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, st>>>(H0, s->a_hi, s->a_lo, n4);
    ast_launch_counter += 1;
}

void tc_layer_fwd(TcState* s, int l, const float* bd, const float* br, uint32_t* zmask_out, uint32_t* hmask_out,
                  int f_slot, cudaStream_t st) {
    FwdArgs a;
    memset(&a, 0, sizeof(a));
    a.tmA_hi = s->tm_a_hi; a.tmA_lo = s->tm_a_lo; a.tmH = s->tm_H; a.tmW1 = s->tm_w1f; a.tmW2 = s->tm_w2f;
    a.w1_row = l * 3 * 2 * AST_C; a.w2_row = l * 2 * AST_C;
    a.T = s->T; a.shift = 1 << (l % 10); a.l = l; a.a_in = l & 1; a.a_out = (l & 1) ^ 1;
    a.bias1 = bd; a.bias2 = br; a.zmask_out = zmask_out; a.hmask_out = hmask_out;
    a.f_slot = (s->n_fslots > 0) ? f_slot : -1;
    if (a.f_slot >= 0) a.tmF = s->tm_F;
    launch_pdl(tc_fwd_kernel, dim3(s->R / 128), dim3(FWD_THREADS), FWD_SMEM, st, a);
    ast_launch_counter += 1;
}

// ---- whole-stack forward (stack_fwd.cuh) ------------------------------------------------------------
template <typename Kern, typename Args>
static cudaError_t launch_coop(Kern kernel, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const Args& args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeCooperative;           // all CTAs co-resident, or the launch fails
    at[0].val.cooperative = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, args);
}
// ---- segments and groups (stack_fwd.cuh) -----------------------------------------------------------------------
// One segment = all layers when every pair fits one tile per SM and nobody exchanges halos; else one segment per
// 10-layer stack (dilations 1..512, sum 1023: the halo a group needs on each inner side is 8 tiles; 1 tile for a last
// segment of <= 6 layers).
void tc_plan_segments(TcState* s, bool per_stack) {
    const bool split = per_stack || s->T / 128 > TC_GROUP_CAP(s);
    s->nseg = 0;
    for (int lo = 0; lo < s->NL; lo = s->seg_hi[s->nseg - 1]) {
        const int hi = split ? (lo + 10 < s->NL ? lo + 10 : s->NL) : s->NL;
        int sum = 0;
        for (int l = lo; l < hi; ++l) sum += stack_dil(l);
        s->seg_lo[s->nseg] = lo; s->seg_hi[s->nseg] = hi; s->seg_halo[s->nseg] = (sum + 127) / 128;
        ++s->nseg;
    }
}
int tc_segments(const TcState* s) { return s->nseg; }
int tc_segment_end(const TcState* s, int k) { return s->seg_hi[k]; }
bool tc_stack_ok(const TcState* s) { return s && s->coop && s->NL <= 32 && s->n_sm >= 64; }
void tc_set_transposed_extracts(TcState* s, void* Ht, int n_slots) { s->Ht = static_cast<__half*>(Ht); s->n_tslots = n_slots; }
unsigned* tc_stack_err(TcState* s) { return s->flags + (size_t)4 * (s->R / 128); }

void tc_set_peer_geometry(TcState* s, int own_lo, int own_hi, unsigned* sig_local) {
    s->peer.own_lo = own_lo; s->peer.own_hi = own_hi; s->peer.sig_local = sig_local;
    s->peer.d_step = sig_local + 4;                  // the caller's buffer: 4 signal words, then the two evaluation counters
}
void tc_clear_peers(TcState* s) {
    s->peer.on = false;
    for (int sd = 0; sd < 2; ++sd) { s->peer.H[sd] = s->peer.g[sd] = s->peer.g_alt[sd] = nullptr; s->peer.sig[sd] = nullptr; }
}
void tc_set_peer(TcState* s, int side, const float* H, const float* g, const float* g_alt, unsigned* sig, int peer_R, int row_off) {
    s->peer.on = true;
    s->peer.H[side] = H; s->peer.g[side] = g; s->peer.g_alt[side] = g_alt; s->peer.sig[side] = sig;
    s->peer.R[side] = peer_R; s->peer.row_off[side] = row_off;
}
// The StackPeer block of segment k.  order = position of the segment in its sweep (forward: k, backward: nseg-1-k); a
// segment waits for the signal the neighbours' previous segment of the SAME evaluation sent, and (unless it is the last of
// the sweep) signals them in turn.  src_of(side) = the neighbour's buffer this segment resumes from.
__global__ void tc_peer_step_bump_kernel(unsigned* step) { *step += 1u; }
template <typename SrcOf>
static StackPeer tc_peer_block(TcState* s, int order, int dir, int halo_in, int halo_out, SrcOf src_of, cudaStream_t st) {
    StackPeer pr;
    memset(&pr, 0, sizeof(pr));
    if (!s->peer.on) return pr;
    // the evaluation number lives in device memory and is counted on the stream: eager launches and replays of a captured
    // graph (HaloShardedStep.capture) number their evaluations alike on every rank
    if (order == 0) { tc_peer_step_bump_kernel<<<1, 1, 0, st>>>(s->peer.d_step + dir); ast_launch_counter += 1; }
    pr.step = s->peer.d_step + dir;
    const int lo = s->peer.own_lo, hi = s->peer.own_hi;
    if (order > 0) {
        pr.wait_seq = (unsigned)order;
        for (int sd = 0; sd < 2; ++sd) {
            if (!s->peer.sig[sd]) continue;
            pr.src[sd] = src_of(sd);
            pr.row_off[sd] = s->peer.row_off[sd];
            pr.sig_wait[sd] = s->peer.sig_local + sd * 2 + dir;
        }
        pr.t0[0] = lo - halo_out; pr.t1[0] = lo - halo_in;          // rows before the owned range ...
        pr.t0[1] = hi + halo_in;  pr.t1[1] = hi + halo_out;         // ... and after it
    }
    if (order + 1 < s->nseg) {
        pr.signal_seq = (unsigned)order + 1u;
        for (int sd = 0; sd < 2; ++sd) pr.sig_out[sd] = s->peer.sig[sd] ? s->peer.sig[sd] + (1 - sd) * 2 + dir : nullptr;
    }
    return pr;
}
#define TC_HALO_ROWS 1024

struct TcLaunch { int tile0, ntiles, grp_t0, grp_t1, wr_t0, wr_t1; };
// The launches of one segment: whole pairs while a pair fits the SMs, else groups of <= n_sm consecutive tiles of one pair
// with `halo` tiles per inner side that are computed but not written.
// [r0, r1): the tiles of a pair that have to be evaluated at all (the backward of a time shard that exchanges halos needs
// only own +- 1024 rows of its window); outside the whole-pair case the range's ends behave like clip ends.
static int tc_plan_launches(const TcState* s, int halo, TcLaunch* out, int max_out, int r0 = 0, int r1 = -1) {
    const int tpp = s->T / 128, cap = TC_GROUP_CAP(s);
    if (r1 < 0) r1 = tpp;
    int n = 0;
    if (tpp <= cap && r0 == 0 && r1 == tpp) {
        const int ppl = cap / tpp;
        for (int p0 = 0; p0 < s->B && n < max_out; p0 += ppl) {
            const int np = s->B - p0 < ppl ? s->B - p0 : ppl;
            out[n++] = TcLaunch{p0 * tpp, np * tpp, 0, s->T, 0, s->T};
        }
        return n;
    }
    for (int p = 0; p < s->B; ++p)
        for (int a = r0; a < r1 && n < max_out;) {
            const int lo = a - halo > r0 ? a - halo : r0;
            const int hi = lo + cap < r1 ? lo + cap : r1;
            const int e = hi == r1 ? r1 : hi - halo;
            out[n++] = TcLaunch{p * tpp + lo, hi - lo, lo * 128, hi * 128, a * 128, e * 128};
            a = e;
        }
    return n;
}
#define TC_MAX_LAUNCHES 256

int tc_forward_stack(TcState* s, const float* x, const float* W0, const float* b0, const float* bd, const float* br,
                     float* H, uint32_t* zmask, uint32_t* hmask, const int* f_slot, const unsigned char* export_h,
                     int seg, cudaStream_t st) {
    FwdStackArgs a;
    memset(&a, 0, sizeof(a));
    a.tmA_hi = s->tm_a_hi; a.tmA_lo = s->tm_a_lo; a.tmW1 = s->tm_w1f; a.tmW2 = s->tm_w2f;
    a.x = x; a.W0 = W0; a.b0 = b0; a.bd = bd; a.br = br; a.H = H; a.zmask = zmask; a.hmask = hmask;
    a.a_hi = s->a_hi; a.a_lo = s->a_lo; a.F = s->F; a.F_lo = s->F_lo;
    a.flags = s->flags; a.err = tc_stack_err(s); a.epoch = tc_stack_err(s) + 2; a.ll = s->ll_f;
    a.T = s->T; a.R = s->R; a.NL = s->NL; a.tiles_per_pair = s->T / 128; a.ntiles = s->R / 128;
    for (int l = 0; l < 32; ++l) a.f_slot[l] = (signed char)((s->n_fslots > 0 && l < s->NL) ? f_slot[l] : -1);
    a.Ht = s->Ht; a.n_tslots = s->n_tslots;
    for (int l = 0; l < 32; ++l) a.t_slot[l] = (signed char)((s->Ht && l < s->NL) ? f_slot[l] : -1);
    for (int i = 0; i <= s->NL; ++i) a.export_h[i] = export_h[i];
    for (int k = 0; k + 1 < s->nseg; ++k) a.export_h[s->seg_hi[k]] = 1;      // the next segment resumes from H[seg_hi]
    TcLaunch ln[TC_MAX_LAUNCHES];
    cudaError_t e = cudaSuccess;
    for (int k = (seg < 0 ? 0 : seg); k < (seg < 0 ? s->nseg : seg + 1) && e == cudaSuccess; ++k) {
        a.l_begin = s->seg_lo[k]; a.l_end = s->seg_hi[k];
        {   // publish i (relu(h_i)) is consumed by layer i: its dilation decides how many warps release per K-half
            unsigned acc = 0;
            for (int i = 0; i < 34; ++i) {
                if (i >= a.l_begin && i < a.l_end) acc += (unsigned)stack_npub(stack_dil(i));
                a.need[i] = (unsigned short)acc;
            }
        }
        const int nl = tc_plan_launches(s, s->seg_halo[k], ln, TC_MAX_LAUNCHES);
        const bool disjoint = s->T / 128 <= TC_GROUP_CAP(s);       // whole-pair launches: one memset covers them all
        if (disjoint) cudaMemsetAsync(s->flags, 0, 2 * (size_t)(s->R / 128) * sizeof(unsigned), st);
        // forward: the stream is exact on own +- 1024 after a segment; the 1024 rows beyond come from the neighbours' H[l_begin]
        const StackPeer pr = tc_peer_block(s, k, 0, TC_HALO_ROWS, 2 * TC_HALO_ROWS, [&](int sd) {
            return s->peer.H[sd] + (size_t)a.l_begin * s->peer.R[sd] * AST_C; }, st);
        for (int i = 0; i < nl && e == cudaSuccess; ++i) {
            a.peer = pr;
            if (i + 1 < nl) a.peer.signal_seq = 0;                 // only the segment's last launch signals
            a.tile0 = ln[i].tile0; a.grp_t0 = ln[i].grp_t0; a.grp_t1 = ln[i].grp_t1; a.wr_t0 = ln[i].wr_t0; a.wr_t1 = ln[i].wr_t1;
            if (!disjoint) cudaMemsetAsync(s->flags + (size_t)2 * ln[i].tile0, 0, (size_t)2 * ln[i].ntiles * sizeof(unsigned), st);
            e = launch_coop(tc_fwd_stack_kernel, dim3(ln[i].ntiles), dim3(F2_THREADS), F2_SMEM, st, a);
            ast_launch_counter += 1;
        }
    }
    return e == cudaSuccess ? 0 : -1;
}

int tc_backward_stack(TcState* s, const uint32_t* hmask, const uint32_t* zmask, const int* df_idx, const int* sg_slot,
                      int sg_L, float* g, float* g_alt, const float* W0, float* pdots, int seg, cudaStream_t st) {
    BwdStackArgs a;
    memset(&a, 0, sizeof(a));
    a.tmDz = s->tm_dz; a.tmW1 = s->tm_w1b; a.tmW2 = s->tm_w2b; a.tmDF = s->tm_dF;
    a.tmF = s->n_fslots > 0 ? s->tm_F : s->tm_dz; a.tmS = s->n_fslots > 0 ? s->tm_ssym : s->tm_w2b;   // unused without slots
    a.dz = s->dz; a.hmask = hmask; a.zmask = zmask; a.W0 = W0; a.pdots = pdots;
    const size_t nt = (size_t)(s->R / 128);
    a.flags = s->flags + 2 * nt; a.err = tc_stack_err(s) + 1; a.epoch = tc_stack_err(s) + 4; a.ll = s->ll_b;
    a.T = s->T; a.R = s->R; a.NL = s->NL; a.tiles_per_pair = s->T / 128; a.ntiles = s->R / 128; a.sg_L = sg_L;
    for (int l = 0; l < 33; ++l) { a.df_idx[l] = -1; a.sg_slot[l] = -1; }
    for (int l = 0; l <= s->NL; ++l) {
        a.sg_slot[l] = (signed char)((s->n_fslots > 0 && df_idx[l] < 0) ? sg_slot[l] : -1);
        a.df_idx[l] = (signed char)df_idx[l];
    }
    TcLaunch ln[TC_MAX_LAUNCHES];
    cudaError_t e = cudaSuccess;
    for (int k = (seg < 0 ? s->nseg - 1 : seg); k >= (seg < 0 ? 0 : seg) && e == cudaSuccess; --k) {
        a.l_top = k == s->nseg - 1 ? s->NL : s->seg_hi[k]; a.l_bot = s->seg_lo[k];
        a.g = (k & 1) ? g_alt : g; a.g_in = (k & 1) ? g : g_alt;      // segment k leaves its result in buffer k & 1 (segment 0: g)
        {   // dz_m is consumed by boundary m; this launch publishes m = l_top-1 down to l_bot (l_bot-1 = -1 never exists)
            unsigned acc = 0;
            for (int m = 32; m >= 0; --m) {
                if (m < a.l_top && m >= a.l_bot && m < s->NL) acc += (unsigned)stack_npub(stack_dil(m));
                a.need[m] = (unsigned short)acc;
            }
        }
        // a time shard with halo exchange: only own +- 1024 rows carry exact masks and head gradients, and only the owned rows
        // are kept -- the outer 1024 rows of the window are the forward's business
        int r0 = 0, r1 = -1;
        if (s->peer.sig_local) {
            r0 = (s->peer.own_lo - TC_HALO_ROWS) / 128; r1 = (s->peer.own_hi + TC_HALO_ROWS) / 128;
            if (r0 < 0) r0 = 0;
            if (r1 > s->T / 128) r1 = s->T / 128;
        }
        const int nl = tc_plan_launches(s, s->seg_halo[k], ln, TC_MAX_LAUNCHES, r0, r1);
        const bool disjoint = s->T / 128 <= TC_GROUP_CAP(s) && r0 == 0 && (r1 < 0 || r1 == s->T / 128);   // whole-pair launches
        cudaMemsetAsync(s->flags + 2 * nt, 0, 2 * nt * sizeof(unsigned), st);
        // backward: the running gradient is exact on the owned rows after a segment; the 1024 rows on either side come from the
        // neighbours' result of the previous segment (their buffer of the same parity as this rank's g_in)
        const StackPeer pr = tc_peer_block(s, s->nseg - 1 - k, 1, 0, TC_HALO_ROWS, [&](int sd) {
            return (k & 1) ? s->peer.g[sd] : s->peer.g_alt[sd]; }, st);
        for (int i = 0; i < nl && e == cudaSuccess; ++i) {
            a.peer = pr;
            if (i + 1 < nl) a.peer.signal_seq = 0;
            a.tile0 = ln[i].tile0; a.grp_t0 = ln[i].grp_t0; a.grp_t1 = ln[i].grp_t1; a.wr_t0 = ln[i].wr_t0; a.wr_t1 = ln[i].wr_t1;
            if (i > 0 && !disjoint) cudaMemsetAsync(s->flags + 2 * nt + (size_t)2 * ln[i].tile0, 0, (size_t)2 * ln[i].ntiles * sizeof(unsigned), st);
            if (s->split) {
                BwdSplitArgs q;
                memset(&q, 0, sizeof(q));
                q.b = a; q.tmDzLo = s->tm_dz_lo; q.tmW1Lo = s->tm_w1b_lo; q.tmW2Lo = s->tm_w2b_lo; q.dz_lo = s->dz_lo; q.dF = s->dF;
                e = launch_coop(tc_bwd_stack_split_kernel, dim3(ln[i].ntiles), dim3(F2_THREADS), S2_SMEM, st, q);
            } else
            e = launch_coop(tc_bwd_stack_kernel, dim3(ln[i].ntiles), dim3(F2_THREADS), B2_SMEM, st, a);
            ast_launch_counter += 1;
        }
    }
    return e == cudaSuccess ? 0 : -1;
}

void tc_layer_bwd(TcState* s, int l, const uint32_t* hmask_in, int df_idx, int sg_slot, int sg_L, const uint32_t* zmask_in,
                  cudaStream_t st) {
    BwdArgs a;
    memset(&a, 0, sizeof(a));
    if (l == s->NL) s->cur = 0;                 // top of the stack: (re)start the dz ping-pong
    a.tmDz = s->tm_dz; a.tmG = s->tm_g; a.tmDF = s->tm_dF;
    a.tmW1 = s->tm_w1b; a.tmW2 = s->tm_w2b;
    a.T = s->T; a.df_idx = df_idx; a.hmask_in = hmask_in; a.zmask_in = zmask_in;
    a.sg_slot = (s->n_fslots > 0 && df_idx < 0) ? sg_slot : -1; a.sg_L = sg_L;
    if (a.sg_slot >= 0) { a.tmF = s->tm_F; a.tmS = s->tm_ssym; }
    a.dz_in = s->cur; a.dz_out = s->cur ^ 1;
    if (l < s->NL) { a.do_gemm1 = 1; a.w1_row = l * 3 * AST_C; a.shift = -(1 << (l % 10)); }
    else a.g_init = 1;
    if (l >= 1) { a.do_gemm2 = 1; a.w2_row = (l - 1) * AST_C; }
    launch_pdl(tc_bwd_kernel, dim3(s->R / 128), dim3(BWD_THREADS), BWD_SMEM, st, a);
    ast_launch_counter += 1;
    s->cur ^= 1;
}

// ---- Gatys head ---------------------------------------------------------------------------------
void tc_gram(TcState* s, int L, int nchunk, int chunk_tiles, int tile_lo, int tile_hi, cudaStream_t st) {
    GramTcArgs a;
    memset(&a, 0, sizeof(a));
    a.tmF = s->tm_F; a.tmP = s->tm_gpart;
    if (s->split) a.tmFlo = s->tm_F_lo;
    a.L = L; a.chunk_tiles = chunk_tiles; a.tile_lo = tile_lo; a.tile_hi = tile_hi;
    launch_gram_tc(a, nchunk, L, s->B, s->split, st);
}
__nv_bfloat16* tc_ssym_bf16(TcState* s) { return s ? s->ssym : nullptr; }
__nv_bfloat16* tc_ssym_bf16_lo(TcState* s) { return s ? s->ssym_lo : nullptr; }
void tc_style_grad(TcState* s, int L, const int* head_of_slot, cudaStream_t st) {
    SGradTcArgs a;
    memset(&a, 0, sizeof(a));
    a.tmF = s->tm_F; a.tmS = s->tm_ssym; a.tmDF = s->tm_dF_st;
    if (s->split) { a.tmFlo = s->tm_F_lo; a.tmSlo = s->tm_ssym_lo; }
    a.T = s->T; a.L = L;
    for (int i = 0; i < L && i < 32; ++i) {
        a.head_of_slot[i] = head_of_slot[i];
        if (head_of_slot[i] >= 0) a.slot_of[a.n_active++] = i;
    }
    if (a.n_active > 0) launch_sgrad_tc(a, s->R / 128, s->split, st);
}
