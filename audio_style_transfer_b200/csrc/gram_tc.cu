// gram_tc.cu -- the Gatys head on tensor cores (sm_100a): G_l = F_l^T F_l (methods.py:71,73) as a
// tcgen05 split-K contraction over time, and its gradient dF_l = F_l (dG_l + dG_l^T).
// Operands are the bf16 copies of the extracts that the forward layer kernel emits ([t][c] row-major):
//   Gram:       A = B = the SAME smem tile read MN-major (M = N = channel, K = time)
//   style grad: A = the tile read K-major (M = time, K = channel), B = Ssym_l (symmetric, so its
//               row-major bf16 image is already the K-major B operand)
// Deterministic: every CTA writes its own partial Gram; gram_reduce_kernel sums them in fixed order.
#include "tc.cuh"
#include "ptx.cuh"

#include <cuda.h>
#include <algorithm>

#define TILE16 16384
#define GB_FULL(i)  (8u * (i))
#define GB_EMPTY(i) (32u + 8u * (i))
#define GB_ACC  64u
#define GB_TMEM 72u

// =============================================================================================
// Gram partials: grid (nchunk, L, B); CTA = tiles [tile_lo, tile_hi) of its chunk, accumulated in TMEM
// =============================================================================================
#define GRAM_NS 4
#define GRAM_STAGE (2 * TILE16)
#define GRAM_SMEM (GRAM_NS * GRAM_STAGE + 4 * TILE16 + 1024 + 128)

__global__ void __launch_bounds__(192, 1) gram_tc_kernel(const __grid_constant__ GramTcArgs p) {
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t base = (smem_u32(tc_smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = tc_smem_raw + (base - smem_u32(tc_smem_raw));
    const uint32_t s_out = base + GRAM_NS * GRAM_STAGE, bar = s_out + 4 * TILE16;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int l = blockIdx.y, b = blockIdx.z;
    const int lo = max(p.tile_lo, (int)blockIdx.x * p.chunk_tiles);
    const int hi = min(p.tile_hi, ((int)blockIdx.x + 1) * p.chunk_tiles);
    const int ntile = max(hi - lo, 0);

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.tmF); tma_prefetch_desc(&p.tmP);
        for (int i = 0; i < GRAM_NS; ++i) { mbar_init(bar + GB_FULL(i), 1); mbar_init(bar + GB_EMPTY(i), 1); }
        mbar_init(bar + GB_ACC, 1);
        mbar_init_fence();
    }
    if (warp == 1) tmem_alloc<128>(bar + GB_TMEM);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t acc = *reinterpret_cast<volatile uint32_t*>(gen + (bar + GB_TMEM - base));
    pdl_launch_dependents();
    pdl_wait();

    if (warp == 0) {
        if (lane == 0) {
            for (int it = 0; it < ntile; ++it) {
                const int slot = it % GRAM_NS;
                mbar_wait(bar + GB_EMPTY(slot), ((it / GRAM_NS) & 1) ^ 1);
                const uint32_t st = base + slot * GRAM_STAGE, fb = bar + GB_FULL(slot);
                mbar_expect_tx(fb, 2 * TILE16);
                tma_load_4d(st, &p.tmF, fb, 0, (lo + it) * 128, b, l);
                tma_load_4d(st + TILE16, &p.tmF, fb, 64, (lo + it) * 128, b, l);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(128, 128, true, true);
            for (int it = 0; it < ntile; ++it) {
                const int slot = it % GRAM_NS;
                mbar_wait(bar + GB_FULL(slot), (it / GRAM_NS) & 1);
                tc_fence_after();
                const uint32_t st = base + slot * GRAM_STAGE;
#pragma unroll
                for (int ks = 0; ks < 8; ++ks) {
                    const uint64_t d = umma_desc_mnmajor_sw128(st + ks * 2048, TILE16);
                    umma_f16(acc, d, d, idesc, (it > 0 || ks > 0) ? 1u : 0u);
                }
                umma_commit(bar + GB_EMPTY(slot));
            }
            umma_commit(bar + GB_ACC);
        }
        __syncwarp();
    } else {
        const int q = warp & 3, r = q * 32 + lane;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        float v[32];
        mbar_wait(bar + GB_ACC, 0);
        tc_fence_after();
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
            if (ntile > 0) tmem_ld32(acc + lane_off + cc * 32, v);
            else {
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = 0.f;
            }
            uint8_t* o = gen + (s_out - base) + cc * TILE16;
#pragma unroll
            for (int i = 0; i < 8; ++i)
                *reinterpret_cast<float4*>(o + sw128_off(r, i)) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
            const int prow = (((b * p.L + l) * (int)gridDim.x) + (int)blockIdx.x) * 128 + q * 32;
#pragma unroll
            for (int cc = 0; cc < 4; ++cc) tma_store_2d(&p.tmP, s_out + cc * TILE16 + q * 4096, cc * 32, prow);
            tma_store_commit();
            tma_store_wait_all();
        }
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<128>(acc);
}

// The accurate tier (AST_PREC_TC3): F = F_hi + F_lo as two bf16 copies; G = F_hi^T F_hi + F_hi^T F_lo + F_lo^T F_hi (the lo x lo
// term is below fp32 rounding), three MMAs per k-step into the same TMEM accumulator.  Two stages of {hi, lo} tiles (64 KB).
#define GRAMS_NS 2
#define GRAMS_STAGE (4 * TILE16)
#define GRAMS_SMEM (GRAMS_NS * GRAMS_STAGE + 4 * TILE16 + 1024 + 128)

__global__ void __launch_bounds__(192, 1) gram_tc_split_kernel(const __grid_constant__ GramTcArgs p) {
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t base = (smem_u32(tc_smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = tc_smem_raw + (base - smem_u32(tc_smem_raw));
    const uint32_t s_out = base + GRAMS_NS * GRAMS_STAGE, bar = s_out + 4 * TILE16;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int l = blockIdx.y, b = blockIdx.z;
    const int lo = max(p.tile_lo, (int)blockIdx.x * p.chunk_tiles);
    const int hi = min(p.tile_hi, ((int)blockIdx.x + 1) * p.chunk_tiles);
    const int ntile = max(hi - lo, 0);

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.tmF); tma_prefetch_desc(&p.tmFlo); tma_prefetch_desc(&p.tmP);
        for (int i = 0; i < GRAMS_NS; ++i) { mbar_init(bar + GB_FULL(i), 1); mbar_init(bar + GB_EMPTY(i), 1); }
        mbar_init(bar + GB_ACC, 1);
        mbar_init_fence();
    }
    if (warp == 1) tmem_alloc<128>(bar + GB_TMEM);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t acc = *reinterpret_cast<volatile uint32_t*>(gen + (bar + GB_TMEM - base));

    if (warp == 0) {
        if (lane == 0) {
            for (int it = 0; it < ntile; ++it) {
                const int slot = it % GRAMS_NS;
                mbar_wait(bar + GB_EMPTY(slot), ((it / GRAMS_NS) & 1) ^ 1);
                const uint32_t st = base + slot * GRAMS_STAGE, fb = bar + GB_FULL(slot);
                mbar_expect_tx(fb, 4 * TILE16);
                tma_load_4d(st, &p.tmF, fb, 0, (lo + it) * 128, b, l);
                tma_load_4d(st + TILE16, &p.tmF, fb, 64, (lo + it) * 128, b, l);
                tma_load_4d(st + 2 * TILE16, &p.tmFlo, fb, 0, (lo + it) * 128, b, l);
                tma_load_4d(st + 3 * TILE16, &p.tmFlo, fb, 64, (lo + it) * 128, b, l);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(128, 128, true, true);
            for (int it = 0; it < ntile; ++it) {
                const int slot = it % GRAMS_NS;
                mbar_wait(bar + GB_FULL(slot), (it / GRAMS_NS) & 1);
                tc_fence_after();
                const uint32_t st = base + slot * GRAMS_STAGE;
#pragma unroll
                for (int ks = 0; ks < 8; ++ks) {
                    const uint64_t dh = umma_desc_mnmajor_sw128(st + ks * 2048, TILE16);
                    const uint64_t dl = umma_desc_mnmajor_sw128(st + 2 * TILE16 + ks * 2048, TILE16);
                    umma_f16(acc, dh, dh, idesc, (it > 0 || ks > 0) ? 1u : 0u);
                    umma_f16(acc, dh, dl, idesc, 1u);
                    umma_f16(acc, dl, dh, idesc, 1u);
                }
                umma_commit(bar + GB_EMPTY(slot));
            }
            umma_commit(bar + GB_ACC);
        }
        __syncwarp();
    } else {
        const int q = warp & 3, r = q * 32 + lane;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        float v[32];
        mbar_wait(bar + GB_ACC, 0);
        tc_fence_after();
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
            if (ntile > 0) tmem_ld32(acc + lane_off + cc * 32, v);
            else {
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = 0.f;
            }
            uint8_t* o = gen + (s_out - base) + cc * TILE16;
#pragma unroll
            for (int i = 0; i < 8; ++i)
                *reinterpret_cast<float4*>(o + sw128_off(r, i)) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
            const int prow = (((b * p.L + l) * (int)gridDim.x) + (int)blockIdx.x) * 128 + q * 32;
#pragma unroll
            for (int cc = 0; cc < 4; ++cc) tma_store_2d(&p.tmP, s_out + cc * TILE16 + q * 4096, cc * 32, prow);
            tma_store_commit();
            tma_store_wait_all();
        }
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<128>(acc);
}

void launch_gram_tc(const GramTcArgs& a, int nchunk, int L, int B, bool split, cudaStream_t s) {
    if (split) gram_tc_split_kernel<<<dim3(nchunk, L, B), 192, GRAMS_SMEM, s>>>(a);
    else launch_pdl(gram_tc_kernel, dim3(nchunk, L, B), dim3(192), GRAM_SMEM, s, a);
    ast_launch_counter += 1;
}

// =============================================================================================
// style gradient: dF[head(l)][tile] = F_l[tile] . Ssym_l      grid (R/128, L)
// =============================================================================================
#define SG_SMEM (4 * TILE16 + 1024 + 128)

__global__ void __launch_bounds__(192, 3) sgrad_tc_kernel(const __grid_constant__ SGradTcArgs p) {
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t base = (smem_u32(tc_smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = tc_smem_raw + (base - smem_u32(tc_smem_raw));
    const uint32_t bar = base + 4 * TILE16;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row0 = blockIdx.x * 128, l = p.slot_of[blockIdx.y];      // slots fused into the backward kernel are not launched
    const int b = row0 / p.T, t0 = row0 - b * p.T;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.tmF); tma_prefetch_desc(&p.tmS); tma_prefetch_desc(&p.tmDF);
        mbar_init(bar + GB_FULL(0), 1); mbar_init(bar + GB_ACC, 1);
        mbar_init_fence();
    }
    if (warp == 1) tmem_alloc<128>(bar + GB_TMEM);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t acc = *reinterpret_cast<volatile uint32_t*>(gen + (bar + GB_TMEM - base));
    pdl_launch_dependents();
    pdl_wait();

    if (warp == 0) {
        if (lane == 0) {
            const uint32_t fb = bar + GB_FULL(0);
            mbar_expect_tx(fb, 4 * TILE16);
            tma_load_4d(base, &p.tmF, fb, 0, t0, b, l);
            tma_load_4d(base + TILE16, &p.tmF, fb, 64, t0, b, l);
            tma_load_2d(base + 2 * TILE16, &p.tmS, fb, 0, (b * p.L + l) * 128);
            tma_load_2d(base + 3 * TILE16, &p.tmS, fb, 64, (b * p.L + l) * 128);
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(128, 128);
            mbar_wait(bar + GB_FULL(0), 0);
            tc_fence_after();
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
                    umma_f16(acc, umma_desc_kmajor_sw128(base + j * TILE16 + kk * 32),
                             umma_desc_kmajor_sw128(base + (2 + j) * TILE16 + kk * 32), idesc, (j > 0 || kk > 0) ? 1u : 0u);
            umma_commit(bar + GB_ACC);
        }
        __syncwarp();
    } else {
        const int q = warp & 3, r = q * 32 + lane;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        float v[32];
        mbar_wait(bar + GB_ACC, 0);                 // MMAs complete: the operand tiles are dead, reuse them as staging
        tc_fence_after();
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
            tmem_ld32(acc + lane_off + cc * 32, v);
            uint8_t* o = gen + cc * TILE16;
#pragma unroll
            for (int i = 0; i < 8; ++i)
                *reinterpret_cast<float4*>(o + sw128_off(r, i)) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
#pragma unroll
            for (int cc = 0; cc < 4; ++cc)
                tma_store_4d(&p.tmDF, base + cc * TILE16 + q * 4096, cc * 32, t0 + q * 32, b, p.head_of_slot[l]);
            tma_store_commit();
            tma_store_wait_all();
        }
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<128>(acc);
}

// The accurate tier: dF = F_hi S_hi + F_lo S_hi + F_hi S_lo for every slot.  Persistent: grid (CTAs per slot, slots); a CTA keeps
// its slot's Ssym hi / lo resident (64 KB) and walks its tiles with two {F_hi, F_lo} buffers (64 KB each) and two TMEM
// accumulators, so the loads of tile i + 1, the MMAs of tile i and the TMA stores of tile i - 1 overlap; the fp32 result of a
// tile is staged in the F buffer it came from.  (First version: one CTA per (tile, slot), load -> MMA -> store in sequence,
// Ssym re-read per tile: 137 us for 462 MB at cfg1 = 0.51 of the HBM bandwidth.)
#define SGS_S     0                       // Ssym: tiles 0,1 = hi K-halves, 2,3 = lo
#define SGS_F(i)  ((4 + 4 * (i)) * TILE16)  // F buffer i: tiles 0,1 = hi K-halves, 2,3 = lo
#define SGS_SMEM (12 * TILE16 + 1024 + 128)
#define SB_SFULL      0u
#define SB_FFULL(i)   (8u + 8u * (i))
#define SB_FEMPTY(i)  (24u + 8u * (i))
#define SB_AFULL(i)   (40u + 8u * (i))
#define SB_AEMPTY(i)  (56u + 8u * (i))
#define SB_TMEM       72u

__global__ void __launch_bounds__(192, 1) sgrad_tc_split_kernel(const __grid_constant__ SGradTcArgs p, int nx) {
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t base = (smem_u32(tc_smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = tc_smem_raw + (base - smem_u32(tc_smem_raw));
    const uint32_t bar = base + 12 * TILE16;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int l = p.slot_of[blockIdx.y];
    // blockIdx.x = pair * nx + part: a CTA stays inside one pair (one Ssym per CTA)
    const int tpp = p.T / 128, pair = (int)blockIdx.x / nx, part = (int)blockIdx.x % nx;
    const int per = (tpp + nx - 1) / nx;
    const int tile_lo = pair * tpp + min(part * per, tpp), tile_hi = pair * tpp + min((part + 1) * per, tpp);
    const int n = max(tile_hi - tile_lo, 0);

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.tmF); tma_prefetch_desc(&p.tmS); tma_prefetch_desc(&p.tmFlo); tma_prefetch_desc(&p.tmSlo);
        tma_prefetch_desc(&p.tmDF);
        mbar_init(bar + SB_SFULL, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar + SB_FFULL(i), 1); mbar_init(bar + SB_FEMPTY(i), 4);
            mbar_init(bar + SB_AFULL(i), 1); mbar_init(bar + SB_AEMPTY(i), 4);
        }
        mbar_init_fence();
    }
    if (warp == 1) tmem_alloc<256>(bar + SB_TMEM);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t acc0 = *reinterpret_cast<volatile uint32_t*>(gen + (bar + SB_TMEM - base));

    if (warp == 0) {
        if (lane == 0 && n > 0) {
            const int b0 = pair;
            mbar_expect_tx(bar + SB_SFULL, 4 * TILE16);
            for (int j = 0; j < 2; ++j) {
                tma_load_2d(base + SGS_S + j * TILE16, &p.tmS, bar + SB_SFULL, j * 64, (b0 * p.L + l) * 128);
                tma_load_2d(base + SGS_S + (2 + j) * TILE16, &p.tmSlo, bar + SB_SFULL, j * 64, (b0 * p.L + l) * 128);
            }
            for (int it = 0; it < n; ++it) {
                const int i = it & 1;
                const int row0 = (tile_lo + it) * 128, b = row0 / p.T, t0 = row0 - b * p.T;
                mbar_wait(bar + SB_FEMPTY(i), ((uint32_t)(it >> 1) & 1u) ^ 1u);       // the stores of tile it - 2 have read the buffer
                const uint32_t fb = bar + SB_FFULL(i), f = base + SGS_F(i);
                mbar_expect_tx(fb, 4 * TILE16);
                for (int j = 0; j < 2; ++j) {
                    tma_load_4d(f + j * TILE16, &p.tmF, fb, j * 64, t0, b, l);
                    tma_load_4d(f + (2 + j) * TILE16, &p.tmFlo, fb, j * 64, t0, b, l);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && n > 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(128, 128);
            mbar_wait(bar + SB_SFULL, 0);
            for (int it = 0; it < n; ++it) {
                const int i = it & 1;
                const uint32_t ph = (uint32_t)(it >> 1) & 1u;
                mbar_wait(bar + SB_FFULL(i), ph);
                mbar_wait(bar + SB_AEMPTY(i), ph ^ 1u);                               // the epilogue of tile it - 2 has drained the accumulator
                tc_fence_after();
                const uint32_t f = base + SGS_F(i), acc = acc0 + (uint32_t)(i * 128);
#pragma unroll
                for (int j = 0; j < 2; ++j)
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) {
                        const uint64_t fh = umma_desc_kmajor_sw128(f + j * TILE16 + kk * 32);
                        const uint64_t fl = umma_desc_kmajor_sw128(f + (2 + j) * TILE16 + kk * 32);
                        const uint64_t sh = umma_desc_kmajor_sw128(base + SGS_S + j * TILE16 + kk * 32);
                        const uint64_t sl = umma_desc_kmajor_sw128(base + SGS_S + (2 + j) * TILE16 + kk * 32);
                        umma_f16(acc, fh, sh, idesc, (j > 0 || kk > 0) ? 1u : 0u);
                        umma_f16(acc, fl, sh, idesc, 1u);
                        umma_f16(acc, fh, sl, idesc, 1u);
                    }
                umma_commit(bar + SB_AFULL(i));
            }
        }
        __syncwarp();
    } else {
        const int q = warp & 3, r = q * 32 + lane;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        float v[32];
        for (int it = 0; it < n; ++it) {
            const int i = it & 1;
            const int row0 = (tile_lo + it) * 128, b = row0 / p.T, t0 = row0 - b * p.T;
            mbar_wait(bar + SB_AFULL(i), (uint32_t)(it >> 1) & 1u);                   // MMAs complete: the F buffer is dead, reuse it as staging
            tc_fence_after();
            uint8_t* o = gen + SGS_F(i);
#pragma unroll
            for (int cc = 0; cc < 4; ++cc) {
                tmem_ld32(acc0 + (uint32_t)(i * 128) + lane_off + cc * 32, v);
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    *reinterpret_cast<float4*>(o + cc * TILE16 + sw128_off(r, k)) = make_float4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
            }
            tc_fence_before();
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(bar + SB_AEMPTY(i));
#pragma unroll
                for (int cc = 0; cc < 4; ++cc)
                    tma_store_4d(&p.tmDF, base + SGS_F(i) + cc * TILE16 + q * 4096, cc * 32, t0 + q * 32, b, p.head_of_slot[l]);
                tma_store_commit();
                tma_store_wait_read();                                                // the stores have read the staging: the buffer may be refilled
                mbar_arrive(bar + SB_FEMPTY(i));
            }
            __syncwarp();
        }
        if (lane == 0) tma_store_wait_all();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<256>(acc0);
}

void launch_sgrad_tc(const SGradTcArgs& a, int ntiles, bool split, cudaStream_t s) {
    if (split) {
        // CTAs per (slot, pair): about one wave in all
        const int tpp = a.T / 128, npair = ntiles / tpp;
        const int nx = std::min(tpp, std::max(1, 148 / std::max(1, a.n_active * npair)));
        sgrad_tc_split_kernel<<<dim3(nx * npair, a.n_active), 192, SGS_SMEM, s>>>(a, nx);
    }
    else launch_pdl(sgrad_tc_kernel, dim3(ntiles, a.n_active), dim3(192), SG_SMEM, s, a);
    ast_launch_counter += 1;
}

int setup_gram_tc_kernels() {
    cudaError_t e = cudaFuncSetAttribute(gram_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GRAM_SMEM);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(gram_tc_split_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GRAMS_SMEM);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(sgrad_tc_split_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SGS_SMEM);
    if (e != cudaSuccess) return (int)e;
    return (int)cudaFuncSetAttribute(sgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SG_SMEM);
}
