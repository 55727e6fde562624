// stack_bwd.cuh -- the WHOLE backward sweep through the dilated stack (boundaries NL .. 0) in ONE cooperative
// launch; the counterpart of stack_fwd.cuh (same hand-off scheme, see there).  Included by kernels_tc.cu.
//   boundary l:  g += conv3_d^T(dz_l) * [h_l > 0] + dF_l (+ F_l . Ssym_l, fused Gatys style gradient)
//                dz_{l-1} = (g Wr_{l-1}^T) * [z_{l-1} > 0]
// CTA i owns time tile i for all boundaries:
//   * the fp32 running gradient g lives in TMEM (128 columns) and is updated in place by the epilogue; it is
//     written to global memory once, after boundary 0, for the start-conv adjoint;
//   * the CTA's own bf16 dz tile stays in shared memory (centre tap of the next boundary's transposed conv);
//     the +-d taps are the neighbours' tiles: published with full-sector stores + one release per warp,
//     consumed after an acquire poll of at most two per-tile flags;
//   * head gradients (fp32 dF tile, or the bf16 F / Ssym operands of the fused style gradient) do not depend
//     on the sweep: they are prefetched into their own 64 KB region one boundary ahead;
//   * weights and neighbour tiles stream through an 8 x 16 KB ring: all six weight tiles of a boundary are requested
//     before the neighbour flags are polled, and all four neighbour tiles right after.
#pragma once

struct alignas(64) BwdStackArgs {
    CUtensorMap tmDz;                // bf16 dz ping-pong [2][B][T][128], box {64,128,1,1}
    CUtensorMap tmW1, tmW2;          // backward weight operand rows (tc_weights_bwd_kernel), box {64,128}
    CUtensorMap tmDF;                // fp32 head gradients [n_head][B][T][128], box {32,128,1,1}
    CUtensorMap tmF, tmS;            // fused style gradient operands: bf16 F [L][B][T][128], bf16 Ssym rows
    __nv_bfloat16* dz;               // the buffer behind tmDz (published with plain stores)
    float* g;                        // [R][128]: receives the final running gradient (input of the start-conv adjoint) ...
    const float* W0; float* pdots;   // ... unless pdots != NULL: then p_k[t] = sum_c W0[k,0,c] g[t,c] goes to pdots[k][R] instead
    const uint32_t* hmask; const uint32_t* zmask;        // [NL+1][R][4], [NL][R][4]
    unsigned* flags; unsigned* err;
    int T, R, NL, tiles_per_pair, sg_L;
    int tile0;                       // global index of this launch's first tile (launches cover whole pairs)
    signed char df_idx[33];          // per boundary l: dF buffer added there, or -1
    signed char sg_slot[33];         // per boundary l: style slot whose F . Ssym is added there, or -1
};

// The own dz tile (centre-tap operand, read at the START of a boundary's transposed conv) and bf16(g') (operand of the 1x1
// GEMM, written after that conv has completed and dead once the 1x1 GEMM is done, i.e. before the next dz tile is staged)
// are never live together: they share one 32 KB buffer, which leaves room for an 8-slot ring.
#define BS_DZ 0                                   // [K-half] x [128][64] bf16: own dz tile / bf16(g')
#define BS_RT BS_DZ
#define BS_IO 32768                               // 64 KB: fp32 dF tile (4 column groups) or F (2) + Ssym (2) bf16 tiles
#define BS_RING 98304                             // 8 slots x 16 KB
#define BS_NSLOT 8
#define BS_BAR (BS_RING + BS_NSLOT * TILE16)
#define BS_SMEM (BS_BAR + 1024 + 1024)
#define BSB_FULL(i) (8u * (i))
#define BSB_EMPTY(i) (64u + 8u * (i))
#define BSB_DZ 128u
#define BSB_ACC1 136u
#define BSB_RT(j) (144u + 8u * (j))
#define BSB_ACC2 160u
#define BSB_IOFULL 168u
#define BSB_IOFREE 176u
#define BSB_ANTI 184u
#define BSB_TMEM 192u

// timeline (scripts/timeline_stack.py): detailed stamps of boundary g_tl_layer in block 2 of the buffer, one stamp per
// boundary (publish done) in block 3
#define TLB(i) do { if (g_timeline && l == g_tl_layer) g_timeline[((size_t)2 * gridDim.x + blockIdx.x) * 64 + (i)] = clock64(); } while (0)
#define TLC() do { if (g_timeline && l < 64) g_timeline[((size_t)3 * gridDim.x + blockIdx.x) * 64 + l] = clock64(); } while (0)
// bf16(v[0..7]) -> one 16-byte chunk at byte offset `off` of a swizzled tile
__device__ __forceinline__ void stage8_bf16(uint8_t* tile, uint32_t off, const float* v) {
    *reinterpret_cast<uint4*>(tile + off) =
        make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
}
// far-tap test of the anti-dependency (see the producer): taps of layer m reach beyond the adjacent tiles or wrap
__device__ __forceinline__ bool bs_anti_needed(int m) { const int e = m % 10; return e == 8 || e == 9 || e == 0; }

__global__ void __launch_bounds__(FWD_THREADS, 1) tc_bwd_stack_kernel(const __grid_constant__ BwdStackArgs p) {
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t base = (smem_u32(tc_smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = tc_smem_raw + (base - smem_u32(tc_smem_raw));
    const uint32_t s_dz = base + BS_DZ, s_rt = base + BS_RT, s_io = base + BS_IO, s_ring = base + BS_RING, bar = base + BS_BAR;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = p.tile0 + (int)blockIdx.x;
    const int row0 = tile * 128;
    const int b = row0 / p.T, t0 = row0 - b * p.T;
    const int NL = p.NL;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.tmDz); tma_prefetch_desc(&p.tmW1); tma_prefetch_desc(&p.tmW2); tma_prefetch_desc(&p.tmDF);
        tma_prefetch_desc(&p.tmF); tma_prefetch_desc(&p.tmS);
        for (int i = 0; i < BS_NSLOT; ++i) { mbar_init(bar + BSB_FULL(i), 1); mbar_init(bar + BSB_EMPTY(i), 1); }
        mbar_init(bar + BSB_DZ, 16); mbar_init(bar + BSB_ACC1, 1);
        mbar_init(bar + BSB_RT(0), 8); mbar_init(bar + BSB_RT(1), 8); mbar_init(bar + BSB_ACC2, 1);
        mbar_init(bar + BSB_IOFULL, 1); mbar_init(bar + BSB_IOFREE, 16); mbar_init(bar + BSB_ANTI, 1);
        mbar_init_fence();
    }
    if (warp == 1) tmem_alloc<512>(bar + BSB_TMEM);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gen + (bar + BSB_TMEM - base));
    const uint32_t acc1 = tmem, acc2 = tmem + 128, acc3 = tmem + 256, gacc = tmem + 384;

    if (warp == 0) {
        // ================================ TMA producer ========================================
        if (lane == 0) {
            uint32_t nfill = 0, io_n = 0;
            auto acquire = [&]() -> uint32_t {               // next ring slot, armed for 16 KB
                const uint32_t slot = nfill % BS_NSLOT, use = nfill / BS_NSLOT;
                if (use > 0) mbar_wait(bar + BSB_EMPTY(slot), (use - 1u) & 1u);
                ++nfill;
                mbar_expect_tx(bar + BSB_FULL(slot), TILE16);
                return slot;
            };
            for (int l = NL; l >= 0; --l) {
                const int s = NL - l;                         // publishes seen so far per tile: dz_l is publish #s
                const bool g1 = l < NL, g2 = l >= 1;
                const int d = 1 << (l % 10);
                const int df = p.df_idx[l], sg = p.sg_slot[l];
                TLB(0);
                if (df >= 0 || sg >= 0) {                     // head gradient operands: independent of the sweep
                    if (io_n > 0) mbar_wait(bar + BSB_IOFREE, (io_n - 1u) & 1u);
                    ++io_n;
                    mbar_expect_tx(bar + BSB_IOFULL, 4 * TILE16);
                    if (sg >= 0) {
                        for (int j = 0; j < 2; ++j) {
                            tma_load_4d(s_io + j * TILE16, &p.tmF, bar + BSB_IOFULL, j * 64, t0, b, sg);
                            tma_load_2d(s_io + (2 + j) * TILE16, &p.tmS, bar + BSB_IOFULL, j * 64, (b * p.sg_L + sg) * 128);
                        }
                    } else {
                        for (int cg = 0; cg < 4; ++cg) tma_load_4d(s_io + cg * TILE16, &p.tmDF, bar + BSB_IOFULL, cg * 32, t0, b, df);
                    }
                }
                if (g1) {
                    const int w1row = l * 3 * AST_C;
                    auto load_w1 = [&](int k, int j) {
                        const uint32_t slot = acquire();
                        tma_load_2d(s_ring + slot * TILE16, &p.tmW1, bar + BSB_FULL(slot), j * 64, w1row + k * 128);
                    };
                    // Fill order (= the MMA issuer's take order): all weight tiles of this boundary's transposed conv (centre
                    // tap, then every live side tap: at most 6 of the 8 slots, none of them waits for a neighbour), then
                    // ONE poll of all the neighbour flags, then the operand tiles of the side taps.
                    TLB(1);
                    load_w1(1, 0);                            // centre tap: the dz tile is resident
                    load_w1(1, 1);
                    const bool live0 = tap_live(t0, 0, -d, p.T), live2 = tap_live(t0, 2, -d, p.T);
                    if (live0) { load_w1(0, 0); load_w1(0, 1); }
                    if (live2) { load_w1(2, 0); load_w1(2, 1); }
                    TLB(2);
                    if (live0 || live2) {
                        stack_wait_rows2(p.flags, p.tiles_per_pair, p.T, p.err, b, live0 ? t0 + d : t0 - d, live2 ? t0 - d : t0 + d,
                                         FS_PUB * (unsigned)s);
                        fence_proxy_async_all();
                        TLB(3);
                        for (int k = 0; k < 3; k += 2) {
                            if (!(k == 0 ? live0 : live2)) continue;
                            const int t = t0 - (k - 1) * d;
                            for (int j = 0; j < 2; ++j) {
                                const uint32_t slot = acquire();
                                tma_load_4d(s_ring + slot * TILE16, &p.tmDz, bar + BSB_FULL(slot), j * 64, t, b, s & 1);
                            }
                        }
                        TLB(4);
                    }
                }
                if (g2) {
                    for (int j = 0; j < 2; ++j) {
                        const uint32_t slot = acquire();
                        tma_load_2d(s_ring + slot * TILE16, &p.tmW2, bar + BSB_FULL(slot), j * 64, (l - 1) * AST_C);
                    }
                    // Anti-dependency of this boundary's publish: dz_{l-1} overwrites ping-pong buffer (s&1)^1 = dz_{l+1},
                    // whose readers (tiles whose +-d_{l+1} taps cover this tile) are done once they have published dz_l.
                    // The adjacent tiles were waited for above unless d_l >= 128; the far taps need the extra look.
                    if (l + 1 < NL && bs_anti_needed(l + 1)) {
                        const int dp = 1 << ((l + 1) % 10);
                        stack_wait_rows2(p.flags, p.tiles_per_pair, p.T, p.err, b, t0 - dp, t0 + dp, FS_PUB * (unsigned)s);
                        mbar_arrive(bar + BSB_ANTI);
                    }
                    TLB(8);
                }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ==========================================
        if (lane == 0) {
            uint32_t nuse = 0, dz_n = 0, rt_n = 0, io_n = 0;
            auto take = [&]() -> uint32_t {
                const uint32_t slot = nuse % BS_NSLOT;
                mbar_wait(bar + BSB_FULL(slot), (nuse / BS_NSLOT) & 1u);
                ++nuse;
                return slot;
            };
            auto mma4 = [&](uint32_t acc, uint32_t a, uint32_t w, uint32_t accum0) {
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
                    umma_f16(acc, umma_desc_kmajor_sw128(a + kk * 32), umma_desc_kmajor_sw128(w + kk * 32), IDESC, (kk > 0) ? 1u : accum0);
            };
            for (int l = NL; l >= 0; --l) {
                const bool g1 = l < NL, g2 = l >= 1;
                const int d = 1 << (l % 10);
                const int df = p.df_idx[l], sg = p.sg_slot[l];
                if (g1) {
                    mbar_wait(bar + BSB_DZ, dz_n & 1u);       // own dz_l staged
                    ++dz_n;
                    tc_fence_after();
                    TLB(10);
                    for (int j = 0; j < 2; ++j) {
                        const uint32_t sw = take();
                        tc_fence_after();
                        TLB(11 + j);
                        mma4(acc1, s_dz + j * TILE16, s_ring + sw * TILE16, j > 0 ? 1u : 0u);
                        umma_commit(bar + BSB_EMPTY(sw));
                    }
                    uint32_t sw[2][2];
                    const bool live0 = tap_live(t0, 0, -d, p.T), live2 = tap_live(t0, 2, -d, p.T);
                    if (live0) { sw[0][0] = take(); sw[0][1] = take(); }        // weights first (producer fill order)
                    if (live2) { sw[1][0] = take(); sw[1][1] = take(); }
                    for (int k = 0; k < 3; k += 2) {
                        if (!(k == 0 ? live0 : live2)) continue;
                        for (int j = 0; j < 2; ++j) {
                            const uint32_t sa = take();
                            tc_fence_after();
                            TLB(13 + k + j);
                            mma4(acc1, s_ring + sa * TILE16, s_ring + sw[k >> 1][j] * TILE16, 1u);
                            umma_commit(bar + BSB_EMPTY(sa));
                            umma_commit(bar + BSB_EMPTY(sw[k >> 1][j]));
                        }
                    }
                }
                if (df >= 0 || sg >= 0) {
                    if (sg >= 0) {                            // acc3 = F_tile . Ssym
                        mbar_wait(bar + BSB_IOFULL, io_n & 1u);
                        tc_fence_after();
                        mma4(acc3, s_io, s_io + 2 * TILE16, 0u);
                        mma4(acc3, s_io + TILE16, s_io + 3 * TILE16, 1u);
                    }
                    ++io_n;
                }
                if (g1 || sg >= 0) umma_commit(bar + BSB_ACC1);
                TLB(17);
                if (g2) {
                    for (int j = 0; j < 2; ++j) {
                        mbar_wait(bar + BSB_RT(j), rt_n & 1u);
                        const uint32_t sw = take();
                        tc_fence_after();
                        TLB(18 + j);
                        mma4(acc2, s_rt + j * TILE16, s_ring + sw * TILE16, j > 0 ? 1u : 0u);
                        umma_commit(bar + BSB_EMPTY(sw));
                    }
                    ++rt_n;
                    umma_commit(bar + BSB_ACC2);
                    TLB(20);
                }
            }
        }
        __syncwarp();
    } else {
        // ================================ epilogue warps ======================================
        const int q = warp & 3, cc = (warp - 2) >> 2, half = cc >> 1;
        const int r = q * 32 + lane;
        const size_t row = (size_t)row0 + r;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        uint8_t* dzt = gen + BS_DZ + half * TILE16;
        uint8_t* rtt = gen + BS_RT + half * TILE16;
        unsigned* my_flag = p.flags + tile;
        uint32_t acc1_n = 0, acc2_n = 0, io_n = 0, anti_n = 0;
        for (int l = NL; l >= 0; --l) {
            const int s = NL - l;
            const bool g1 = l < NL, g2 = l >= 1;
            const int df = p.df_idx[l], sg = p.sg_slot[l];
            const bool has_io = df >= 0 || sg >= 0;
            const uint32_t hm = g1 ? p.hmask[((size_t)l * p.R + row) * AST_MW + cc] : 0u;
            const uint32_t zm = g2 ? p.zmask[((size_t)(l - 1) * p.R + row) * AST_MW + cc] : 0u;
            // ---- epilogue 1: g' = g + acc1 * [h>0] (+ acc3) + dF -> TMEM; bf16(g') -> A operand of the 1x1 GEMM ----
            if (g1 || sg >= 0) { mbar_wait(bar + BSB_ACC1, acc1_n & 1u); ++acc1_n; tc_fence_after(); }
            if (threadIdx.x == 128) TLB(30);
            if (has_io) { mbar_wait(bar + BSB_IOFULL, io_n & 1u); ++io_n; }
            if (threadIdx.x == 128) TLB(31);
#pragma unroll
            for (int sb = 0; sb < 2; ++sb) {
                const uint32_t col = lane_off + cc * 32 + sb * 16;
                float v[16];
                if (g1) {
                    tmem_ld16_nowait(acc1 + col, v);
                    tmem_wait_ld();
#pragma unroll
                    for (int e = 0; e < 16; ++e) if (!((hm >> (sb * 16 + e)) & 1u)) v[e] = 0.f;
                } else {
#pragma unroll
                    for (int e = 0; e < 16; ++e) v[e] = 0.f;
                }
                if (sg >= 0) {
                    float w[16];
                    tmem_ld16_nowait(acc3 + col, w);
                    tmem_wait_ld();
#pragma unroll
                    for (int e = 0; e < 16; ++e) v[e] += w[e];
                }
                if (g1) {                                     // l == NL starts the running gradient
                    float w[16];
                    tmem_ld16_nowait(gacc + col, w);
                    tmem_wait_ld();
#pragma unroll
                    for (int e = 0; e < 16; ++e) v[e] += w[e];
                }
                if (df >= 0) {
                    const uint8_t* dcg = gen + BS_IO + cc * TILE16;
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float4 t = *reinterpret_cast<const float4*>(dcg + sw128_off(r, sb * 4 + e));
                        v[4 * e] += t.x; v[4 * e + 1] += t.y; v[4 * e + 2] += t.z; v[4 * e + 3] += t.w;
                    }
                }
                tmem_st16(gacc + col, v);
                if (l == 0 && !p.pdots) {
                    float4* out = reinterpret_cast<float4*>(p.g + row * AST_C + cc * 32 + sb * 16);
#pragma unroll
                    for (int e = 0; e < 4; ++e) out[e] = make_float4(v[4 * e], v[4 * e + 1], v[4 * e + 2], v[4 * e + 3]);
                }
                if (g2) {
#pragma unroll
                    for (int c = 0; c < 2; ++c) stage8_bf16(rtt, sw128_off(r, (cc & 1) * 4 + sb * 2 + c), &v[8 * c]);
                }
            }
            if (has_io) { __syncwarp(); if (lane == 0) mbar_arrive(bar + BSB_IOFREE); }
            if (g2) {
                fence_proxy_async_smem();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar + BSB_RT(half));
            }
            tmem_wait_st();
            if (threadIdx.x == 128) TLB(32);
            if (l == 0 && p.pdots) {
                // start-conv adjoint dots p_k[t] = sum_c W0[k,0,c] g[t,c], from the final running gradient in TMEM (kept
                // out of the per-boundary loop above: executed once).  Each thread covers its 32 columns,
                // the 4 column groups of a row meet in smem (the dz / g' buffer is dead: boundary 0's transposed conv has
                // completed and there is no 1x1 GEMM below it); fixed order -> reproducible
                float pk0 = 0.f, pk1 = 0.f, pk2 = 0.f;
#pragma unroll 1
                for (int sb = 0; sb < 2; ++sb) {
                    float v[16];
                    tmem_ld16_nowait(gacc + lane_off + cc * 32 + sb * 16, v);
                    tmem_wait_ld();
                    const float* w = p.W0 + cc * 32 + sb * 16;
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        pk0 = fmaf(v[e], __ldg(w + e), pk0);
                        pk1 = fmaf(v[e], __ldg(w + AST_C + e), pk1);
                        pk2 = fmaf(v[e], __ldg(w + 2 * AST_C + e), pk2);
                    }
                }
                float* ps = reinterpret_cast<float*>(gen + BS_RT);
                ps[(0 * 4 + cc) * 128 + r] = pk0; ps[(1 * 4 + cc) * 128 + r] = pk1; ps[(2 * 4 + cc) * 128 + r] = pk2;
                named_bar_sync(1, 512);
                if (cc < 3) {
                    const float* q4 = ps + cc * 4 * 128 + r;
                    p.pdots[(size_t)cc * p.R + row] = (q4[0] + q4[128]) + (q4[256] + q4[384]);
                }
            }
            if (!g2) break;
            // ---- epilogue 2: dz_{l-1} = acc2 * [z>0] -> bf16: own centre-tap operand (smem) + published tile ----
            mbar_wait(bar + BSB_ACC2, acc2_n & 1u);
            ++acc2_n;
            tc_fence_after();
            if (threadIdx.x == 128) TLB(33);
#pragma unroll
            for (int sb = 0; sb < 2; ++sb) {
                float v[16];
                tmem_ld16_nowait(acc2 + lane_off + cc * 32 + sb * 16, v);
                tmem_wait_ld();
#pragma unroll
                for (int e = 0; e < 16; ++e) if (!((zm >> (sb * 16 + e)) & 1u)) v[e] = 0.f;
#pragma unroll
                for (int c = 0; c < 2; ++c) stage8_bf16(dzt, sw128_off(r, (cc & 1) * 4 + sb * 2 + c), &v[8 * c]);
            }
            fence_proxy_async_smem();
            tc_fence_before();
            __syncwarp();
            if (threadIdx.x == 128) TLB(34);
            if (l - 1 >= 0 && l >= 1) {
                if (l + 1 < NL && bs_anti_needed(l + 1)) { mbar_wait(bar + BSB_ANTI, anti_n & 1u); ++anti_n; }   // see producer
                // each warp publishes the region it staged itself: 32 rows x 64 bytes (see stack_fwd.cuh)
                const int ch = (cc & 1) * 4 + (lane & 3);
                const size_t gbase = ((size_t)((s & 1) ^ 1) * p.R + row0) * AST_C + half * 64 + ch * 8;
#pragma unroll
                for (int it = 0; it < 4; ++it) {
                    const int rr = q * 32 + it * 8 + (lane >> 2);
                    *reinterpret_cast<uint4*>(p.dz + gbase + (size_t)rr * AST_C) = *reinterpret_cast<const uint4*>(dzt + sw128_off(rr, ch));
                }
                __syncwarp();
                // release before the DZ arrival: cumulative flag counts then imply every warp finished this publish
                if (lane == 0) { red_release_gpu_add(my_flag, 1u); mbar_arrive(bar + BSB_DZ); }
                if (threadIdx.x == 128) { TLB(35); TLC(); }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<512>(tmem);
}
