// stack_bwd.cuh -- the WHOLE backward sweep through the dilated stack (boundaries NL .. 0) in ONE cooperative
// launch; the counterpart of stack_fwd.cuh (same hand-off and K-half pipelining, see there).  Included by kernels_tc.cu.
//   boundary l:  g += conv3_d^T(dz_l) * [h_l > 0] + dF_l (+ F_l . Ssym_l, fused Gatys style gradient)
//                dz_{l-1} = (g Wr_{l-1}^T) * [z_{l-1} > 0]
// CTA i owns time tile i for all boundaries:
//   * the fp32 running gradient g lives in TMEM (128 columns) and is updated in place by the epilogue; it leaves the
//     SM once, after boundary 0, as the start-conv adjoint's channel dots (or as g itself);
//   * dz (bf16) is kept as a 192-row operand window per K-half: rows [32, 160) the own tile, the halo rows the d
//     boundary rows of the neighbours.  Near boundaries (d <= 32): the three taps of the transposed conv are UMMA
//     descriptors into the window at row offsets 32+d, 32, 32-d; tiles hand over only their first / last d rows as
//     tagged LL lines that the epilogue warps of the neighbour poll and import (no fence, no flag: stack_fwd.cuh).
//     Far boundaries: whole tiles are published (full-line copies + ONE release per K-half) and streamed through the ring;
//   * bf16(g') (A operand of the 1x1^T GEMM) overlays the window centre: dead before the next dz tile is staged;
//   * head gradients (fp32 dF tile, or the bf16 F / Ssym operands of the fused style gradient) do not depend on the
//     sweep: prefetched into their own 64 KB region one boundary ahead;
//   * weights (and far-tap tiles) stream through a 7 x 16 KB ring.
// Segments and groups (see stack_fwd.cuh): one launch runs the boundaries l_top .. l_bot on a group of tiles.
//   l_top < NL ("resume"): boundary l_top starts from the running gradient in global memory (p.g, which already holds
//                          that boundary's conv^T and head terms) and only forms dz_{l_top-1};
//   l_bot > 0  ("stop")  : boundary l_bot adds its conv^T and head terms, writes g to global memory and ends there.
#pragma once

struct alignas(64) BwdStackArgs {
    CUtensorMap tmDz;                // bf16 dz ping-pong [2][B][T][128], box {64,128,1,1}: far taps
    CUtensorMap tmW1, tmW2;          // backward weight operand rows (tc_weights_bwd_kernel), box {64,128}
    CUtensorMap tmDF;                // fp32 head gradients [n_head][B][T][128], box {32,128,1,1}
    CUtensorMap tmF, tmS;            // fused style gradient operands: bf16 F [L][B][T][128], bf16 Ssym rows
    __nv_bfloat16* dz;               // the buffer behind tmDz (published with plain stores)
    const float* g_in;               // [R][128]: running gradient a resumed launch (l_top < NL) starts from; never the same
                                     // buffer as g: groups of one segment overlap, a later group still reads rows an earlier one owns
    float* g;                        // [R][128]: receives the final running gradient (input of the start-conv adjoint) ...
    const float* W0; float* pdots;   // ... unless pdots != NULL: then p_k[t] = sum_c W0[k,0,c] g[t,c] goes to pdots[k][R] instead
    const uint32_t* hmask; const uint32_t* zmask;        // [NL+1][R][4], [NL][R][4]
    uint4* ll;                       // near hand-off lines [2][ntiles][2 sides][2 K-halves][4 cc][4][32 rows] x 16 B
    unsigned* flags; unsigned* err;  // [ntiles][2]
    unsigned* epoch;                 // [0] launch epoch (tags), [1] CTAs done (see stack_fwd.cuh)
    int T, R, NL, tiles_per_pair, ntiles, sg_L;
    int tile0;                       // global index of this launch's first tile
    int l_top, l_bot;                // boundaries of this launch, l_top >= l_bot
    int grp_t0, grp_t1;              // in-pair time range covered by this launch's tiles (hand-off edges)
    int wr_t0, wr_t1;                // tiles with t0 in this range write the running gradient / the start-conv dots
    StackPeer peer;                  // peer-mapped halo rows of the running gradient (stack_fwd.cuh)
    unsigned short need[33];         // need[m]: flag value once dz_m (consumed by boundary m) has been published
    signed char df_idx[33];          // per boundary l: dF buffer added there, or -1
    signed char sg_slot[33];         // per boundary l: style slot whose F . Ssym is added there, or -1
};

#define B2_WBUF F2_WBUF                           // one K-half window buffer: [192 rows][128 B] bf16 dz
#define B2_IO (2 * B2_WBUF)                       // 64 KB: fp32 dF tile (4 column groups) or F (2) + Ssym (2) bf16 tiles
#define B2_RING (B2_IO + 65536)
#define B2_NSLOT 7
#define B2_BAR (B2_RING + B2_NSLOT * TILE16)
#define B2_SMEM (B2_BAR + 1024 + 1024)
#define B2B_FULL(i) (8u * (i))
#define B2B_EMPTY(i) (64u + 8u * (i))
#define B2B_DZ(j) (128u + 8u * (j))
#define B2B_HALO(j) (144u + 8u * (j))
#define B2B_ACC1 160u
#define B2B_RT(j) (168u + 8u * (j))
#define B2B_ACC2 184u
#define B2B_IOFULL 192u
#define B2B_IOFREE 200u
#define B2B_ANTI 208u
#define B2B_TMEM 216u
#define B2B_ABORT 224u

#ifdef AST_EXPERIMENTS
// timeline (scripts/timeline_stack.py): detailed stamps of boundary g_tl_layer in block 2 of the buffer, one stamp per
// boundary (publish done) in block 3
#define TLB(i) do { if (g_timeline && l == g_tl_layer) g_timeline[((size_t)2 * gridDim.x + blockIdx.x) * 64 + (i)] = clock64(); } while (0)
#define TLC() do { if (g_timeline && l < 64) g_timeline[((size_t)3 * gridDim.x + blockIdx.x) * 64 + l] = clock64(); } while (0)
#else
#define TLB(i) do { } while (0)
#define TLC() do { } while (0)
#endif

__global__ void __launch_bounds__(F2_THREADS, 1) tc_bwd_stack_kernel(const __grid_constant__ BwdStackArgs p) {
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t base = (smem_u32(tc_smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = tc_smem_raw + (base - smem_u32(tc_smem_raw));
    const uint32_t s_io = base + B2_IO, s_ring = base + B2_RING, bar = base + B2_BAR;
    volatile unsigned* ab = reinterpret_cast<volatile unsigned*>(gen + B2_BAR + B2B_ABORT);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = p.tile0 + (int)blockIdx.x;
    const int row0 = tile * 128;
    const int b = row0 / p.T, t0 = row0 - b * p.T;
    const int NL = p.NL, LT = p.l_top, LBT = p.l_bot;
    const bool resume = LT < NL, stop = LBT > 0;
    const int gt0 = p.grp_t0, gt1 = p.grp_t1;
    const uint32_t epoch = *reinterpret_cast<const volatile unsigned*>(p.epoch);
    // what boundary l does in THIS launch: conv^T + head terms (g1), the 1x1^T that forms dz_{l-1} (g2)
    auto has_g1 = [&](int l) { return l < NL && !(resume && l == LT); };
    auto has_g2 = [&](int l) { return l >= 1 && !(stop && l == LBT); };
    auto df_of = [&](int l) { return (resume && l == LT) ? -1 : (int)p.df_idx[l]; };
    auto sg_of = [&](int l) { return (resume && l == LT) ? -1 : (int)p.sg_slot[l]; };

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.tmDz); tma_prefetch_desc(&p.tmW1); tma_prefetch_desc(&p.tmW2); tma_prefetch_desc(&p.tmDF);
        tma_prefetch_desc(&p.tmF); tma_prefetch_desc(&p.tmS);
        for (int i = 0; i < B2_NSLOT; ++i) { mbar_init(bar + B2B_FULL(i), 1); mbar_init(bar + B2B_EMPTY(i), 1); }
        for (int j = 0; j < 2; ++j) { mbar_init(bar + B2B_DZ(j), 16); mbar_init(bar + B2B_HALO(j), 16); mbar_init(bar + B2B_RT(j), 16); }
        mbar_init(bar + B2B_ACC1, 1); mbar_init(bar + B2B_ACC2, 1);
        mbar_init(bar + B2B_IOFULL, 1); mbar_init(bar + B2B_IOFREE, 16); mbar_init(bar + B2B_ANTI, 1);
        *ab = 0u;
        mbar_init_fence();
    }
    if (warp == 1) tmem_alloc<512>(bar + B2B_TMEM);
    if (warp >= 2) {
        // halo rows of the dz window start as zeros and stay zero where the pair ends (the transposed conv's padding)
        const int e = (warp - 2) * 32 + lane;
        for (int i = e; i < 2 * 2 * F2_HALO * 8; i += 512) {
            const int buf = i / (2 * F2_HALO * 8), rem = i % (2 * F2_HALO * 8), hr = rem >> 3, ch = rem & 7;
            const int wr = hr < F2_HALO ? hr : 128 + hr;
            *reinterpret_cast<uint4*>(gen + buf * B2_WBUF + wr * 128 + ch * 16) = make_uint4(0u, 0u, 0u, 0u);
        }
        fence_proxy_async_smem();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gen + B2_BAR + B2B_TMEM);
    const uint32_t acc1 = tmem, acc2 = tmem + 128, acc3 = tmem + 256, gacc = tmem + 384;
    auto win = [&](int j) -> uint32_t { return base + (uint32_t)j * B2_WBUF; };

    if (warp == 0) {
        // ================================ ring producer: weights, far-tap tiles, head-gradient operands ===========
        if (lane == 0) {
            uint32_t nfill = 0, io_n = 0;
            auto acquire = [&]() -> uint32_t {               // next ring slot, armed for 16 KB
                const uint32_t slot = nfill % B2_NSLOT, use = nfill / B2_NSLOT;
                if (use > 0) stack_mbar_wait(bar + B2B_EMPTY(slot), (use - 1u) & 1u, ab, p.err);
                ++nfill;
                mbar_expect_tx(bar + B2B_FULL(slot), TILE16);
                return slot;
            };
            for (int l = LT; l >= LBT; --l) {
                const bool g1 = has_g1(l), g2 = has_g2(l);
                const int d = stack_dil(l);
                const int df = df_of(l), sg = sg_of(l);
                TLB(0);
                if (df >= 0 || sg >= 0) {                     // head gradient operands: independent of the sweep
                    if (io_n > 0) stack_mbar_wait(bar + B2B_IOFREE, (io_n - 1u) & 1u, ab, p.err);
                    ++io_n;
                    mbar_expect_tx(bar + B2B_IOFULL, 4 * TILE16);
                    if (sg >= 0) {
                        for (int j = 0; j < 2; ++j) {
                            tma_load_4d(s_io + j * TILE16, &p.tmF, bar + B2B_IOFULL, j * 64, t0, b, sg);
                            tma_load_2d(s_io + (2 + j) * TILE16, &p.tmS, bar + B2B_IOFULL, j * 64, (b * p.sg_L + sg) * 128);
                        }
                    } else {
                        for (int cg = 0; cg < 4; ++cg) tma_load_4d(s_io + cg * TILE16, &p.tmDF, bar + B2B_IOFULL, cg * 32, t0, b, df);
                    }
                }
                if (g1) {
                    const int w1row = l * 3 * AST_C;
                    auto load_w1 = [&](int k, int j) {
                        const uint32_t slot = acquire();
                        tma_load_2d(s_ring + slot * TILE16, &p.tmW1, bar + B2B_FULL(slot), j * 64, w1row + k * 128);
                    };
                    load_w1(1, 0);                            // centre tap, both K-halves: the dz tile is resident
                    load_w1(1, 1);
                    if (stack_near(d)) {
                        for (int j = 0; j < 2; ++j) { load_w1(0, j); load_w1(2, j); }
                    } else {
                        // all weight tiles of this boundary's transposed conv (at most 6 of the 7 slots, none of them waits for a
                        // neighbour), ONE poll of the neighbour flags, then the tap tiles
                        const bool live0 = stack_tap_live(t0, 0, -d, gt0, gt1), live2 = stack_tap_live(t0, 2, -d, gt0, gt1);
                        if (live0) { load_w1(0, 0); load_w1(0, 1); }
                        if (live2) { load_w1(2, 0); load_w1(2, 1); }
                        if (live0 || live2) {
                            stack_wait_rows2(p.flags, p.tiles_per_pair, gt0, gt1, b, live0 ? t0 + d : t0 - d, live2 ? t0 - d : t0 + d,
                                             p.need[l], ab, p.err);
                            fence_proxy_async_all();
                            TLB(3);
                            for (int k = 0; k < 3; k += 2) {
                                if (!(k == 0 ? live0 : live2)) continue;
                                for (int j = 0; j < 2; ++j) {
                                    const uint32_t slot = acquire();
                                    tma_load_4d(s_ring + slot * TILE16, &p.tmDz, bar + B2B_FULL(slot), j * 64, t0 - (k - 1) * d, b, (NL - l) & 1);
                                }
                            }
                        }
                    }
                }
                if (g2) {
                    for (int j = 0; j < 2; ++j) {
                        const uint32_t slot = acquire();
                        tma_load_2d(s_ring + slot * TILE16, &p.tmW2, bar + B2B_FULL(slot), j * 64, (l - 1) * AST_C);
                    }
                    TLB(8);
                    // Anti-dependency of a FAR publish of dz_{l-1} (whole tile -> the ping-pong buffer that holds dz_{l+1} if that
                    // was far too): the tiles whose +-d_{l+1} taps read this tile are done once they have published dz_l.
                    if (l + 1 < LT && !stack_near(stack_dil(l - 1)) && !stack_near(stack_dil(l + 1))) {
                        const int dp = stack_dil(l + 1);
                        stack_wait_rows2(p.flags, p.tiles_per_pair, gt0, gt1, b, t0 - dp, t0 + dp, p.need[l], ab, p.err);
                        mbar_arrive(bar + B2B_ANTI);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ==========================================
        if (lane == 0) {
            uint32_t nuse = 0, dz_n = 0, rt_n = 0, io_n = 0, n_near = 0;
            auto take = [&]() -> uint32_t {
                const uint32_t slot = nuse % B2_NSLOT;
                stack_mbar_wait(bar + B2B_FULL(slot), (nuse / B2_NSLOT) & 1u, ab, p.err);
                ++nuse;
                return slot;
            };
            auto mma4 = [&](uint32_t acc, uint32_t a, uint32_t w, uint32_t accum0) {
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
                    umma_f16(acc, umma_desc_kmajor_sw128(a + kk * 32), umma_desc_kmajor_sw128(w + kk * 32), IDESC, (kk > 0) ? 1u : accum0);
            };
            for (int l = LT; l >= LBT; --l) {
                const bool g1 = has_g1(l), g2 = has_g2(l);
                const int d = stack_dil(l);
                const int df = df_of(l), sg = sg_of(l);
                if (g1) {
                    for (int j = 0; j < 2; ++j) {             // centre tap, both K-halves
                        stack_mbar_wait(bar + B2B_DZ(j), dz_n & 1u, ab, p.err);          // own dz_l, K-half j, staged
                        tc_fence_after();
                        TLB(10 + j);
                        const uint32_t sw = take();
                        tc_fence_after();
                        mma4(acc1, win(j) + F2_CENTRE, s_ring + sw * TILE16, j > 0 ? 1u : 0u);
                        umma_commit(bar + B2B_EMPTY(sw));
                    }
                    if (stack_near(d)) {
                        for (int j = 0; j < 2; ++j) {
                            stack_mbar_wait(bar + B2B_HALO(j), n_near & 1u, ab, p.err);
                            tc_fence_after();
                            TLB(12 + j);
#pragma unroll
                            for (int k = 0; k < 3; k += 2) {
                                const uint32_t sw = take();
                                tc_fence_after();
                                mma4(acc1, win(j) + (uint32_t)(F2_HALO - (k - 1) * d) * 128u, s_ring + sw * TILE16, 1u);
                                umma_commit(bar + B2B_EMPTY(sw));
                            }
                        }
                        ++n_near;
                    } else {
                        uint32_t sw[2][2];
                        const bool live0 = stack_tap_live(t0, 0, -d, gt0, gt1), live2 = stack_tap_live(t0, 2, -d, gt0, gt1);
                        if (live0) { sw[0][0] = take(); sw[0][1] = take(); }        // weights first (producer fill order)
                        if (live2) { sw[1][0] = take(); sw[1][1] = take(); }
                        for (int k = 0; k < 3; k += 2) {
                            if (!(k == 0 ? live0 : live2)) continue;
                            for (int j = 0; j < 2; ++j) {
                                const uint32_t sa = take();
                                tc_fence_after();
                                TLB(14 + k + j);
                                mma4(acc1, s_ring + sa * TILE16, s_ring + sw[k >> 1][j] * TILE16, 1u);
                                umma_commit(bar + B2B_EMPTY(sa));
                                umma_commit(bar + B2B_EMPTY(sw[k >> 1][j]));
                            }
                        }
                    }
                    ++dz_n;
                }
                if (df >= 0 || sg >= 0) {
                    if (sg >= 0) {                            // acc3 = F_tile . Ssym
                        stack_mbar_wait(bar + B2B_IOFULL, io_n & 1u, ab, p.err);
                        tc_fence_after();
                        mma4(acc3, s_io, s_io + 2 * TILE16, 0u);
                        mma4(acc3, s_io + TILE16, s_io + 3 * TILE16, 1u);
                    }
                    ++io_n;
                }
                if (g1 || sg >= 0) umma_commit(bar + B2B_ACC1);
                TLB(18);
                if (g2) {
                    for (int j = 0; j < 2; ++j) {
                        stack_mbar_wait(bar + B2B_RT(j), rt_n & 1u, ab, p.err);
                        const uint32_t sw = take();
                        tc_fence_after();
                        TLB(19 + j);
                        mma4(acc2, win(j) + F2_CENTRE, s_ring + sw * TILE16, j > 0 ? 1u : 0u);
                        umma_commit(bar + B2B_EMPTY(sw));
                    }
                    ++rt_n;
                    umma_commit(bar + B2B_ACC2);
                    TLB(21);
                }
            }
        }
        __syncwarp();
    } else {
        // ================================ epilogue warps ======================================
        const int q = warp & 3, cc = (warp - 2) >> 2;                // TMEM lane quadrant, 16-column group of a K-half
        const int r = q * 32 + lane;
        const size_t row = (size_t)row0 + r;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        const uint32_t sw0 = F2_CENTRE + sw128_off(r, 2 * cc), sw1 = F2_CENTRE + sw128_off(r, 2 * cc + 1);
        unsigned* my_flag = p.flags + (size_t)tile * 2;
        const uint16_t* hmask16 = reinterpret_cast<const uint16_t*>(p.hmask);
        const uint16_t* zmask16 = reinterpret_cast<const uint16_t*>(p.zmask);
        const bool has_up = t0 > gt0, has_dn = t0 + 128 < gt1;
        const bool wr = t0 >= p.wr_t0 && t0 < p.wr_t1;
        uint32_t acc1_n = 0, acc2_n = 0, io_n = 0, anti_n = 0;
        const int peer_side = resume ? stack_peer_side(p.peer, t0, ab, p.err) : -1;
        // LL item of (ping-pong pp, tile tl, side sd, K-half j, cc, row slot): lines [4][LL_ROWS], slot fastest (stack_fwd.cuh)
        auto ll_item = [&](int pp, int tl, int sd, int j, int slot, int c) -> uint4* {
            return p.ll + ((((((size_t)pp * p.ntiles + tl) * 2 + sd) * 2 + j) * 4 + c) * 4) * LL_ROWS + slot;
        };
        for (int l = LT; l >= LBT; --l) {
            const bool g1 = has_g1(l), g2 = has_g2(l);
            const int df = df_of(l), sg = sg_of(l);
            const bool has_io = df >= 0 || sg >= 0;
            uint32_t hm[2] = {0u, 0u}, zm[2] = {0u, 0u};
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                if (g1) hm[j] = hmask16[((size_t)l * p.R + row) * 8 + j * 4 + cc];
                if (g2) zm[j] = zmask16[((size_t)(l - 1) * p.R + row) * 8 + j * 4 + cc];
            }
            // ---- epilogue 1: g' = g + acc1 * [h>0] (+ acc3) + dF -> TMEM; bf16(g') -> A operand of the 1x1^T GEMM ----
            if (g1 || sg >= 0) { stack_mbar_wait(bar + B2B_ACC1, acc1_n & 1u, ab, p.err); ++acc1_n; tc_fence_after(); }
            if (threadIdx.x == 64) TLB(30);
            if (has_io) { stack_mbar_wait(bar + B2B_IOFULL, io_n & 1u, ab, p.err); ++io_n; }
            if (threadIdx.x == 64) TLB(31);
            float pk0 = 0.f, pk1 = 0.f, pk2 = 0.f;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                if (j == 1) named_bar_sync(1 + q, 128);       // K-half 0 of the whole quadrant before anyone starts K-half 1
                const int colc = j * 64 + cc * 16;
                const uint32_t col = lane_off + colc;
                float v[16];
                if (g1) {
                    tmem_ld16_nowait(acc1 + col, v);
                    tmem_wait_ld();
#pragma unroll
                    for (int e = 0; e < 16; ++e) if (!((hm[j] >> e) & 1u)) v[e] = 0.f;
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
                } else if (resume && l == LT) {               // ... a resumed launch takes it from global memory (own or neighbour's)
                    const float* gin = peer_side < 0 ? p.g_in + row * AST_C + colc
                                                     : p.peer.src[peer_side] + (size_t)(t0 + r + p.peer.row_off[peer_side]) * AST_C + colc;
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float4 t = peer_side < 0 ? *reinterpret_cast<const float4*>(gin + 4 * e) : ld_relaxed_sys_f4(gin + 4 * e);
                        v[4 * e] = t.x; v[4 * e + 1] = t.y; v[4 * e + 2] = t.z; v[4 * e + 3] = t.w;
                    }
                }
                if (df >= 0) {
                    const uint8_t* dcg = gen + B2_IO + (j * 2 + (cc >> 1)) * TILE16;
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float4 t = *reinterpret_cast<const float4*>(dcg + sw128_off(r, (cc & 1) * 4 + e));
                        v[4 * e] += t.x; v[4 * e + 1] += t.y; v[4 * e + 2] += t.z; v[4 * e + 3] += t.w;
                    }
                }
                tmem_st16(gacc + col, v);
                if (g2) {
                    uint8_t* w = gen + j * B2_WBUF;
                    *reinterpret_cast<uint4*>(w + sw0) =
                        make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
                    *reinterpret_cast<uint4*>(w + sw1) =
                        make_uint4(pack_bf16(v[8], v[9]), pack_bf16(v[10], v[11]), pack_bf16(v[12], v[13]), pack_bf16(v[14], v[15]));
                    fence_proxy_async_smem();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar + B2B_RT(j));
                }
                if (l == LBT) {
                    if (stop || !p.pdots) {
                        if (wr) {
                            float4* out = reinterpret_cast<float4*>(p.g + row * AST_C + colc);
#pragma unroll
                            for (int e = 0; e < 4; ++e) out[e] = make_float4(v[4 * e], v[4 * e + 1], v[4 * e + 2], v[4 * e + 3]);
                        }
                    } else {
                        // start-conv adjoint dots p_k[t] = sum_c W0[k,0,c] g[t,c] over this thread's columns
                        const float* w = p.W0 + colc;
#pragma unroll
                        for (int e = 0; e < 16; ++e) {
                            pk0 = fmaf(v[e], __ldg(w + e), pk0);
                            pk1 = fmaf(v[e], __ldg(w + AST_C + e), pk1);
                            pk2 = fmaf(v[e], __ldg(w + 2 * AST_C + e), pk2);
                        }
                    }
                }
            }
            if (has_io) { __syncwarp(); if (lane == 0) mbar_arrive(bar + B2B_IOFREE); }
            tmem_wait_st();
            if (threadIdx.x == 64) TLB(32);
            if (l == 0 && !stop && p.pdots) {
                // the 4 column groups of a row meet in smem (the ring is dead: boundary 0 has no 1x1 GEMM and its conv has
                // completed); fixed order -> reproducible
                float* ps = reinterpret_cast<float*>(gen + B2_RING);
                ps[(0 * 4 + cc) * 128 + r] = pk0; ps[(1 * 4 + cc) * 128 + r] = pk1; ps[(2 * 4 + cc) * 128 + r] = pk2;
                named_bar_sync(5, 512);
                if (cc < 3 && wr) {
                    const float* q4 = ps + cc * 4 * 128 + r;
                    p.pdots[(size_t)cc * p.R + row] = (q4[0] + q4[128]) + (q4[256] + q4[384]);
                }
            }
            if (!g2) break;
            // ---- epilogue 2: dz_{l-1} = acc2 * [z>0] -> bf16: own window centre + hand-off to the neighbours ----
            stack_mbar_wait(bar + B2B_ACC2, acc2_n & 1u, ab, p.err);
            ++acc2_n;
            tc_fence_after();
            if (threadIdx.x == 64) TLB(33);
            const int m = l - 1;                              // the boundary that consumes dz_{l-1}
            const int dn = stack_dil(m);
            const bool near = stack_near(dn);
            const uint32_t tag = (epoch << 6) | (uint32_t)(m + 1);
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                if (j == 1) named_bar_sync(1 + q, 128);
                const int colc = j * 64 + cc * 16;
                float v[16];
                tmem_ld16_nowait(acc2 + lane_off + colc, v);
                tmem_wait_ld();
#pragma unroll
                for (int e = 0; e < 16; ++e) if (!((zm[j] >> e) & 1u)) v[e] = 0.f;
                uint32_t P[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) P[e] = pack_bf16(v[2 * e], v[2 * e + 1]);
                uint8_t* w = gen + j * B2_WBUF;
                *reinterpret_cast<uint4*>(w + sw0) = make_uint4(P[0], P[1], P[2], P[3]);
                *reinterpret_cast<uint4*>(w + sw1) = make_uint4(P[4], P[5], P[6], P[7]);
                fence_proxy_async_smem();
                tc_fence_before();
                if (near) {
                    if (r < dn || r >= 128 - dn) {
                        const int sd = r < dn ? 0 : 1, slot = r < dn ? r : r - (128 - dn);
                        uint4* d0 = ll_item((NL - m) & 1, tile, sd, j, slot, cc);
#pragma unroll
                        for (int e = 0; e < 4; ++e) ll_store(d0 + e * LL_ROWS, P[2 * e], P[2 * e + 1], tag);
                    }
                    __syncwarp();
                    if (lane == 0) {
                        if (q == 0 || q == 3) red_relaxed_gpu_add(my_flag + j, 1u);      // bookkeeping for the anti-dependency polls only
                        mbar_arrive(bar + B2B_DZ(j));
                    }
                } else {
                    if (j == 0 && l + 1 < LT && !stack_near(stack_dil(l + 1))) { stack_mbar_wait(bar + B2B_ANTI, anti_n & 1u, ab, p.err); ++anti_n; }
                    named_bar_sync(1 + q, 128);
                    // far publish: the four warps of a row quadrant copy full 128-byte lines of the staged K-half, then ONE release
                    const uint8_t* wc = gen + j * B2_WBUF + F2_CENTRE;
                    const size_t g0 = ((size_t)((NL - m) & 1) * p.R + row0) * AST_C + j * 64 + (lane & 7) * 8;
#pragma unroll
                    for (int it = 0; it < 2; ++it) {
                        const int rr = q * 32 + cc * 8 + it * 4 + (lane >> 3);
                        *reinterpret_cast<uint4*>(p.dz + g0 + (size_t)rr * AST_C) = *reinterpret_cast<const uint4*>(wc + sw128_off(rr, lane & 7));
                    }
                    named_bar_sync(5, 512);
                    if (warp == 2 && lane == 0) red_release_gpu_add(my_flag + j, 1u);
                    if (lane == 0) mbar_arrive(bar + B2B_DZ(j));
                }
                if (threadIdx.x == 64) { TLB(34 + j); }
            }
            if (near) {
                // near hand-off, consumer side: the dn boundary rows of both neighbours -> window halo (see stack_fwd.cuh)
                const int e = (warp - 2) * 32 + lane;
                const int slot = e % dn, rest = e / dn, c = rest & 3, hs = rest >> 2;      // slot fastest: coalesced
                const bool active = hs < 2 && (hs == 0 ? has_up : has_dn);
                const uint4* src[2];
                uint4 vv[2][4];
                if (active) {                          // both K-halves' lines are requested at once
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        src[j] = ll_item((NL - m) & 1, hs == 0 ? tile - 1 : tile + 1, hs == 0 ? 1 : 0, j, slot, c);
                        ll_fetch(src[j], vv[j]);
                    }
                }
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    if (active) {
                        uint32_t w[8];
                        ll_wait_item(src[j], vv[j], tag, w, ab, p.err);
                        const int wr = hs == 0 ? F2_HALO - dn + slot : F2_HALO + 128 + slot;
                        uint8_t* wb = gen + j * B2_WBUF + wr * 128;
                        *reinterpret_cast<uint4*>(wb + (((2 * c) ^ (wr & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
                        *reinterpret_cast<uint4*>(wb + (((2 * c + 1) ^ (wr & 7)) << 4)) = make_uint4(w[4], w[5], w[6], w[7]);
                        fence_proxy_async_smem();
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar + B2B_HALO(j));
                    if (threadIdx.x == 64) TLB(24 + j);
                }
            }
            if (threadIdx.x == 64) TLC();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(p.epoch + 1, 1u) == gridDim.x - 1) { p.epoch[1] = 0u; __threadfence(); atomicAdd(p.epoch, 1u); stack_peer_signal(p.peer); }
    }
    if (warp == 1) tmem_dealloc<512>(tmem);
}
