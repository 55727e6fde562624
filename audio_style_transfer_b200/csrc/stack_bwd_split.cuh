// stack_bwd_split.cuh -- the backward sweep of stack_bwd.cuh with 2-term bf16 operand splits (hi + lo, three MMAs per
// product: hi hi + lo hi + hi lo, the lo lo term dropped), for the accurate tensor-core tier AST_PREC_TC3.  Included by
// kernels_tc.cu after stack_fwd.cuh / stack_bwd.cuh (same hand-off protocols, segments, groups and peer halos).
//   boundary l:  g += conv3_d^T(dz_l) * [h_l > 0] + dF_l          dz_{l-1} = (g Wr_{l-1}^T) * [z_{l-1} > 0]
// What differs from the bf16 kernel: dz and bf16(g') are kept as hi / lo pairs -- the operand windows, the tagged
// hand-off lines and the far tiles carry both parts, exactly as relu(h) does in the forward kernel, and the shared-memory
// budget is the forward kernel's (4 x 24 KB windows + 4 x 32 KB ring).  There is no room for a head-gradient staging area:
// the fp32 head gradients dF_l are read from global memory by the epilogue (every head of this tier leaves its gradient
// in a dF buffer; nothing is fused into this kernel).  fp32 accumulate in TMEM; the running gradient stays fp32 in TMEM.
// Reference: the TF gradient ops this replaces are Conv2DBackpropInput / ReluGrad / AddN behind tf.gradients
// (methods.py:133-137 -> ScipyOptimizerInterface), formulas SURVEY.md 8(a).
#pragma once

struct alignas(64) BwdSplitArgs {
    BwdStackArgs b;                  // tmDz / tmW1 / tmW2 / dz are the HI parts; tmDF, tmF, tmS, sg_slot are unused
    CUtensorMap tmDzLo, tmW1Lo, tmW2Lo;
    __nv_bfloat16* dz_lo;            // [2][R][128] ping-pong, lo part of the far tiles
    const float* dF;                 // fp32 head gradients [n_head][R][128]
};

#define S2_WBUF F2_WBUF
#define S2_RING (4 * S2_WBUF)
#define S2_NSLOT 4
#define S2_SLOT 32768                             // hi [128][64] | lo [128][64]
#define S2_BAR (S2_RING + S2_NSLOT * S2_SLOT)
#define S2_SMEM (S2_BAR + 1024 + 1024)

__global__ void __launch_bounds__(F2_THREADS, 1) tc_bwd_stack_split_kernel(const __grid_constant__ BwdSplitArgs q_) {
    const BwdStackArgs& p = q_.b;
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t base = (smem_u32(tc_smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = tc_smem_raw + (base - smem_u32(tc_smem_raw));
    const uint32_t s_ring = base + S2_RING, bar = base + S2_BAR;
    volatile unsigned* ab = reinterpret_cast<volatile unsigned*>(gen + S2_BAR + B2B_ABORT);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = p.tile0 + (int)blockIdx.x;
    const int row0 = tile * 128;
    const int b = row0 / p.T, t0 = row0 - b * p.T;
    const int NL = p.NL, LT = p.l_top, LBT = p.l_bot;
    const bool resume = LT < NL, stop = LBT > 0;
    const int gt0 = p.grp_t0, gt1 = p.grp_t1;
    const uint32_t epoch = *reinterpret_cast<const volatile unsigned*>(p.epoch);
    auto has_g1 = [&](int l) { return l < NL && !(resume && l == LT); };
    auto has_g2 = [&](int l) { return l >= 1 && !(stop && l == LBT); };
    auto df_of = [&](int l) { return (resume && l == LT) ? -1 : (int)p.df_idx[l]; };

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.tmDz); tma_prefetch_desc(&q_.tmDzLo); tma_prefetch_desc(&p.tmW1); tma_prefetch_desc(&q_.tmW1Lo);
        tma_prefetch_desc(&p.tmW2); tma_prefetch_desc(&q_.tmW2Lo);
        for (int i = 0; i < S2_NSLOT; ++i) { mbar_init(bar + B2B_FULL(i), 1); mbar_init(bar + B2B_EMPTY(i), 1); }
        for (int j = 0; j < 2; ++j) { mbar_init(bar + B2B_DZ(j), 16); mbar_init(bar + B2B_HALO(j), 16); mbar_init(bar + B2B_RT(j), 16); }
        mbar_init(bar + B2B_ACC1, 1); mbar_init(bar + B2B_ACC2, 1); mbar_init(bar + B2B_ANTI, 1);
        *ab = 0u;
        mbar_init_fence();
    }
    if (warp == 1) tmem_alloc<512>(bar + B2B_TMEM);
    if (warp >= 2) {
        // halo rows of the four window buffers start as zeros and stay zero where the pair (or the group) ends
        const int e = (warp - 2) * 32 + lane;
        for (int i = e; i < 4 * 2 * F2_HALO * 8; i += 512) {
            const int buf = i / (2 * F2_HALO * 8), rem = i % (2 * F2_HALO * 8), hr = rem >> 3, ch = rem & 7;
            const int wr = hr < F2_HALO ? hr : 128 + hr;
            *reinterpret_cast<uint4*>(gen + buf * S2_WBUF + wr * 128 + ch * 16) = make_uint4(0u, 0u, 0u, 0u);
        }
        fence_proxy_async_smem();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gen + S2_BAR + B2B_TMEM);
    const uint32_t acc1 = tmem, acc2 = tmem + 128, gacc = tmem + 384;
    auto win = [&](int j, int part) -> uint32_t { return base + (uint32_t)(j * 2 + part) * S2_WBUF; };

    if (warp == 0) {
        // ================================ ring producer: weights (hi | lo) and far-tap tiles (hi | lo) ================
        if (lane == 0) {
            uint32_t nfill = 0;
            auto acquire = [&]() -> uint32_t {               // next ring slot, armed for 32 KB
                const uint32_t slot = nfill % S2_NSLOT, use = nfill / S2_NSLOT;
                if (use > 0) stack_mbar_wait(bar + B2B_EMPTY(slot), (use - 1u) & 1u, ab, p.err);
                ++nfill;
                mbar_expect_tx(bar + B2B_FULL(slot), 2 * TILE16);
                return slot;
            };
            for (int l = LT; l >= LBT; --l) {
                const bool g1 = has_g1(l), g2 = has_g2(l);
                const int d = stack_dil(l);
                if (g1) {
                    const int w1row = l * 3 * AST_C;
                    auto load_w1 = [&](int k, int j) {
                        const uint32_t slot = acquire();
                        tma_load_2d(s_ring + slot * S2_SLOT, &p.tmW1, bar + B2B_FULL(slot), j * 64, w1row + k * 128);
                        tma_load_2d(s_ring + slot * S2_SLOT + TILE16, &q_.tmW1Lo, bar + B2B_FULL(slot), j * 64, w1row + k * 128);
                    };
                    load_w1(1, 0);                            // centre tap, both K-halves: the dz tile is resident
                    load_w1(1, 1);
                    if (stack_near(d)) {
                        for (int j = 0; j < 2; ++j) { load_w1(0, j); load_w1(2, j); }
                    } else {
                        // per live tap: weights of both K-halves, then (after ONE poll of the neighbour flags) its tiles
                        const bool live0 = stack_tap_live(t0, 0, -d, gt0, gt1), live2 = stack_tap_live(t0, 2, -d, gt0, gt1);
                        bool polled = false;
                        for (int k = 0; k < 3; k += 2) {
                            if (!(k == 0 ? live0 : live2)) continue;
                            for (int j = 0; j < 2; ++j) {
                                load_w1(k, j);
                                if (!polled) {
                                    stack_wait_rows2(p.flags, p.tiles_per_pair, gt0, gt1, b, live0 ? t0 + d : t0 - d, live2 ? t0 - d : t0 + d,
                                                     p.need[l], ab, p.err);
                                    fence_proxy_async_all();
                                    polled = true;
                                }
                                const uint32_t slot = acquire();
                                tma_load_4d(s_ring + slot * S2_SLOT, &p.tmDz, bar + B2B_FULL(slot), j * 64, t0 - (k - 1) * d, b, (NL - l) & 1);
                                tma_load_4d(s_ring + slot * S2_SLOT + TILE16, &q_.tmDzLo, bar + B2B_FULL(slot), j * 64, t0 - (k - 1) * d, b, (NL - l) & 1);
                            }
                        }
                    }
                }
                if (g2) {
                    for (int j = 0; j < 2; ++j) {
                        const uint32_t slot = acquire();
                        tma_load_2d(s_ring + slot * S2_SLOT, &p.tmW2, bar + B2B_FULL(slot), j * 64, (l - 1) * AST_C);
                        tma_load_2d(s_ring + slot * S2_SLOT + TILE16, &q_.tmW2Lo, bar + B2B_FULL(slot), j * 64, (l - 1) * AST_C);
                    }
                    // anti-dependency of a FAR publish of dz_{l-1} (see stack_bwd.cuh)
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
            uint32_t nuse = 0, dz_n = 0, rt_n = 0, n_near = 0;
            auto take = [&]() -> uint32_t {
                const uint32_t slot = nuse % S2_NSLOT;
                stack_mbar_wait(bar + B2B_FULL(slot), (nuse / S2_NSLOT) & 1u, ab, p.err);
                ++nuse;
                return slot;
            };
            // 12 MMAs of one K-half of one tap: (a_hi + a_lo) x (w_hi + w_lo) without the lo x lo term
            auto mma_slice = [&](uint32_t acc, uint32_t a_hi0, uint32_t a_lo0, uint32_t w, uint32_t accum0) {
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    const uint64_t a_hi = umma_desc_kmajor_sw128(a_hi0 + kk * 32);
                    const uint64_t a_lo = umma_desc_kmajor_sw128(a_lo0 + kk * 32);
                    const uint64_t w_hi = umma_desc_kmajor_sw128(w + kk * 32);
                    const uint64_t w_lo = umma_desc_kmajor_sw128(w + TILE16 + kk * 32);
                    umma_f16(acc, a_hi, w_hi, IDESC, (kk > 0) ? 1u : accum0);
                    umma_f16(acc, a_lo, w_hi, IDESC, 1u);
                    umma_f16(acc, a_hi, w_lo, IDESC, 1u);
                }
            };
            for (int l = LT; l >= LBT; --l) {
                const bool g1 = has_g1(l), g2 = has_g2(l);
                const int d = stack_dil(l);
                if (g1) {
                    for (int j = 0; j < 2; ++j) {             // centre tap, both K-halves
                        stack_mbar_wait(bar + B2B_DZ(j), dz_n & 1u, ab, p.err);
                        tc_fence_after();
                        const uint32_t sw = take();
                        tc_fence_after();
                        mma_slice(acc1, win(j, 0) + F2_CENTRE, win(j, 1) + F2_CENTRE, s_ring + sw * S2_SLOT, j > 0 ? 1u : 0u);
                        umma_commit(bar + B2B_EMPTY(sw));
                    }
                    if (stack_near(d)) {
                        for (int j = 0; j < 2; ++j) {
                            stack_mbar_wait(bar + B2B_HALO(j), n_near & 1u, ab, p.err);
                            tc_fence_after();
#pragma unroll
                            for (int k = 0; k < 3; k += 2) {
                                const uint32_t sw = take();
                                tc_fence_after();
                                const uint32_t ro = (uint32_t)(F2_HALO - (k - 1) * d) * 128u;
                                mma_slice(acc1, win(j, 0) + ro, win(j, 1) + ro, s_ring + sw * S2_SLOT, 1u);
                                umma_commit(bar + B2B_EMPTY(sw));
                            }
                        }
                        ++n_near;
                    } else {
                        const bool live0 = stack_tap_live(t0, 0, -d, gt0, gt1), live2 = stack_tap_live(t0, 2, -d, gt0, gt1);
                        for (int k = 0; k < 3; k += 2) {      // producer fill order: W A W A per live tap
                            if (!(k == 0 ? live0 : live2)) continue;
                            for (int j = 0; j < 2; ++j) {
                                const uint32_t sw = take();
                                const uint32_t sa = take();
                                tc_fence_after();
                                mma_slice(acc1, s_ring + sa * S2_SLOT, s_ring + sa * S2_SLOT + TILE16, s_ring + sw * S2_SLOT, 1u);
                                umma_commit(bar + B2B_EMPTY(sw));
                                umma_commit(bar + B2B_EMPTY(sa));
                            }
                        }
                    }
                    ++dz_n;
                    umma_commit(bar + B2B_ACC1);
                }
                if (g2) {
                    for (int j = 0; j < 2; ++j) {
                        stack_mbar_wait(bar + B2B_RT(j), rt_n & 1u, ab, p.err);
                        const uint32_t sw = take();
                        tc_fence_after();
                        mma_slice(acc2, win(j, 0) + F2_CENTRE, win(j, 1) + F2_CENTRE, s_ring + sw * S2_SLOT, j > 0 ? 1u : 0u);
                        umma_commit(bar + B2B_EMPTY(sw));
                    }
                    ++rt_n;
                    umma_commit(bar + B2B_ACC2);
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
        uint32_t acc1_n = 0, acc2_n = 0, anti_n = 0;
        const int peer_side = resume ? stack_peer_side(p.peer, t0, ab, p.err) : -1;
        // LL item of (ping-pong pp, tile tl, side sd, K-half j, cc, part, row slot): the forward kernel's layout
        auto ll_item = [&](int pp, int tl, int sd, int j, int slot, int c, int part) -> uint4* {
            return p.ll + ((((((size_t)pp * p.ntiles + tl) * 2 + sd) * 2 + j) * 8 + c * 2 + part) * 4) * LL_ROWS + slot;
        };
        // v[0..15] -> bf16 hi / lo, two 16-byte chunks each into the centre of window (j, hi / lo)
        auto stage16 = [&](int j, const float (&v)[16], uint32_t (&Hh)[8], uint32_t (&Ll)[8]) {
#pragma unroll
            for (int e = 0; e < 8; ++e) split_bf16x2(v[2 * e], v[2 * e + 1], Hh[e], Ll[e]);
            uint8_t* whi = gen + (j * 2 + 0) * S2_WBUF;
            uint8_t* wlo = gen + (j * 2 + 1) * S2_WBUF;
            *reinterpret_cast<uint4*>(whi + sw0) = make_uint4(Hh[0], Hh[1], Hh[2], Hh[3]);
            *reinterpret_cast<uint4*>(whi + sw1) = make_uint4(Hh[4], Hh[5], Hh[6], Hh[7]);
            *reinterpret_cast<uint4*>(wlo + sw0) = make_uint4(Ll[0], Ll[1], Ll[2], Ll[3]);
            *reinterpret_cast<uint4*>(wlo + sw1) = make_uint4(Ll[4], Ll[5], Ll[6], Ll[7]);
        };
        for (int l = LT; l >= LBT; --l) {
            const bool g1 = has_g1(l), g2 = has_g2(l);
            const int df = df_of(l);
            uint32_t hm[2] = {0u, 0u}, zm[2] = {0u, 0u};
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                if (g1) hm[j] = hmask16[((size_t)l * p.R + row) * 8 + j * 4 + cc];
                if (g2) zm[j] = zmask16[((size_t)(l - 1) * p.R + row) * 8 + j * 4 + cc];
            }
            // ---- epilogue 1: g' = g + acc1 * [h>0] + dF -> TMEM; hi/lo(g') -> A operand of the 1x1^T GEMM ----
            if (g1) { stack_mbar_wait(bar + B2B_ACC1, acc1_n & 1u, ab, p.err); ++acc1_n; tc_fence_after(); }
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
                    float w[16];
                    tmem_ld16_nowait(gacc + col, w);
                    tmem_wait_ld();
#pragma unroll
                    for (int e = 0; e < 16; ++e) v[e] += w[e];
                } else if (resume && l == LT) {               // a resumed launch takes g from global memory (own or neighbour's)
                    const float* gin = peer_side < 0 ? p.g_in + row * AST_C + colc
                                                     : p.peer.src[peer_side] + (size_t)(t0 + r + p.peer.row_off[peer_side]) * AST_C + colc;
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float4 t = peer_side < 0 ? *reinterpret_cast<const float4*>(gin + 4 * e) : ld_relaxed_sys_f4(gin + 4 * e);
                        v[4 * e] = t.x; v[4 * e + 1] = t.y; v[4 * e + 2] = t.z; v[4 * e + 3] = t.w;
                    }
                } else {                                      // l == NL starts the running gradient
#pragma unroll
                    for (int e = 0; e < 16; ++e) v[e] = 0.f;
                }
                if (df >= 0) {
                    const float4* dfp = reinterpret_cast<const float4*>(q_.dF + ((size_t)df * p.R + row) * AST_C + colc);
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float4 t = __ldg(dfp + e);
                        v[4 * e] += t.x; v[4 * e + 1] += t.y; v[4 * e + 2] += t.z; v[4 * e + 3] += t.w;
                    }
                }
                tmem_st16(gacc + col, v);
                if (g2) {
                    uint32_t Hh[8], Ll[8];
                    stage16(j, v, Hh, Ll);
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
                        const float* w = p.W0 + colc;         // start-conv adjoint dots over this thread's columns
#pragma unroll
                        for (int e = 0; e < 16; ++e) {
                            pk0 = fmaf(v[e], __ldg(w + e), pk0);
                            pk1 = fmaf(v[e], __ldg(w + AST_C + e), pk1);
                            pk2 = fmaf(v[e], __ldg(w + 2 * AST_C + e), pk2);
                        }
                    }
                }
            }
            tmem_wait_st();
            if (l == 0 && !stop && p.pdots) {
                float* ps = reinterpret_cast<float*>(gen + S2_RING);  // the ring is dead: boundary 0 has no 1x1 GEMM and its conv has completed
                ps[(0 * 4 + cc) * 128 + r] = pk0; ps[(1 * 4 + cc) * 128 + r] = pk1; ps[(2 * 4 + cc) * 128 + r] = pk2;
                named_bar_sync(5, 512);
                if (cc < 3 && wr) {
                    const float* q4 = ps + cc * 4 * 128 + r;
                    p.pdots[(size_t)cc * p.R + row] = (q4[0] + q4[128]) + (q4[256] + q4[384]);
                }
            }
            if (!g2) break;
            // ---- epilogue 2: dz_{l-1} = acc2 * [z>0] -> hi/lo: own window centre + hand-off to the neighbours ----
            stack_mbar_wait(bar + B2B_ACC2, acc2_n & 1u, ab, p.err);
            ++acc2_n;
            tc_fence_after();
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
                uint32_t Hh[8], Ll[8];
                stage16(j, v, Hh, Ll);
                fence_proxy_async_smem();
                tc_fence_before();
                if (near) {
                    if (r < dn || r >= 128 - dn) {
                        const int sd = r < dn ? 0 : 1, slot = r < dn ? r : r - (128 - dn);
                        uint4* d0 = ll_item((NL - m) & 1, tile, sd, j, slot, cc, 0);
#pragma unroll
                        for (int e = 0; e < 4; ++e) ll_store(d0 + e * LL_ROWS, Hh[2 * e], Hh[2 * e + 1], tag);
#pragma unroll
                        for (int e = 0; e < 4; ++e) ll_store(d0 + (4 + e) * LL_ROWS, Ll[2 * e], Ll[2 * e + 1], tag);
                    }
                    __syncwarp();
                    if (lane == 0) {
                        if (q == 0 || q == 3) red_relaxed_gpu_add(my_flag + j, 1u);      // bookkeeping for the anti-dependency polls only
                        mbar_arrive(bar + B2B_DZ(j));
                    }
                } else {
                    if (j == 0 && l + 1 < LT && !stack_near(stack_dil(l + 1))) { stack_mbar_wait(bar + B2B_ANTI, anti_n & 1u, ab, p.err); ++anti_n; }
                    named_bar_sync(1 + q, 128);
                    // far publish: the four warps of a row quadrant copy full 128-byte lines of the staged K-half (hi and lo), then ONE release
                    const uint8_t* whi = gen + (j * 2 + 0) * S2_WBUF + F2_CENTRE;
                    const uint8_t* wlo = gen + (j * 2 + 1) * S2_WBUF + F2_CENTRE;
                    const size_t g0 = ((size_t)((NL - m) & 1) * p.R + row0) * AST_C + j * 64 + (lane & 7) * 8;
#pragma unroll
                    for (int it = 0; it < 2; ++it) {
                        const int rr = q * 32 + cc * 8 + it * 4 + (lane >> 3);
                        const uint32_t so = sw128_off(rr, lane & 7);
                        *reinterpret_cast<uint4*>(p.dz + g0 + (size_t)rr * AST_C) = *reinterpret_cast<const uint4*>(whi + so);
                        *reinterpret_cast<uint4*>(q_.dz_lo + g0 + (size_t)rr * AST_C) = *reinterpret_cast<const uint4*>(wlo + so);
                    }
                    named_bar_sync(5, 512);
                    if (warp == 2 && lane == 0) red_release_gpu_add(my_flag + j, 1u);
                    if (lane == 0) mbar_arrive(bar + B2B_DZ(j));
                }
            }
            if (near) {
                // near hand-off, consumer side: the dn boundary rows of both neighbours -> window halo (stack_fwd.cuh: import_halo)
                const int e = (warp - 2) * 32 + lane;
                const int slot = e % dn, rest = e / dn, part = rest & 1, c = (rest >> 1) & 3, hs = rest >> 3;      // slot fastest: coalesced
                const bool active = hs < 2 && (hs == 0 ? has_up : has_dn);
                const uint4* src[2];
                uint4 vv[2][4];
                if (active) {                          // both K-halves' lines are requested at once
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        src[j] = ll_item((NL - m) & 1, hs == 0 ? tile - 1 : tile + 1, hs == 0 ? 1 : 0, j, slot, c, part);
                        ll_fetch(src[j], vv[j]);
                    }
                }
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    if (active) {
                        uint32_t w[8];
                        ll_wait_item(src[j], vv[j], tag, w, ab, p.err);
                        const int wr_ = hs == 0 ? F2_HALO - dn + slot : F2_HALO + 128 + slot;
                        uint8_t* wb = gen + (j * 2 + part) * S2_WBUF + wr_ * 128;
                        *reinterpret_cast<uint4*>(wb + (((2 * c) ^ (wr_ & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
                        *reinterpret_cast<uint4*>(wb + (((2 * c + 1) ^ (wr_ & 7)) << 4)) = make_uint4(w[4], w[5], w[6], w[7]);
                        fence_proxy_async_smem();
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar + B2B_HALO(j));
                }
            }
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
