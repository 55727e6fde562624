// stack_fwd2.cuh -- whole-stack forward with TWO pairs per CTA ("duo"): CTA = the same time tile of two independent
// pairs.  Included by kernels_tc.cu after stack_fwd.cuh (uses its argument struct and helpers).
//
// STATUS: EXPERIMENT, opt-in with AST_B200_DUO=1, correct (bit-identical to stack_fwd.cuh) but NOT faster: at T=16384,
// B=2 the forward takes 0.709 ms against 0.690 ms for two single-pair launches.  Without a resident operand tile every
// K-slice is a 64 KB unit through a ring that holds two of them; a unit takes ~1.7-2.2k cycles from request to MMA start
// (scripts/experiments/tma_rate.cu: 16 KB tiles take ~500 cycles each and throughput only scales with the number in
// flight -- 32 / 60 / 74 / 87 B/clk/SM at 1 / 2 / 4 / 8 tiles), against 768 cycles of MMA work per unit.  Making several
// pairs per CTA pay needs a deeper ring than the 128 KB left next to the staging buffers.  Visiting the taps in a
// different order per CTA (L2 hot-spot theory) made it slower, not faster.
//
// Why: with one tile per SM (stack_fwd.cuh) the tensor pipe idles ~70 % of a layer -- during the epilogues and the
// hand-off to the neighbour tiles.  Two pairs are independent, so here the work items (layer l, pair 0), (layer l,
// pair 1), (layer l+1, pair 0), ... are pipelined through the same warps: while the epilogue warps finish item n
// (residual update, publish), the MMA warp already runs the dilated conv of item n+1, and by the time item n+2 (the
// same pair, next layer) starts, its neighbours' tiles were published a whole item ago -- no flag wait on the
// critical path.
//   TMEM (512 columns): h of pair 0 | h of pair 1 | acc1 | acc2 (the accumulators are shared: the items alternate).
//   smem: one staging region (relu(z) operand of the 1x1 GEMM, then the relu(h') hi/lo tile on its way to global
//   memory), the extract staging tile, and the 4 x 32 KB operand ring.  No operand tile is resident: the centre tap is
//   loaded like the side taps (the tile was published anyway).
//   Ring slots are handed out from a free list that the producer thread and the MMA thread both simulate (same
//   deterministic greedy schedule: fill the next unit when enough slots are free, else consume the oldest unit), so a
//   unit's two slots are requested together as soon as ANY two slots are free.
#pragma once

#define F2_RT 0                                   // [K-half][hi, lo] x [128][64] bf16: staging (see above)
#define F2_STG 65536
#define F2_RING 98304
#define F2B_FULL(i) (8u * (i))
#define F2B_EMPTY(i) (32u + 8u * (i))
#define F2B_ACC1 64u
#define F2B_ACC2 72u
#define F2B_RT(j) (80u + 8u * (j))
#define F2B_DONE 96u
#define F2B_ANTI(pa) (104u + 8u * (pa))            // one per pair: the producer may run a whole item ahead of the epilogue
#define F2B_TMEM 120u

// free list of ring slots, simulated identically by the producer and the MMA thread
struct RingSim {
    uint32_t q;        // queue of slot ids, 4 bits each, oldest in the low bits
    int n;
    __device__ __forceinline__ void init() { q = 0x3210u; n = 4; }
    __device__ __forceinline__ uint32_t pop() { const uint32_t s = q & 15u; q >>= 4; --n; return s; }
    __device__ __forceinline__ void push(uint32_t s) { q |= s << (4 * n); ++n; }
};
// units of one work item: up to 6 K-slices of the dilated conv (weights + operand: 2 slots) then the two K-halves of the
// 1x1 weights (1 slot each)
struct F2Units {
    int ntap; int tapk[3];
    __device__ __forceinline__ int count() const { return 2 * ntap + 2; }
    __device__ __forceinline__ int need(int u) const { return u < 2 * ntap ? 2 : 1; }
};
__device__ __forceinline__ F2Units f2_units(int t0, int d, int T) {
    F2Units u;
    u.ntap = 1; u.tapk[0] = 1;                                  // centre first: its rows never leave the clip
    if (tap_live(t0, 0, d, T)) u.tapk[u.ntap++] = 0;
    if (tap_live(t0, 2, d, T)) u.tapk[u.ntap++] = 2;
    return u;
}

__global__ void __launch_bounds__(FWD_THREADS, 1) tc_fwd_stack2_kernel(const __grid_constant__ FwdStackArgs p) {
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t base = (smem_u32(tc_smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = tc_smem_raw + (base - smem_u32(tc_smem_raw));
    const uint32_t s_rt = base + F2_RT, s_stg = base + F2_STG, s_ring = base + F2_RING, bar = base + FS_BAR;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tpp = p.tiles_per_pair;
    const int ti = (int)blockIdx.x % tpp, gi = (int)blockIdx.x / tpp;
    const int pairA = p.tile0 / tpp + 2 * gi;                    // this CTA's pairs: pairA, pairA + 1
    const int t0 = ti * 128;
    const int NL = p.NL;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.tmA_hi); tma_prefetch_desc(&p.tmA_lo); tma_prefetch_desc(&p.tmW1); tma_prefetch_desc(&p.tmW2);
        tma_prefetch_desc(&p.tmF);
        for (int i = 0; i < 4; ++i) { mbar_init(bar + F2B_FULL(i), 1); mbar_init(bar + F2B_EMPTY(i), 1); }
        mbar_init(bar + F2B_ACC1, 1); mbar_init(bar + F2B_ACC2, 1);
        mbar_init(bar + F2B_RT(0), 8); mbar_init(bar + F2B_RT(1), 8);
        mbar_init(bar + F2B_DONE, 16); mbar_init(bar + F2B_ANTI(0), 1); mbar_init(bar + F2B_ANTI(1), 1);
        mbar_init_fence();
    }
    if (warp == 1) tmem_alloc<512>(bar + F2B_TMEM);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gen + (bar + F2B_TMEM - base));
    const uint32_t acc1 = tmem, acc2 = tmem + 128;

    if (warp == 0) {
        // ================================ TMA producer ========================================
        if (lane == 0) {
            RingSim ring; ring.init();
            uint32_t fills[4] = {0, 0, 0, 0};
            auto arm = [&](uint32_t slot) {                  // wait until the slot's previous contents were consumed; arm for 32 KB
                if (fills[slot] > 0) mbar_wait(bar + F2B_EMPTY(slot), (fills[slot] - 1u) & 1u);
                ++fills[slot];
                mbar_expect_tx(bar + F2B_FULL(slot), 2 * TILE16);
            };
            for (int l = 0; l < NL; ++l) {
                const int d = 1 << (l % 10);
                const int w1row = l * 3 * 2 * AST_C, w2row = l * 2 * AST_C;
                const F2Units un = f2_units(t0, d, p.T);
                const int nu = un.count();
                for (int pa = 0; pa < 2; ++pa) {
                    const int b = pairA + pa;
                    bool polled = false;
                    TLD(50 + pa * 5);
                    uint32_t used[8][2];                     // slots of every unit (for the simulated frees)
                    int nf = 0, nc = 0;
                    while (nc < nu) {
                        if (nf < nu && ring.n >= un.need(nf)) {
                            const int u = nf++;
                            if (u < 2 * un.ntap) {
                                const int k = un.tapk[u >> 1], j = u & 1;
                                const uint32_t sw = ring.pop(), sa = ring.pop();
                                used[u][0] = sw; used[u][1] = sa;
                                arm(sw);
                                if (u == 0) TLD(51 + pa * 5);
                                tma_load_2d(s_ring + sw * FS_SLOT, &p.tmW1, bar + F2B_FULL(sw), j * 64, w1row + (k * 2 + 0) * 128);
                                tma_load_2d(s_ring + sw * FS_SLOT + TILE16, &p.tmW1, bar + F2B_FULL(sw), j * 64, w1row + (k * 2 + 1) * 128);
                                if (!polled) {               // relu(h_l) of this pair: own tile and the +-d neighbours are published
                                    stack_wait_rows2(p.flags, tpp, p.T, p.err, b, t0, t0, FS_PUB * (unsigned)(l + 1));
                                    stack_wait_rows2(p.flags, tpp, p.T, p.err, b, t0 - d, t0 + d, FS_PUB * (unsigned)(l + 1));
                                    fence_proxy_async_all();
                                    polled = true;
                                    TLD(52 + pa * 5);
                                }
                                arm(sa);
                                if (u == 0) TLD(53 + pa * 5);
                                const int t = t0 + (k - 1) * d;
                                tma_load_4d(s_ring + sa * FS_SLOT, &p.tmA_hi, bar + F2B_FULL(sa), j * 64, t, b, l & 1);
                                tma_load_4d(s_ring + sa * FS_SLOT + TILE16, &p.tmA_lo, bar + F2B_FULL(sa), j * 64, t, b, l & 1);
                            } else {
                                const int j = u - 2 * un.ntap;
                                const uint32_t sw = ring.pop();
                                used[u][0] = sw; used[u][1] = 0xFu;
                                arm(sw);
                                tma_load_2d(s_ring + sw * FS_SLOT, &p.tmW2, bar + F2B_FULL(sw), j * 64, w2row);
                                tma_load_2d(s_ring + sw * FS_SLOT + TILE16, &p.tmW2, bar + F2B_FULL(sw), j * 64, w2row + 128);
                            }
                        } else {                             // the MMA thread consumes the oldest unit: its slots return to the list
                            const int u = nc++;
                            ring.push(used[u][0]);
                            if (used[u][1] != 0xFu) ring.push(used[u][1]);
                        }
                    }
                    // anti-dependency of this item's publish (see stack_fwd.cuh): far taps of layer l-1
                    if (l >= 1 && l + 1 < NL && ((l - 1) % 10) >= 7) {
                        const int dp = 1 << ((l - 1) % 10);
                        stack_wait_rows2(p.flags, tpp, p.T, p.err, b, t0 - dp, t0 + dp, FS_PUB * (unsigned)(l + 1));
                        mbar_arrive(bar + F2B_ANTI(pa));
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ==========================================
        if (lane == 0) {
            RingSim ring; ring.init();
            uint32_t uses[4] = {0, 0, 0, 0};
            auto ready = [&](uint32_t slot) { mbar_wait(bar + F2B_FULL(slot), uses[slot] & 1u); ++uses[slot]; };
            auto mma_slice = [&](uint32_t acc, uint32_t a, uint32_t w, uint32_t accum0) {
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    const uint64_t a_hi = umma_desc_kmajor_sw128(a + kk * 32);
                    const uint64_t a_lo = umma_desc_kmajor_sw128(a + TILE16 + kk * 32);
                    const uint64_t w_hi = umma_desc_kmajor_sw128(w + kk * 32);
                    const uint64_t w_lo = umma_desc_kmajor_sw128(w + TILE16 + kk * 32);
                    umma_f16(acc, a_hi, w_hi, IDESC, (kk > 0) ? 1u : accum0);
                    umma_f16(acc, a_lo, w_hi, IDESC, 1u);
                    umma_f16(acc, a_hi, w_lo, IDESC, 1u);
                }
            };
            uint32_t n = 0;                                   // work item index = 2 l + pair
            for (int l = 0; l < NL; ++l) {
                const int d = 1 << (l % 10);
                const F2Units un = f2_units(t0, d, p.T);
                const int nu = un.count();
                for (int pa = 0; pa < 2; ++pa, ++n) {
                    uint32_t used[8][2];
                    int nf = 0, nc = 0;
                    while (nc < nu) {
                        if (nf < nu && ring.n >= un.need(nf)) {            // same schedule as the producer: learn the slot ids
                            const int u = nf++;
                            used[u][0] = ring.pop();
                            used[u][1] = u < 2 * un.ntap ? ring.pop() : 0xFu;
                        } else {
                            const int u = nc++;
                            const uint32_t sw = used[u][0], sa = used[u][1];
                            if (u < 2 * un.ntap) {
                                ready(sw); ready(sa);
                                tc_fence_after();
                                TLD(pa * 10 + u);
                                mma_slice(acc1, s_ring + sa * FS_SLOT, s_ring + sw * FS_SLOT, u > 0 ? 1u : 0u);
                                umma_commit(bar + F2B_EMPTY(sw));
                                umma_commit(bar + F2B_EMPTY(sa));
                                ring.push(sw); ring.push(sa);
                                if (u == 2 * un.ntap - 1) umma_commit(bar + F2B_ACC1);
                            } else {
                                const int j = u - 2 * un.ntap;
                                mbar_wait(bar + F2B_RT(j), n & 1u);          // relu(z) K-half j staged (also: acc1 has been read)
                                if (j == 0 && n > 0) mbar_wait(bar + F2B_DONE, (n - 1u) & 1u);   // acc2 of the previous item has been read
                                ready(sw);
                                tc_fence_after();
                                TLD(pa * 10 + u);
                                mma_slice(acc2, s_rt + j * 2 * TILE16, s_ring + sw * FS_SLOT, j > 0 ? 1u : 0u);
                                umma_commit(bar + F2B_EMPTY(sw));
                                ring.push(sw);
                                if (j == 1) umma_commit(bar + F2B_ACC2);
                            }
                        }
                    }
                }
            }
        }
        __syncwarp();
    } else {
        // ================================ epilogue warps ======================================
        const int q = warp & 3, cc = (warp - 2) >> 2, half = cc >> 1;
        const int r = q * 32 + lane;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        uint8_t* act = gen + F2_RT + half * 2 * TILE16;              // K-half `half`: hi, lo
        const bool storer = (q == 0) && ((cc & 1) == 0) && (lane == 0);
        uint32_t anti_n[2] = {0, 0};

        auto put16 = [&](uint32_t hacc, int sb, const float (&v)[16], uint32_t& word, bool more) {
            if (more) {
#pragma unroll
                for (int c = 0; c < 2; ++c) stage8_relu_hi_lo(act, act + TILE16, sw128_off(r, (cc & 1) * 4 + sb * 2 + c), &v[8 * c]);
            }
            tmem_st16(hacc + lane_off + cc * 32 + sb * 16, v);
#pragma unroll
            for (int e = 0; e < 16; ++e) or_bit_if_pos(word, v[e], 1u << (sb * 16 + e));
        };
        // h_i of pair pa is staged: publish the operand tile (each warp its own region), then the non-critical outputs
        auto publish = [&](int i, int pa, uint32_t word) {
            const int b = pairA + pa;
            const size_t row0 = (size_t)b * p.T + t0, row = row0 + r;
            const uint32_t hacc = tmem + 256 + pa * 128;
            unsigned* my_flag = p.flags + (size_t)b * tpp + ti;
            const bool more = i < NL;
            const int f = i >= 1 ? (int)p.f_slot[i - 1] : -1;
            const bool exp = p.export_h[i] != 0;
            if (more) {
                __syncwarp();
                if (i >= 2 && ((i - 2) % 10) >= 7) { mbar_wait(bar + F2B_ANTI(pa), anti_n[pa] & 1u); ++anti_n[pa]; }
                const int ch = (cc & 1) * 4 + (lane & 3);
                const size_t gbase = ((size_t)(i & 1) * p.R + row0) * AST_C + half * 64 + ch * 8;
#pragma unroll
                for (int it = 0; it < 4; ++it) {
                    const int rr = q * 32 + it * 8 + (lane >> 2);
                    const uint4 H = *reinterpret_cast<const uint4*>(act + sw128_off(rr, ch));
                    const uint4 L = *reinterpret_cast<const uint4*>(act + TILE16 + sw128_off(rr, ch));
                    *reinterpret_cast<uint4*>(p.a_hi + gbase + (size_t)rr * AST_C) = H;
                    *reinterpret_cast<uint4*>(p.a_lo + gbase + (size_t)rr * AST_C) = L;
                }
                __syncwarp();
                // the next GEMM1 of this pair polls this tile's own flag (centre tap), so "count >= 16 k" still implies that
                // all 16 warps finished publish k before any of them can reach publish k+1
                if (lane == 0) red_release_gpu_add(my_flag, 1u);
            }
            p.hmask[((size_t)i * p.R + row) * AST_MW + cc] = word;
            tmem_wait_st();
            if (f >= 0 || exp) {
#pragma unroll
                for (int sb = 0; sb < 2; ++sb) {
                    float h[16];
                    tmem_ld16_nowait(hacc + lane_off + cc * 32 + sb * 16, h);
                    tmem_wait_ld();
                    if (exp) {
                        float4* out = reinterpret_cast<float4*>(p.H + ((size_t)i * p.R + row) * AST_C + cc * 32 + sb * 16);
#pragma unroll
                        for (int e = 0; e < 4; ++e) out[e] = make_float4(h[4 * e], h[4 * e + 1], h[4 * e + 2], h[4 * e + 3]);
                    }
                    if (f >= 0) {
#pragma unroll
                        for (int c = 0; c < 2; ++c)
                            *reinterpret_cast<uint4*>(gen + F2_STG + half * TILE16 + sw128_off(r, (cc & 1) * 4 + sb * 2 + c)) =
                                make_uint4(pack_bf16(h[8 * c], h[8 * c + 1]), pack_bf16(h[8 * c + 2], h[8 * c + 3]),
                                           pack_bf16(h[8 * c + 4], h[8 * c + 5]), pack_bf16(h[8 * c + 6], h[8 * c + 7]));
                    }
                }
                if (f >= 0) {
                    fence_proxy_async_smem();
                    named_bar_sync(1 + half, 256);
                    if (storer) {
                        tma_store_4d(&p.tmF, s_stg + half * TILE16, half * 64, t0, b, f);
                        tma_store_commit();
                        tma_store_wait_read();
                    }
                    named_bar_sync(1 + half, 256);
                }
            }
        };

        // ---- pseudo layer: start conv of both pairs ----
        for (int pa = 0; pa < 2; ++pa) {
            const int t = t0 + r;
            const float* xb = p.x + (size_t)(pairA + pa) * p.T;
            const float xm = (t > 0) ? xb[t - 1] * (1.0f / 128.0f) : 0.f;
            const float x0 = xb[t] * (1.0f / 128.0f);
            const float xp = (t + 1 < p.T) ? xb[t + 1] * (1.0f / 128.0f) : 0.f;
            uint32_t word = 0;
#pragma unroll
            for (int sb = 0; sb < 2; ++sb) {
                float v[16];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int c4 = cc * 8 + sb * 4 + e;
                    const float4 w0 = __ldg(reinterpret_cast<const float4*>(p.W0) + c4);
                    const float4 w1 = __ldg(reinterpret_cast<const float4*>(p.W0 + AST_C) + c4);
                    const float4 w2 = __ldg(reinterpret_cast<const float4*>(p.W0 + 2 * AST_C) + c4);
                    const float4 bb = __ldg(reinterpret_cast<const float4*>(p.b0) + c4);
                    v[4 * e] = fmaf(w2.x, xp, fmaf(w1.x, x0, fmaf(w0.x, xm, bb.x)));
                    v[4 * e + 1] = fmaf(w2.y, xp, fmaf(w1.y, x0, fmaf(w0.y, xm, bb.y)));
                    v[4 * e + 2] = fmaf(w2.z, xp, fmaf(w1.z, x0, fmaf(w0.z, xm, bb.z)));
                    v[4 * e + 3] = fmaf(w2.w, xp, fmaf(w1.w, x0, fmaf(w0.w, xm, bb.w)));
                }
                put16(tmem + 256 + pa * 128, sb, v, word, true);
            }
            publish(0, pa, word);
        }
        uint32_t n = 0;
        for (int l = 0; l < NL; ++l) {
            for (int pa = 0; pa < 2; ++pa, ++n) {
                const uint32_t par = n & 1u;
                const size_t row = (size_t)(pairA + pa) * p.T + t0 + r;
                const uint32_t hacc = tmem + 256 + pa * 128;
                // ---- epilogue 1: z = acc1 + bd -> sign mask, relu(z) hi/lo -> A operand of the 1x1 GEMM ----
                mbar_wait(bar + F2B_ACC1, par);
                tc_fence_after();
                if (threadIdx.x == 128) TLD(30 + pa * 10);
                {
                    uint32_t word = 0;
#pragma unroll
                    for (int sb = 0; sb < 2; ++sb) {
                        float v[16];
                        tmem_ld16_nowait(acc1 + lane_off + cc * 32 + sb * 16, v);
                        float4 bb[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) bb[e] = __ldg(reinterpret_cast<const float4*>(p.bd + (size_t)l * AST_C) + cc * 8 + sb * 4 + e);
                        tmem_wait_ld();
#pragma unroll
                        for (int e = 0; e < 4; ++e) { v[4 * e] += bb[e].x; v[4 * e + 1] += bb[e].y; v[4 * e + 2] += bb[e].z; v[4 * e + 3] += bb[e].w; }
#pragma unroll
                        for (int e = 0; e < 16; ++e) or_bit_if_pos(word, v[e], 1u << (sb * 16 + e));
#pragma unroll
                        for (int c = 0; c < 2; ++c) stage8_relu_hi_lo(act, act + TILE16, sw128_off(r, (cc & 1) * 4 + sb * 2 + c), &v[8 * c]);
                    }
                    fence_proxy_async_smem();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar + F2B_RT(half));
                    p.zmask[((size_t)l * p.R + row) * AST_MW + cc] = word;
                }
                if (threadIdx.x == 128) TLD(31 + pa * 10);
                // ---- epilogue 2: h' = h + (acc2 + br) ----
                mbar_wait(bar + F2B_ACC2, par);
                tc_fence_after();
                if (threadIdx.x == 128) TLD(32 + pa * 10);
                {
                    uint32_t word = 0;
                    const bool more = l + 1 < NL;
#pragma unroll
                    for (int sb = 0; sb < 2; ++sb) {
                        float v[16], h[16];
                        tmem_ld16_nowait(acc2 + lane_off + cc * 32 + sb * 16, v);
                        tmem_ld16_nowait(hacc + lane_off + cc * 32 + sb * 16, h);
                        float4 bb[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) bb[e] = __ldg(reinterpret_cast<const float4*>(p.br + (size_t)l * AST_C) + cc * 8 + sb * 4 + e);
                        tmem_wait_ld();
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            v[4 * e] = h[4 * e] + (v[4 * e] + bb[e].x); v[4 * e + 1] = h[4 * e + 1] + (v[4 * e + 1] + bb[e].y);
                            v[4 * e + 2] = h[4 * e + 2] + (v[4 * e + 2] + bb[e].z); v[4 * e + 3] = h[4 * e + 3] + (v[4 * e + 3] + bb[e].w);
                        }
                        put16(hacc, sb, v, word, more);
                    }
                    tc_fence_before();                       // acc2 has been read: the next item's 1x1 GEMM may overwrite it
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar + F2B_DONE);
                    if (threadIdx.x == 128) TLD(33 + pa * 10);
                    publish(l + 1, pa, word);
                    if (threadIdx.x == 128) TLD(34 + pa * 10);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<512>(tmem);
}
