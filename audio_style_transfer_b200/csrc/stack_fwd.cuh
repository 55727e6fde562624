// stack_fwd.cuh -- the WHOLE encoder forward (start conv + every dilated layer, model.py:80-119 of the reference) in
// ONE cooperative launch.  Included by kernels_tc.cu (uses its helpers).  CTA i owns time tile i (128 rows) for all
// layers; used when a pair has at most one tile per SM (the headline shape: T = 16384, 128 tiles on 148 SMs).
//
// What stays on the SM
//   * the fp32 residual stream h: TMEM (128 columns), updated in place by the epilogue; only content / style
//     extracts are exported to global memory;
//   * relu(h) as a bf16 hi/lo operand WINDOW of 256 rows per K-half: rows [64, 192) are the CTA's own tile, rows
//     [64-d, 64) and [192, 192+d) the d boundary rows of the two neighbour tiles.  For d <= 64 ("near" layers, 21
//     of 30) all three taps of the dilated conv are UMMA descriptors into this window at row offsets 64-d, 64, 64+d
//     (K-major SWIZZLE_128B descriptors may start at any row: scripts/experiments/umma_row_offset.cu) -- no operand
//     tile is streamed, the ring carries weights only;
//   * relu(z) (A operand of the 1x1 conv) overlays the centre of the window: it is dead once the 1x1 GEMM is done.
// What crosses SMs
//   * near layers: a tile publishes ONLY its first and last d rows (32-byte stores straight from the epilogue
//     registers into the canonical [2][B][T][128] bf16 buffers + one release per boundary warp); the neighbour polls
//     the flag and pulls the rows into its window halo with two small TMA boxes per part.  Rows outside the pair are
//     zero-filled by TMA: that IS the reference's per-layer 'SAME' padding (masked.py:139);
//   * far layers (d >= 128): whole tiles are published and the +-d taps are streamed through the ring as before.
// Pipelining inside a tile (what hides the epilogues behind the MMAs): everything is split by K-half j (channels
// 64j .. 64j+63).  All 16 epilogue warps first finish K-half 0 of a phase, signal it, then do K-half 1:
//   epilogue 1 (relu(z) hi/lo)  : RT(0) lets the 1x1 GEMM start on K-half 0 while K-half 1 is still being split;
//   epilogue 2 (h', relu(h'))   : ACT(0) + publish(0) let the next layer's centre-tap MMAs on K-half 0 start and the
//                                 neighbours fetch K-half 0 of the halo while K-half 1 is still in the epilogue.
// Roles: warp 0 = weight / far-operand ring producer, warp 1 = MMA issuer, warps 2..17 = epilogue (TMEM lane quadrant
// q = warp % 4, 16-column group cc of each K-half), warp 18 = halo producer (flag polls + halo TMA, anti-dependency
// polls).  All CTAs are co-resident (cooperative launch, grid <= #SMs, 1 CTA/SM); every wait targets work of an
// earlier publish, so the dependency graph is acyclic.  Every spin is bounded: on expiry the CTA raises *err, sets its
// abort word and all further waits fall through, so the kernel always terminates and the host reports AST_ECUDA.
#pragma once

struct alignas(64) FwdStackArgs {
    CUtensorMap tmA_hi, tmA_lo;      // bf16 relu(h) hi / lo [2][B][T][128] ping-pong, box {64,128,1,1}: far taps
    CUtensorMap tmHa_hi[4], tmHa_lo[4];   // same buffers, box {64, 8<<i, 1, 1}: halo rows of the near layers
    CUtensorMap tmW1, tmW2;          // forward weight operand rows (see tc_weights_fwd_kernel), box {64,128}
    const float* x;                  // [B][T] waveform
    const float* W0; const float* b0;            // start conv [3][128], [128]
    const float* bd; const float* br;            // [NL][128] each
    float* H;                        // fp32 residual stream [NL+1][R][128]: only exported layers are written
    uint32_t* zmask; uint32_t* hmask;            // [NL][R][4], [NL+1][R][4]
    __nv_bfloat16* a_hi; __nv_bfloat16* a_lo;    // the buffers behind tmA_* (published with plain stores)
    __nv_bfloat16* F;                // bf16 extract copies [n_style][R][128]
    unsigned* flags;                 // [ntiles][2] (K-half), zeroed before the launch
    unsigned* err;
    int T, R, NL, tiles_per_pair;
    int tile0;                       // global index of this launch's first tile (launches cover whole pairs)
    unsigned short need[34];         // need[i]: flag value once publish i (relu(h_i), i = 0: start conv) is complete
    signed char f_slot[32];          // style slot that receives bf16(H[l+1]), or -1
    unsigned char export_h[33];      // export_h[i] != 0: H[i] is written to global memory (i = 0: start-conv output)
};

#define F2_HALO 64
#define F2_WBUF 32768                             // one (K-half, part) window buffer: [256 rows][128 B]
#define F2_CENTRE (F2_HALO * 128)                 // byte offset of the own tile inside a window buffer
#define F2_RING (4 * F2_WBUF)
#define F2_NSLOT 3
#define F2_SLOT 32768                             // hi [128][64] | lo [128][64]
#define F2_BAR (F2_RING + F2_NSLOT * F2_SLOT)
#define F2_SMEM (F2_BAR + 1024 + 1024)
#define F2_THREADS 608
#define F2B_FULL(i) (8u * (i))
#define F2B_EMPTY(i) (32u + 8u * (i))
#define F2B_ACT(j) (64u + 8u * (j))
#define F2B_HALO(j) (80u + 8u * (j))
#define F2B_ACC1 96u
#define F2B_RT(j) (104u + 8u * (j))
#define F2B_ACC2 120u
#define F2B_ANTI 128u
#define F2B_TMEM 136u
#define F2B_ABORT 144u
#define FS_SPIN_LIMIT (1u << 21)                  // flag polls (one L2 round trip each)
#define FS_WAIT_CYCLES 3000000000ll               // mbarrier waits: ~1.5 s

#ifdef AST_EXPERIMENTS
// detailed stamps of ONE layer (g_tl_layer), second block of the timeline buffer: [gridDim.x*64 + cta*64 + i]
__device__ int g_tl_layer = 12;
#define TLD(i) do { if (g_timeline && l == g_tl_layer) g_timeline[((size_t)gridDim.x + blockIdx.x) * 64 + (i)] = clock64(); } while (0)
#else
#define TLD(i) do { } while (0)
#endif

// dilation of layer l (model.py:98) and the publish geometry that follows from the CONSUMER layer's dilation
__device__ __host__ __forceinline__ int stack_dil(int l) { return 1 << (l % 10); }
__device__ __host__ __forceinline__ bool stack_near(int d) { return d <= F2_HALO; }
__device__ __host__ __forceinline__ int stack_npub(int d) { return d <= 32 ? 8 : 16; }   // warps that release per publish and K-half

// ---- bounded waits ------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool stack_aborted(const volatile unsigned* ab) { return *ab != 0u; }
__device__ __forceinline__ void stack_abort(volatile unsigned* ab, unsigned* err, unsigned code) {
    *ab = 1u;
    atomicExch(err, code);
}
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void stack_mbar_wait(uint32_t bar, uint32_t parity, volatile unsigned* ab, unsigned* err) {
    if (mbar_try(bar, parity)) return;
    const long long t0 = clock64();
    unsigned n = 0;
    while (!mbar_try(bar, parity)) {
        if ((++n & 63u) == 0u) {
            if (stack_aborted(ab)) return;
            if (clock64() - t0 > FS_WAIT_CYCLES) { stack_abort(ab, err, 2u); return; }
        }
    }
}
// Waits until flags f0 and f1 (may be the same word) have reached `need`.  Relaxed polls + one acquire fence: two
// acquire loads would cost two dependent L2 round trips.
__device__ __forceinline__ void stack_wait_flags(const unsigned* f0, const unsigned* f1, unsigned need, volatile unsigned* ab,
                                                 unsigned* err) {
    unsigned spins = 0;
    for (;;) {
        const unsigned v0 = ld_relaxed_gpu(f0), v1 = ld_relaxed_gpu(f1);
        if (v0 >= need && v1 >= need) break;
        if (++spins > FS_SPIN_LIMIT || stack_aborted(ab)) { stack_abort(ab, err, 1u); break; }
    }
    fence_acq_rel_gpu();
}
// Both K-half flags of every tile that holds rows [ta, ta+127] / [tb, tb+127] of pair b (far taps, anti-dependency).
__device__ __forceinline__ void stack_wait_rows2(const unsigned* flags, int tiles_per_pair, int T, int b, int ta, int tb,
                                                 unsigned need, volatile unsigned* ab, unsigned* err) {
    const bool la = !(ta + 127 < 0 || ta >= T), lb = !(tb + 127 < 0 || tb >= T);
    if (!la && !lb) return;
    if (!la) ta = tb;
    if (!lb) tb = ta;
    const unsigned* base = flags + (size_t)b * tiles_per_pair * 2;
    const unsigned* f[4] = {base + 2 * ((ta < 0 ? 0 : ta) >> 7), base + 2 * ((ta + 127 >= T ? T - 1 : ta + 127) >> 7),
                            base + 2 * ((tb < 0 ? 0 : tb) >> 7), base + 2 * ((tb + 127 >= T ? T - 1 : tb + 127) >> 7)};
    unsigned spins = 0;
    for (;;) {
        bool ok = true;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            unsigned vx, vy;
            asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];" : "=r"(vx), "=r"(vy) : "l"(f[i]) : "memory");
            ok = ok && vx >= need && vy >= need;
        }
        if (ok) break;
        if (++spins > FS_SPIN_LIMIT || stack_aborted(ab)) { stack_abort(ab, err, 1u); break; }
    }
    fence_acq_rel_gpu();
}

__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

__global__ void __launch_bounds__(F2_THREADS, 1) tc_fwd_stack_kernel(const __grid_constant__ FwdStackArgs p) {
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t base = (smem_u32(tc_smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = tc_smem_raw + (base - smem_u32(tc_smem_raw));
    const uint32_t s_ring = base + F2_RING, bar = base + F2_BAR;
    volatile unsigned* ab = reinterpret_cast<volatile unsigned*>(gen + F2_BAR + F2B_ABORT);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = p.tile0 + (int)blockIdx.x;
    const int row0 = tile * 128;
    const int b = row0 / p.T, t0 = row0 - b * p.T;
    const int NL = p.NL;
    if (threadIdx.x == 0) TL(0);

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.tmA_hi); tma_prefetch_desc(&p.tmA_lo); tma_prefetch_desc(&p.tmW1); tma_prefetch_desc(&p.tmW2);
        for (int i = 0; i < 4; ++i) { tma_prefetch_desc(&p.tmHa_hi[i]); tma_prefetch_desc(&p.tmHa_lo[i]); }
        for (int i = 0; i < F2_NSLOT; ++i) { mbar_init(bar + F2B_FULL(i), 1); mbar_init(bar + F2B_EMPTY(i), 1); }
        for (int j = 0; j < 2; ++j) { mbar_init(bar + F2B_ACT(j), 16); mbar_init(bar + F2B_HALO(j), 1); mbar_init(bar + F2B_RT(j), 16); }
        mbar_init(bar + F2B_ACC1, 1); mbar_init(bar + F2B_ACC2, 1); mbar_init(bar + F2B_ANTI, 1);
        *ab = 0u;
        mbar_init_fence();
    }
    if (warp == 1) tmem_alloc<512>(bar + F2B_TMEM);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gen + F2_BAR + F2B_TMEM);
    const uint32_t acc1 = tmem, acc2 = tmem + 128, hacc = tmem + 256;
    auto win = [&](int j, int part) -> uint32_t { return base + (uint32_t)(j * 2 + part) * F2_WBUF; };

    if (warp == 0) {
        // ================================ ring producer: weights (+ far-tap operand tiles) ========================
        if (lane == 0) {
            uint32_t nfill = 0;
            auto acquire = [&]() -> uint32_t {               // next ring slot, armed for 32 KB
                const uint32_t slot = nfill % F2_NSLOT;
                if (nfill >= F2_NSLOT) stack_mbar_wait(bar + F2B_EMPTY(slot), ((nfill / F2_NSLOT) - 1u) & 1u, ab, p.err);
                ++nfill;
                mbar_expect_tx(bar + F2B_FULL(slot), 2 * TILE16);
                return slot;
            };
            for (int l = 0; l < NL; ++l) {
                const int d = stack_dil(l);
                const int w1row = l * 3 * 2 * AST_C, w2row = l * 2 * AST_C;
                auto load_w1 = [&](int k, int j) {
                    const uint32_t slot = acquire();
                    const uint32_t st = s_ring + slot * F2_SLOT, fb = bar + F2B_FULL(slot);
                    tma_load_2d(st, &p.tmW1, fb, j * 64, w1row + (k * 2 + 0) * 128);
                    tma_load_2d(st + TILE16, &p.tmW1, fb, j * 64, w1row + (k * 2 + 1) * 128);
                };
                TLD(0);
                if (stack_near(d)) {
                    // fill order = the MMA issuer's take order: per K-half the centre, -d and +d tap weights
                    for (int j = 0; j < 2; ++j) { load_w1(1, j); load_w1(0, j); load_w1(2, j); }
                } else {
                    load_w1(1, 0);
                    load_w1(1, 1);
                    const bool live0 = tap_live(t0, 0, d, p.T), live2 = tap_live(t0, 2, d, p.T);
                    bool polled = false;
                    for (int k = 0; k < 3; k += 2) {
                        if (!(k == 0 ? live0 : live2)) continue;
                        const int t = t0 + (k - 1) * d;
                        for (int j = 0; j < 2; ++j) {
                            load_w1(k, j);
                            if (!polled) {                    // relu(h_l) of the tap rows is in global memory (both K-halves)
                                stack_wait_rows2(p.flags, p.tiles_per_pair, p.T, b, live0 ? t0 - d : t0 + d, live2 ? t0 + d : t0 - d,
                                                 p.need[l], ab, p.err);
                                fence_proxy_async_all();
                                polled = true;
                                TLD(2);
                            }
                            const uint32_t slot = acquire();
                            const uint32_t st = s_ring + slot * F2_SLOT, fb = bar + F2B_FULL(slot);
                            tma_load_4d(st, &p.tmA_hi, fb, j * 64, t, b, l & 1);
                            tma_load_4d(st + TILE16, &p.tmA_lo, fb, j * 64, t, b, l & 1);
                        }
                    }
                }
                for (int j = 0; j < 2; ++j) {                 // 1x1 weights, K-half j: hi rows then lo rows
                    const uint32_t slot = acquire();
                    const uint32_t st = s_ring + slot * F2_SLOT, fb = bar + F2B_FULL(slot);
                    tma_load_2d(st, &p.tmW2, fb, j * 64, w2row);
                    tma_load_2d(st + TILE16, &p.tmW2, fb, j * 64, w2row + 128);
                }
                TLD(6);
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ==========================================
        if (lane == 0) {
            uint32_t nuse = 0, n_near = 0;
            auto take = [&]() -> uint32_t {
                const uint32_t slot = nuse % F2_NSLOT;
                stack_mbar_wait(bar + F2B_FULL(slot), (nuse / F2_NSLOT) & 1u, ab, p.err);
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
            for (int l = 0; l < NL; ++l) {
                const uint32_t par = (uint32_t)l & 1u;
                const int d = stack_dil(l);
                if (stack_near(d)) {
                    for (int j = 0; j < 2; ++j) {
                        stack_mbar_wait(bar + F2B_ACT(j), par, ab, p.err);       // own relu(h_l), K-half j, staged
                        tc_fence_after();
                        TLD(8 + j);
                        uint32_t sw = take();
                        tc_fence_after();
                        mma_slice(acc1, win(j, 0) + F2_CENTRE, win(j, 1) + F2_CENTRE, s_ring + sw * F2_SLOT, j > 0 ? 1u : 0u);
                        umma_commit(bar + F2B_EMPTY(sw));
                        stack_mbar_wait(bar + F2B_HALO(j), n_near & 1u, ab, p.err);   // the neighbours' boundary rows landed
                        tc_fence_after();
                        TLD(10 + j);
#pragma unroll
                        for (int k = 0; k < 3; k += 2) {
                            sw = take();
                            tc_fence_after();
                            const uint32_t ro = (uint32_t)(F2_HALO + (k - 1) * d) * 128u;
                            mma_slice(acc1, win(j, 0) + ro, win(j, 1) + ro, s_ring + sw * F2_SLOT, 1u);
                            umma_commit(bar + F2B_EMPTY(sw));
                        }
                    }
                    ++n_near;
                } else {
                    for (int j = 0; j < 2; ++j) {
                        stack_mbar_wait(bar + F2B_ACT(j), par, ab, p.err);
                        tc_fence_after();
                        TLD(8 + j);
                        const uint32_t sw = take();
                        tc_fence_after();
                        mma_slice(acc1, win(j, 0) + F2_CENTRE, win(j, 1) + F2_CENTRE, s_ring + sw * F2_SLOT, j > 0 ? 1u : 0u);
                        umma_commit(bar + F2B_EMPTY(sw));
                    }
                    for (int k = 0; k < 3; k += 2) {
                        if (!tap_live(t0, k, d, p.T)) continue;
                        for (int j = 0; j < 2; ++j) {
                            const uint32_t sw = take();
                            const uint32_t sa = take();
                            tc_fence_after();
                            TLD(11 + k + j);
                            mma_slice(acc1, s_ring + sa * F2_SLOT, s_ring + sa * F2_SLOT + TILE16, s_ring + sw * F2_SLOT, 1u);
                            umma_commit(bar + F2B_EMPTY(sw));
                            umma_commit(bar + F2B_EMPTY(sa));
                        }
                    }
                }
                umma_commit(bar + F2B_ACC1);
                TLD(15);
                for (int j = 0; j < 2; ++j) {
                    stack_mbar_wait(bar + F2B_RT(j), par, ab, p.err);
                    const uint32_t sw = take();
                    tc_fence_after();
                    TLD(16 + j);
                    mma_slice(acc2, win(j, 0) + F2_CENTRE, win(j, 1) + F2_CENTRE, s_ring + sw * F2_SLOT, j > 0 ? 1u : 0u);
                    umma_commit(bar + F2B_EMPTY(sw));
                }
                umma_commit(bar + F2B_ACC2);
                TLD(18);
            }
        }
        __syncwarp();
    } else if (warp == 18) {
        // ================================ halo producer ==========================================
        if (lane == 0) {
            uint32_t n_near = 0;
            const unsigned* fl = p.flags + (size_t)b * p.tiles_per_pair * 2;
            const int ti = t0 >> 7;
            const bool has_up = t0 > 0, has_dn = t0 + 128 < p.T;
            for (int l = 0; l < NL; ++l) {
                const int d = stack_dil(l);
                if (stack_near(d)) {
                    // the halo rows of the previous near layer are dead once that layer's conv MMAs have completed
                    if (l > 0) stack_mbar_wait(bar + F2B_ACC1, (uint32_t)(l - 1) & 1u, ab, p.err);
                    const int bi = d <= 8 ? 0 : (d == 16 ? 1 : (d == 32 ? 2 : 3));
                    const int dbox = 8 << bi;
                    for (int j = 0; j < 2; ++j) {
                        const unsigned* f_up = fl + 2 * (has_up ? ti - 1 : (has_dn ? ti + 1 : ti)) + j;
                        const unsigned* f_dn = fl + 2 * (has_dn ? ti + 1 : (has_up ? ti - 1 : ti)) + j;
                        if (has_up || has_dn) {
                            stack_wait_flags(f_up, f_dn, p.need[l], ab, p.err);
                            fence_proxy_async_all();
                        }
                        TLD(20 + j);
                        const uint32_t hb = bar + F2B_HALO(j);
                        mbar_expect_tx(hb, 4u * (uint32_t)dbox * 128u);
                        const uint32_t top = (uint32_t)(F2_HALO - dbox) * 128u, bot = (uint32_t)(F2_HALO + 128) * 128u;
                        tma_load_4d(win(j, 0) + top, &p.tmHa_hi[bi], hb, j * 64, t0 - dbox, b, l & 1);
                        tma_load_4d(win(j, 1) + top, &p.tmHa_lo[bi], hb, j * 64, t0 - dbox, b, l & 1);
                        tma_load_4d(win(j, 0) + bot, &p.tmHa_hi[bi], hb, j * 64, t0 + 128, b, l & 1);
                        tma_load_4d(win(j, 1) + bot, &p.tmHa_lo[bi], hb, j * 64, t0 + 128, b, l & 1);
                    }
                    ++n_near;
                }
                // Anti-dependency of publish l+1 (relu(h_{l+1}) overwrites ping-pong buffer (l+1)&1 = relu(h_{l-1})): the
                // tiles whose +-d_{l-1} taps read this tile must be past layer l-1, i.e. have published relu(h_l).  For
                // d_{l-1} <= 128 those are the adjacent tiles, already waited for by this layer's taps; the far taps
                // (d_{l-1} >= 128, incl. the 512 -> 1 wrap) need this extra look, done here off the critical path.
                if (l >= 1 && l + 1 < NL && ((l - 1) % 10) >= 7) {
                    const int dp = stack_dil(l - 1);
                    stack_wait_rows2(p.flags, p.tiles_per_pair, p.T, b, t0 - dp, t0 + dp, p.need[l], ab, p.err);
                    mbar_arrive(bar + F2B_ANTI);
                }
            }
            (void)n_near;
        }
    } else {
        // ================================ epilogue warps ======================================
        const int q = warp & 3, cc = (warp - 2) >> 2;                // TMEM lane quadrant, 16-column group of a K-half
        const int r = q * 32 + lane;
        const size_t row = (size_t)row0 + r;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        const uint32_t sw0 = F2_CENTRE + sw128_off(r, 2 * cc), sw1 = F2_CENTRE + sw128_off(r, 2 * cc + 1);
        unsigned* my_flag = p.flags + (size_t)tile * 2;
        uint16_t* hmask16 = reinterpret_cast<uint16_t*>(p.hmask);
        uint16_t* zmask16 = reinterpret_cast<uint16_t*>(p.zmask);
        uint32_t anti_n = 0;

        // relu(v[0..15]) -> bf16 hi / lo of columns 64 j + 16 cc ...: two 16-byte chunks each into window (j, hi/lo)
        auto stage16 = [&](int j, const float (&v)[16], uint32_t (&Hh)[8], uint32_t (&Ll)[8]) {
#pragma unroll
            for (int e = 0; e < 8; ++e) split_bf16x2(fmaxf(v[2 * e], 0.f), fmaxf(v[2 * e + 1], 0.f), Hh[e], Ll[e]);
            uint8_t* whi = gen + (j * 2 + 0) * F2_WBUF;
            uint8_t* wlo = gen + (j * 2 + 1) * F2_WBUF;
            *reinterpret_cast<uint4*>(whi + sw0) = make_uint4(Hh[0], Hh[1], Hh[2], Hh[3]);
            *reinterpret_cast<uint4*>(whi + sw1) = make_uint4(Hh[4], Hh[5], Hh[6], Hh[7]);
            *reinterpret_cast<uint4*>(wlo + sw0) = make_uint4(Ll[0], Ll[1], Ll[2], Ll[3]);
            *reinterpret_cast<uint4*>(wlo + sw1) = make_uint4(Ll[4], Ll[5], Ll[6], Ll[7]);
        };
        // h_i, K-half j (columns 64 j + 16 cc .. + 15 of this thread's row) is in v: TMEM, own operand window, hand-off
        // to the neighbours, sign mask, exports.  Critical outputs first.
        auto emit = [&](int i, int j, const float (&v)[16]) {
            const int l = i - 1;                       // for TLD
            const bool more = i < NL;
            const int col = j * 64 + cc * 16;
            tmem_st16(hacc + lane_off + col, v);
            if (more) {
                uint32_t Hh[8], Ll[8];
                stage16(j, v, Hh, Ll);
                fence_proxy_async_smem();
                tc_fence_before();
                const int dn = stack_dil(i);           // dilation of the layer that consumes relu(h_i)
                const bool near = stack_near(dn);
                if (i >= 2 && j == 0 && ((i - 2) % 10) >= 7) { stack_mbar_wait(bar + F2B_ANTI, anti_n & 1u, ab, p.err); ++anti_n; }
                if (!near || r < dn || r >= 128 - dn) {
                    const size_t g = ((size_t)(i & 1) * p.R + row) * AST_C + col;
                    st_global_v8(p.a_hi + g, Hh[0], Hh[1], Hh[2], Hh[3], Hh[4], Hh[5], Hh[6], Hh[7]);
                    st_global_v8(p.a_lo + g, Ll[0], Ll[1], Ll[2], Ll[3], Ll[4], Ll[5], Ll[6], Ll[7]);
                }
                __syncwarp();
                // Flag counts are cumulative over publishes, so "count >= need[i]" must imply that every releasing warp
                // finished publish i: the ACT arrival (which gates the next layer's conv and with it every warp's publish
                // i+1) comes AFTER this warp's release.
                if (lane == 0) {
#ifdef AST_EXPERIMENTS
                    const bool mute = g_dbg_skip_publish == tile + 1 && i >= 3;     // fault injection (tests/test_gpu_loud.py)
#else
                    const bool mute = false;
#endif
                    if ((dn > 32 || q == 0 || q == 3) && !mute) red_release_gpu_add(my_flag + j, 1u);
                    mbar_arrive(bar + F2B_ACT(j));
                }
                if (threadIdx.x == 64) TLD(36 + j);
            }
            uint32_t word = 0;
#pragma unroll
            for (int e = 0; e < 16; ++e) or_bit_if_pos(word, v[e], 1u << e);
            hmask16[((size_t)i * p.R + row) * 8 + j * 4 + cc] = (uint16_t)word;
            const int f = i >= 1 ? (int)p.f_slot[i - 1] : -1;
            if (p.export_h[i]) {
                float4* out = reinterpret_cast<float4*>(p.H + ((size_t)i * p.R + row) * AST_C + col);
#pragma unroll
                for (int e = 0; e < 4; ++e) out[e] = make_float4(v[4 * e], v[4 * e + 1], v[4 * e + 2], v[4 * e + 3]);
            }
            if (f >= 0)                                // bf16 extract copy: 16 columns = 32 bytes = one full sector per thread
                st_global_v8(p.F + ((size_t)f * p.R + row) * AST_C + col,
                             pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]),
                             pack_bf16(v[8], v[9]), pack_bf16(v[10], v[11]), pack_bf16(v[12], v[13]), pack_bf16(v[14], v[15]));
        };

        // ---- pseudo layer: start conv (model.py:80-93), same arithmetic as startconv_fwd_kernel ----
        {
            const int t = t0 + r;
            const float* xb = p.x + (size_t)b * p.T;
            const float xm = (t > 0) ? xb[t - 1] * (1.0f / 128.0f) : 0.f;
            const float x0 = xb[t] * (1.0f / 128.0f);
            const float xp = (t + 1 < p.T) ? xb[t + 1] * (1.0f / 128.0f) : 0.f;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                float v[16];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int c4 = j * 16 + cc * 4 + e;
                    const float4 w0 = __ldg(reinterpret_cast<const float4*>(p.W0) + c4);
                    const float4 w1 = __ldg(reinterpret_cast<const float4*>(p.W0 + AST_C) + c4);
                    const float4 w2 = __ldg(reinterpret_cast<const float4*>(p.W0 + 2 * AST_C) + c4);
                    const float4 bb = __ldg(reinterpret_cast<const float4*>(p.b0) + c4);
                    v[4 * e] = fmaf(w2.x, xp, fmaf(w1.x, x0, fmaf(w0.x, xm, bb.x)));
                    v[4 * e + 1] = fmaf(w2.y, xp, fmaf(w1.y, x0, fmaf(w0.y, xm, bb.y)));
                    v[4 * e + 2] = fmaf(w2.z, xp, fmaf(w1.z, x0, fmaf(w0.z, xm, bb.z)));
                    v[4 * e + 3] = fmaf(w2.w, xp, fmaf(w1.w, x0, fmaf(w0.w, xm, bb.w)));
                }
                emit(0, j, v);
            }
            tmem_wait_st();
        }
        for (int l = 0; l < NL; ++l) {
            const uint32_t par = (uint32_t)l & 1u;
            // ---- epilogue 1: z = acc1 + bd -> sign mask, relu(z) hi/lo -> A operand of the 1x1 GEMM (window centre) ----
            stack_mbar_wait(bar + F2B_ACC1, par, ab, p.err);
            tc_fence_after();
            if (threadIdx.x == 64 && l < 30) TL(2 + l);
            if (threadIdx.x == 64) TLD(30);
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int col = j * 64 + cc * 16;
                float v[16];
                tmem_ld16_nowait(acc1 + lane_off + col, v);
                float4 bb[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) bb[e] = __ldg(reinterpret_cast<const float4*>(p.bd + (size_t)l * AST_C + col) + e);
                tmem_wait_ld();
#pragma unroll
                for (int e = 0; e < 4; ++e) { v[4 * e] += bb[e].x; v[4 * e + 1] += bb[e].y; v[4 * e + 2] += bb[e].z; v[4 * e + 3] += bb[e].w; }
                uint32_t Hh[8], Ll[8];
                stage16(j, v, Hh, Ll);
                fence_proxy_async_smem();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar + F2B_RT(j));
                uint32_t word = 0;
#pragma unroll
                for (int e = 0; e < 16; ++e) or_bit_if_pos(word, v[e], 1u << e);
                zmask16[((size_t)l * p.R + row) * 8 + j * 4 + cc] = (uint16_t)word;
            }
            if (threadIdx.x == 64) TLD(32);
            // ---- epilogue 2: h' = h + (acc2 + br)  (same association as the reference, model.py:109-114) ----
            stack_mbar_wait(bar + F2B_ACC2, par, ab, p.err);
            tc_fence_after();
            if (threadIdx.x == 64) TLD(33);
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int col = j * 64 + cc * 16;
                float v[16], h[16];
                tmem_ld16_nowait(acc2 + lane_off + col, v);
                tmem_ld16_nowait(hacc + lane_off + col, h);
                float4 bb[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) bb[e] = __ldg(reinterpret_cast<const float4*>(p.br + (size_t)l * AST_C + col) + e);
                tmem_wait_ld();
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    v[4 * e] = h[4 * e] + (v[4 * e] + bb[e].x); v[4 * e + 1] = h[4 * e + 1] + (v[4 * e + 1] + bb[e].y);
                    v[4 * e + 2] = h[4 * e + 2] + (v[4 * e + 2] + bb[e].z); v[4 * e + 3] = h[4 * e + 3] + (v[4 * e + 3] + bb[e].w);
                }
                emit(l + 1, j, v);
            }
            tmem_wait_st();
            if (threadIdx.x == 64) TLD(41);
            if (threadIdx.x == 64 && l < 30) TL(32 + l);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) TL(1);
    if (warp == 1) tmem_dealloc<512>(tmem);
}
