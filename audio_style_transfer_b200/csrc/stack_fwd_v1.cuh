// stack_fwd.cuh -- the WHOLE encoder forward (start conv + every dilated layer) in ONE cooperative launch.
// Included by kernels_tc.cu (uses its helpers).  Used when the clip has at most one 128-row tile per SM, which
// is the headline shape (T = 16384, one pair: 128 tiles on 148 SMs).
//
// Why: with one tile per SM the per-layer kernels cannot overlap anything -- every layer pays a launch
// boundary (~5k cycles to first data), a full store of the tile and a full reload.  Here CTA i owns time tile i
// for all layers:
//   * the fp32 residual stream h never leaves the SM: it lives in TMEM (128 columns) and is updated in place
//     by the epilogue (tcgen05.ld / tcgen05.st); only content/style extracts are exported to global memory;
//   * the CTA's own relu(h) hi/lo operand tile stays in shared memory: the centre tap of the next layer's
//     dilated conv needs no load at all and its MMAs start while the neighbours' tiles are still in flight;
//   * only the +-d taps come from other CTAs: each CTA publishes its bf16 operand tile by TMA store (issued
//     before the rest of the epilogue) and, once the bulk group has completed, bumps a per-tile flag (release);
//     consumers poll the flags of the (at most two) neighbour tiles they need (acquire) and TMA-load the
//     shifted tiles -- no grid-wide barrier, so a slow tile only delays the tiles that depend on it.
//     (Measured alternative: publishing with plain 16-byte global stores from the epilogue threads + one
//     red.release per warp -- 7k cycles from the stores to the flag, worse than the 4k of the TMA store.)
//   * weights stream through a 4 x 32 KB ring that runs ahead across layer boundaries.
// All CTAs are co-resident (cooperative launch, grid <= #SMs, 1 CTA/SM), every wait targets work of an earlier
// layer, so the dependency graph is acyclic; spins are bounded and raise *err instead of hanging.
#pragma once

struct alignas(64) FwdStackArgs {
    CUtensorMap tmA_hi, tmA_lo;      // bf16 relu(h) hi / lo [2][B][T][128] ping-pong, box {64,128,1,1}
    CUtensorMap tmW1, tmW2;          // forward weight operand rows (see tc_weights_fwd_kernel), box {64,128}
    CUtensorMap tmF;                 // bf16 extract copies [n_style][B][T][128]
    const float* x;                  // [B][T] waveform
    const float* W0; const float* b0;            // start conv [3][128], [128]
    const float* bd; const float* br;            // [NL][128] each
    float* H;                        // fp32 residual stream [NL+1][R][128]: only exported layers are written
    uint32_t* zmask; uint32_t* hmask;            // [NL][R][4], [NL+1][R][4]
    __nv_bfloat16* a_hi; __nv_bfloat16* a_lo;    // the buffers behind tmA_hi / tmA_lo (published with plain stores)
    __nv_bfloat16* F;                // the buffer behind tmF: bf16 extract copies [n_style][R][128]
    unsigned* flags;                 // [ntiles], zeroed before the launch; += 1 per epilogue warp per published tile
    unsigned* err;
    int T, R, NL, tiles_per_pair;
    int tile0;                       // global index of this launch's first tile (launches cover whole pairs)
    int nslot;                       // ring slots: 4, or 5 when no extract copy is staged (the 32 KB staging tile joins the ring)
    signed char f_slot[32];          // style slot that receives bf16(H[l+1]), or -1
    unsigned char export_h[33];      // export_h[i] != 0: H[i] is written to global memory (i = 0: start-conv output)
};

#define FS_ACT 0                                  // [K-half][hi, lo] x [128][64] bf16: own operand tile / relu(z)
#define FS_STG 65536                              // [K-half] x [128][64] bf16: staging of the extract copy
#define FS_RING 98304                             // 4 slots x 32 KB
#define FS_SLOT 32768
#define FS_BAR (FS_RING + 4 * FS_SLOT)
#define FS_SMEM (FS_BAR + 1024 + 1024)
#define FSB_FULL(i) (8u * (i))                    // up to 8 ring slots
#define FSB_EMPTY(i) (64u + 8u * (i))
#define FSB_ACT 128u
#define FSB_ACC1 136u
#define FSB_RT(j) (144u + 8u * (j))
#define FSB_ACC2 160u
#define FSB_TMEM 168u
#define FSB_ANTI 176u
#define FS_PUB 16u                                // flag increments per published operand tile (one per epilogue warp)
#define FS_SPIN_LIMIT (1u << 22)
// detailed stamps of ONE layer (g_tl_layer), second half of the timeline buffer: [gridDim.x*64 + cta*64 + i]
__device__ int g_tl_layer = 12;
#define TLD(i) do { if (g_timeline && l == g_tl_layer) g_timeline[((size_t)gridDim.x + blockIdx.x) * 64 + (i)] = clock64(); } while (0)

// Waits until every tile that holds rows [t, t+127] of pair b (clipped to the pair; one or two tiles) has published
// `need` increments.  Both flags are polled in the same loop so that their L2 round trips overlap.
__device__ __forceinline__ void stack_wait_rows(const unsigned* flags, int tiles_per_pair, int T, unsigned* err, int b, int t,
                                                unsigned need) {
    if (t + 127 < 0 || t >= T) return;                       // no such rows in this pair: nothing to wait for
    const int lo = (t < 0 ? 0 : t) >> 7;
    const int hi = ((t + 127 >= T ? T - 1 : t + 127)) >> 7;
    const unsigned* f0 = flags + (size_t)b * tiles_per_pair + lo;
    const unsigned* f1 = flags + (size_t)b * tiles_per_pair + hi;
    unsigned spins = 0;
    for (;;) {
        const unsigned v0 = ld_relaxed_gpu(f0), v1 = ld_relaxed_gpu(f1);
        if (v0 >= need && v1 >= need) break;
        if (++spins > FS_SPIN_LIMIT) { atomicExch(err, 1u); break; }
    }
    fence_acq_rel_gpu();
}
// Same for the rows of two taps (ta, tb) at once: up to four flags polled per round trip.
__device__ __forceinline__ void stack_wait_rows2(const unsigned* flags, int tiles_per_pair, int T, unsigned* err, int b, int ta,
                                                 int tb, unsigned need) {
    const bool la = !(ta + 127 < 0 || ta >= T), lb = !(tb + 127 < 0 || tb >= T);
    if (!la && !lb) return;
    if (!la) ta = tb;
    if (!lb) tb = ta;
    const unsigned* base = flags + (size_t)b * tiles_per_pair;
    const unsigned* f0 = base + ((ta < 0 ? 0 : ta) >> 7);
    const unsigned* f1 = base + ((ta + 127 >= T ? T - 1 : ta + 127) >> 7);
    const unsigned* f2 = base + ((tb < 0 ? 0 : tb) >> 7);
    const unsigned* f3 = base + ((tb + 127 >= T ? T - 1 : tb + 127) >> 7);
    unsigned spins = 0;
    for (;;) {
        const unsigned v0 = ld_relaxed_gpu(f0), v1 = ld_relaxed_gpu(f1), v2 = ld_relaxed_gpu(f2), v3 = ld_relaxed_gpu(f3);
        if (v0 >= need && v1 >= need && v2 >= need && v3 >= need) break;
        if (++spins > FS_SPIN_LIMIT) { atomicExch(err, 1u); break; }
    }
    fence_acq_rel_gpu();
}

__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
// relu(v) -> bf16 hi / lo: 4 x 16-byte chunks each into the swizzled smem tile pair (own centre-tap operand) AND
// into the global operand buffers (64 contiguous bytes per thread and part: two full sectors).  v is not modified.
__device__ __forceinline__ void stage_publish_relu_hi_lo(uint8_t* tile_hi, uint8_t* tile_lo, __nv_bfloat16* ghi,
                                                         __nv_bfloat16* glo, int r, int chunk0, const float (&v)[32]) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t off = sw128_off(r, chunk0 + i);
        float a[8], hi[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) { a[e] = fmaxf(v[8 * i + e], 0.f); hi[e] = bf16_round(a[e]); }
        const uint4 H = make_uint4(pack_bf16(hi[0], hi[1]), pack_bf16(hi[2], hi[3]), pack_bf16(hi[4], hi[5]), pack_bf16(hi[6], hi[7]));
        const uint4 L = make_uint4(pack_bf16(a[0] - hi[0], a[1] - hi[1]), pack_bf16(a[2] - hi[2], a[3] - hi[3]),
                                   pack_bf16(a[4] - hi[4], a[5] - hi[5]), pack_bf16(a[6] - hi[6], a[7] - hi[7]));
        *reinterpret_cast<uint4*>(tile_hi + off) = H;
        *reinterpret_cast<uint4*>(tile_lo + off) = L;
        if (ghi) { reinterpret_cast<uint4*>(ghi)[i] = H; reinterpret_cast<uint4*>(glo)[i] = L; }
    }
}

// relu(v[0..7]) -> bf16 hi / lo, one 16-byte chunk each at byte offset `off` of the swizzled tile pair
__device__ __forceinline__ void stage8_relu_hi_lo(uint8_t* tile_hi, uint8_t* tile_lo, uint32_t off, const float* v) {
    uint32_t P[4], Q[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) split_bf16x2(fmaxf(v[2 * e], 0.f), fmaxf(v[2 * e + 1], 0.f), P[e], Q[e]);
    *reinterpret_cast<uint4*>(tile_hi + off) = make_uint4(P[0], P[1], P[2], P[3]);
    *reinterpret_cast<uint4*>(tile_lo + off) = make_uint4(Q[0], Q[1], Q[2], Q[3]);
}

__global__ void __launch_bounds__(FWD_THREADS, 1) tc_fwd_stack_kernel(const __grid_constant__ FwdStackArgs p) {
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t base = (smem_u32(tc_smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = tc_smem_raw + (base - smem_u32(tc_smem_raw));
    const uint32_t nslot = (uint32_t)p.nslot;
    const uint32_t s_act = base + FS_ACT, s_stg = base + FS_STG, s_ring = base + (nslot == 5 ? FS_STG : FS_RING), bar = base + FS_BAR;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = p.tile0 + (int)blockIdx.x;
    const int row0 = tile * 128;
    const int b = row0 / p.T, t0 = row0 - b * p.T;
    if (threadIdx.x == 0) TL(0);

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.tmA_hi); tma_prefetch_desc(&p.tmA_lo); tma_prefetch_desc(&p.tmW1); tma_prefetch_desc(&p.tmW2);
        tma_prefetch_desc(&p.tmF);
        for (int i = 0; i < 5; ++i) { mbar_init(bar + FSB_FULL(i), 1); mbar_init(bar + FSB_EMPTY(i), 1); }
        mbar_init(bar + FSB_ACT, 16); mbar_init(bar + FSB_ACC1, 1);
        mbar_init(bar + FSB_RT(0), 8); mbar_init(bar + FSB_RT(1), 8); mbar_init(bar + FSB_ACC2, 1);
        mbar_init(bar + FSB_ANTI, 1);
        mbar_init_fence();
    }
    if (warp == 1) tmem_alloc<512>(bar + FSB_TMEM);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gen + (bar + FSB_TMEM - base));
    const uint32_t acc1 = tmem, acc2 = tmem + 128, hacc = tmem + 256;

    if (warp == 0) {
        // ================================ TMA producer ========================================
        if (lane == 0) {
            uint32_t nfill = 0;
            auto acquire = [&]() -> uint32_t {               // next ring slot, armed for 32 KB
                const uint32_t slot = nfill % nslot;
                if (nfill >= nslot) mbar_wait(bar + FSB_EMPTY(slot), ((nfill / nslot) - 1u) & 1u);
                ++nfill;
                mbar_expect_tx(bar + FSB_FULL(slot), 2 * TILE16);
                return slot;
            };
            for (int l = 0; l < p.NL; ++l) {
                const int d = 1 << (l % 10);
                const int w1row = l * 3 * 2 * AST_C, w2row = l * 2 * AST_C;
                auto load_w1 = [&](int k, int j) {
                    const uint32_t slot = acquire();
                    const uint32_t st = s_ring + slot * FS_SLOT, fb = bar + FSB_FULL(slot);
                    tma_load_2d(st, &p.tmW1, fb, j * 64, w1row + (k * 2 + 0) * 128);
                    tma_load_2d(st + TILE16, &p.tmW1, fb, j * 64, w1row + (k * 2 + 1) * 128);
                };
                // Fill order (the MMA issuer takes slots in the same order): centre-tap weights; then per live side tap its two
                // weight slots followed by its two operand slots.  Weights never wait for a neighbour, so they are requested
                // before the flags are polled; all flags this layer needs are polled in ONE loop before the first operand tile.
                TLD(0);
                load_w1(1, 0);                                // centre tap: operand tile is resident, weights only
                load_w1(1, 1);
                TLD(1);
                const bool live0 = tap_live(t0, 0, d, p.T), live2 = tap_live(t0, 2, d, p.T);
                bool polled = false;
                for (int k = 0; k < 3; k += 2) {
                    if (!(k == 0 ? live0 : live2)) continue;
                    const int t = t0 + (k - 1) * d;
                    auto load_a = [&](int j) {
                        const uint32_t slot = acquire();
                        const uint32_t st = s_ring + slot * FS_SLOT, fb = bar + FSB_FULL(slot);
                        tma_load_4d(st, &p.tmA_hi, fb, j * 64, t, b, l & 1);
                        tma_load_4d(st + TILE16, &p.tmA_lo, fb, j * 64, t, b, l & 1);
                    };
                    if (polled && nslot == 5) {
                        // second tap with the 5-slot ring: weight and operand slot of a K-half back to back, so both are
                        // requested as soon as the first tap's K-half 0 has freed its two slots (round-robin slot order)
                        load_w1(k, 0); load_a(0);
                        load_w1(k, 1); load_a(1);
                        TLD(3 + k);
                        continue;
                    }
                    load_w1(k, 0);
                    load_w1(k, 1);
                    if (!polled) {                            // relu(h_l) of the tap rows is in global memory
                        stack_wait_rows2(p.flags, p.tiles_per_pair, p.T, p.err, b, live0 ? t0 - d : t0 + d, live2 ? t0 + d : t0 - d,
                                         FS_PUB * (unsigned)(l + 1));
                        fence_proxy_async_all();
                        polled = true;
                        TLD(2);
                    }
                    load_a(0);
                    load_a(1);
                    TLD(3 + k);
                }
                for (int j = 0; j < 2; ++j) {                 // 1x1 weights, K-half j: hi rows then lo rows
                    const uint32_t slot = acquire();
                    const uint32_t st = s_ring + slot * FS_SLOT, fb = bar + FSB_FULL(slot);
                    tma_load_2d(st, &p.tmW2, fb, j * 64, w2row);
                    tma_load_2d(st + TILE16, &p.tmW2, fb, j * 64, w2row + 128);
                }
                TLD(6);
                // Anti-dependency of the NEXT publish (relu(h_{l+1}) overwrites ping-pong buffer (l+1)&1 = relu(h_{l-1})):
                // the tiles whose +-d_{l-1} taps read this tile must be past layer l-1, i.e. have published relu(h_l).
                // For d_{l-1} < 128 those are the adjacent tiles, already waited for above (d_l taps); only the
                // far taps (d_{l-1} >= 128, incl. the 512 -> 1 wrap) need this extra look, done here off the critical path.
                if (l >= 1 && l + 1 < p.NL && ((l - 1) % 10) >= 7) {
                    const int dp = 1 << ((l - 1) % 10);
                    stack_wait_rows2(p.flags, p.tiles_per_pair, p.T, p.err, b, t0 - dp, t0 + dp, FS_PUB * (unsigned)(l + 1));
                    mbar_arrive(bar + FSB_ANTI);
                }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ==========================================
        if (lane == 0) {
            uint32_t nuse = 0;
            auto take = [&]() -> uint32_t {
                const uint32_t slot = nuse % nslot;
                mbar_wait(bar + FSB_FULL(slot), (nuse / nslot) & 1u);
                ++nuse;
                return slot;
            };
            // 12 MMAs of one K-half: (a_hi + a_lo) x (w_hi + w_lo) without the lo x lo term
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
            for (int l = 0; l < p.NL; ++l) {
                const uint32_t par = (uint32_t)l & 1u;
                const int d = 1 << (l % 10);
                mbar_wait(bar + FSB_ACT, par);                // own relu(h_l) hi/lo staged; h_l stored in TMEM
                tc_fence_after();
                TLD(8);
                for (int j = 0; j < 2; ++j) {
                    const uint32_t sw = take();
                    tc_fence_after();
                    TLD(9 + j);
                    mma_slice(acc1, s_act + j * 2 * TILE16, s_ring + sw * FS_SLOT, j > 0 ? 1u : 0u);
                    umma_commit(bar + FSB_EMPTY(sw));
                }
                bool first = true;
                for (int k = 0; k < 3; k += 2) {
                    if (!tap_live(t0, k, d, p.T)) continue;
                    const bool inter = !first && nslot == 5;              // producer fill order: W W A A, or W A W A (see there)
                    uint32_t sw[2];
                    if (!inter) { sw[0] = take(); sw[1] = take(); }
                    for (int j = 0; j < 2; ++j) {
                        if (inter) sw[j] = take();
                        const uint32_t sa = take();
                        tc_fence_after();
                        TLD(11 + k + j);
                        mma_slice(acc1, s_ring + sa * FS_SLOT, s_ring + sw[j] * FS_SLOT, 1u);
                        umma_commit(bar + FSB_EMPTY(sa));
                        umma_commit(bar + FSB_EMPTY(sw[j]));
                    }
                    first = false;
                }
                umma_commit(bar + FSB_ACC1);
                TLD(15);
                for (int j = 0; j < 2; ++j) {
                    mbar_wait(bar + FSB_RT(j), par);
                    const uint32_t sw = take();
                    tc_fence_after();
                    TLD(16 + j);
                    mma_slice(acc2, s_act + j * 2 * TILE16, s_ring + sw * FS_SLOT, j > 0 ? 1u : 0u);
                    umma_commit(bar + FSB_EMPTY(sw));
                }
                umma_commit(bar + FSB_ACC2);
                TLD(18);
            }
        }
        __syncwarp();
    } else {
        // ================================ epilogue warps ======================================
        const int q = warp & 3, cc = (warp - 2) >> 2, half = cc >> 1;
        const int r = q * 32 + lane;
        const size_t row = (size_t)row0 + r;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        uint8_t* act = gen + FS_ACT + half * 2 * TILE16;             // K-half `half`: hi, lo
        const bool storer = (q == 0) && ((cc & 1) == 0) && (lane == 0);
        unsigned* my_flag = p.flags + tile;
        uint32_t anti_n = 0;

        // The epilogue works in 16-column sub-blocks: with 576 threads the register cap is 96, and two live 32-wide
        // arrays spill to local memory, which the 227 KB smem carve-out leaves almost no L1 for.
        // put16: 16 values of h_i (columns cc*32 + sb*16 ...): relu hi/lo -> own operand tile, h_i -> TMEM, sign bits.
        auto put16 = [&](int sb, const float (&v)[16], uint32_t& word, bool more) {
            if (more) {
#pragma unroll
                for (int c = 0; c < 2; ++c) stage8_relu_hi_lo(act, act + TILE16, sw128_off(r, (cc & 1) * 4 + sb * 2 + c), &v[8 * c]);
            }
            tmem_st16(hacc + lane_off + cc * 32 + sb * 16, v);
#pragma unroll
            for (int e = 0; e < 16; ++e) or_bit_if_pos(word, v[e], 1u << (sb * 16 + e));
        };
        // publish: h_i is staged.  First what the next layer waits for (own centre-tap operand ready, operand tile
        // published for the neighbours), then the non-critical outputs re-read from TMEM: bf16 extract copy, fp32 export.
        auto publish = [&](int i, uint32_t word) {
            const int l = i - 1;                       // for TLD
            const bool more = i < p.NL;
            const int f = i >= 1 ? (int)p.f_slot[i - 1] : -1;
            const bool exp = p.export_h[i] != 0;
            if (more) {
                fence_proxy_async_smem();
                tc_fence_before();
                __syncwarp();
                if (threadIdx.x == 128) TLD(36);
                // Each warp publishes the region it staged itself (its 32 rows x 64 bytes of the hi and of the lo tile): no
                // cross-warp barrier.  A warp instruction moves 8 rows x 64 B = 16 full sectors; one release per warp.
                if (i >= 2 && ((i - 2) % 10) >= 7) { mbar_wait(bar + FSB_ANTI, anti_n & 1u); ++anti_n; }   // see producer
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
                // Flag counts are cumulative over publishes, so "count >= 16 k" must imply that ALL 16 warps finished publish
                // k: the ACT arrival (which gates the next layer's GEMM1 and with it every warp's publish k+1) comes AFTER this
                // warp's release.  The own-centre-tap MMAs it delays have several thousand cycles of slack.
                if (lane == 0) { red_release_gpu_add(my_flag, 1u); mbar_arrive(bar + FSB_ACT); }
                if (threadIdx.x == 128) TLD(37);
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
                    if (f >= 0 && p.nslot == 5) {
                        // bf16 extract copy straight from registers: 16 columns = 32 bytes = one full sector per thread
                        st_global_v8(p.F + ((size_t)f * p.R + row) * AST_C + cc * 32 + sb * 16,
                                     pack_bf16(h[0], h[1]), pack_bf16(h[2], h[3]), pack_bf16(h[4], h[5]), pack_bf16(h[6], h[7]),
                                     pack_bf16(h[8], h[9]), pack_bf16(h[10], h[11]), pack_bf16(h[12], h[13]), pack_bf16(h[14], h[15]));
                    } else if (f >= 0) {
#pragma unroll
                        for (int c = 0; c < 2; ++c)
                            *reinterpret_cast<uint4*>(gen + FS_STG + half * TILE16 + sw128_off(r, (cc & 1) * 4 + sb * 2 + c)) =
                                make_uint4(pack_bf16(h[8 * c], h[8 * c + 1]), pack_bf16(h[8 * c + 2], h[8 * c + 3]),
                                           pack_bf16(h[8 * c + 4], h[8 * c + 5]), pack_bf16(h[8 * c + 6], h[8 * c + 7]));
                    }
                }
                if (f >= 0 && p.nslot != 5) {          // 4-slot ring: staged tile -> TMA store, one thread per half
                    fence_proxy_async_smem();
                    named_bar_sync(1 + half, 256);
                    if (storer) {
                        tma_store_4d(&p.tmF, s_stg + half * TILE16, half * 64, t0, b, f);
                        tma_store_commit();
                        tma_store_wait_read();
                    }
                }
            }
            if (f >= 0 && p.nslot != 5) named_bar_sync(1 + half, 256);             // STG of this half may be overwritten again (ACT: every
                                                                   // warp only ever touches the region it staged itself)
            if (threadIdx.x == 128) TLD(41);
        };

        // ---- pseudo layer: start conv (model.py:80-93), same arithmetic as startconv_fwd_kernel ----
        {
            const int t = t0 + r;
            const float* xb = p.x + (size_t)b * p.T;
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
                put16(sb, v, word, true);
            }
            publish(0, word);
        }
        for (int l = 0; l < p.NL; ++l) {
            const uint32_t par = (uint32_t)l & 1u;
            // ---- epilogue 1: z = acc1 + bd -> sign mask, relu(z) hi/lo -> A operand of the 1x1 GEMM ----
            mbar_wait(bar + FSB_ACC1, par);
            tc_fence_after();
            if (threadIdx.x == 64 && l < 30) TL(2 + l);
            if (threadIdx.x == 128) TLD(30);
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
                if (lane == 0) mbar_arrive(bar + FSB_RT(half));
                p.zmask[((size_t)l * p.R + row) * AST_MW + cc] = word;
            }
            if (threadIdx.x == 128) TLD(32);
            // ---- epilogue 2: h' = h + (acc2 + br)  (same association as the reference, model.py:109-114) ----
            mbar_wait(bar + FSB_ACC2, par);
            tc_fence_after();
            if (threadIdx.x == 128) TLD(33);
            {
                uint32_t word = 0;
                const bool more = l + 1 < p.NL;
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
                    put16(sb, v, word, more);
                }
                if (threadIdx.x == 128) TLD(35);
                publish(l + 1, word);
            }
            if (threadIdx.x == 64 && l < 30) TL(32 + l);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) TL(1);
    if (warp == 1) tmem_dealloc<512>(tmem);
}
