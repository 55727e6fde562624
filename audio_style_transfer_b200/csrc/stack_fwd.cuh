// stack_fwd.cuh -- the WHOLE encoder forward (start conv + every dilated layer, model.py:80-119 of the reference) in
// ONE cooperative launch.  Included by kernels_tc.cu (uses its helpers).  CTA i owns time tile i (128 rows) for all
// layers; used when a pair has at most one tile per SM (the headline shape: T = 16384, 128 tiles on 148 SMs).
//
// What stays on the SM
//   * the fp32 residual stream h: TMEM (128 columns), updated in place by the epilogue; only content / style
//     extracts are exported to global memory;
//   * relu(h) as a bf16 hi/lo operand WINDOW of 192 rows per K-half: rows [32, 160) are the CTA's own tile, rows
//     [32-d, 32) and [160, 160+d) the d boundary rows of the two neighbour tiles.  For d <= 32 ("near" layers, 18
//     of 30) all three taps of the dilated conv are UMMA descriptors into this window at row offsets 32-d, 32, 32+d
//     (K-major SWIZZLE_128B descriptors may start at any row: scripts/experiments/umma_row_offset.cu) -- no operand
//     tile is streamed, the ring carries weights only;
//   * relu(z) (A operand of the 1x1 conv) overlays the centre of the window: it is dead once the 1x1 GEMM is done.
// What crosses SMs
//   * near layers: a tile hands over ONLY its first and last d rows, in "LL" form: every 8-byte unit is {4 bytes of
//     payload, 4-byte tag} (tag = launch epoch | publish index), written with 16-byte volatile stores straight from the
//     epilogue registers.  No fence, no flag: the consumer polls the data itself and a unit is valid when its tag
//     matches (an aligned 8-byte store is single-copy atomic), so the hand-off costs one store -> L2 -> load trip
//     instead of store, membar, red, poll, acquire fence and a TMA load (measured ~5k cycles that way).  The 16
//     epilogue warps import the rows themselves: one 64-byte LL item -> two swizzled 16-byte chunks of the window
//     halo.  Halo rows outside the pair stay zero from the start of the kernel: that IS the reference's per-layer
//     'SAME' padding (masked.py:139);
//   * far layers (d >= 64): whole tiles are published to the canonical [2][B][T][128] bf16 buffers (full 128-byte lines
//     copied from the staged smem tile, ONE gpu-scope release per K-half after a CTA barrier); consumers poll the
//     per-tile flags and stream the shifted tiles through the ring by TMA (rows outside the pair zero-filled by TMA).
// Pipelining inside a tile (what hides the epilogues behind the MMAs): everything is split by K-half j (channels
// 64j .. 64j+63).  All 16 epilogue warps first finish K-half 0 of a phase, signal it, then do K-half 1 (a named barrier
// per row quadrant keeps the hardware's highest-warp-first arbitration from running one warp through both halves):
//   epilogue 1 (relu(z) hi/lo)  : RT(0) lets the 1x1 GEMM start on K-half 0 while K-half 1 is still being split;
//   epilogue 2 (h', relu(h'))   : ACT(0) + hand-off(0) let the next layer's centre-tap MMAs on K-half 0 start and the
//                                 neighbours fetch K-half 0 of the halo while K-half 1 is still in the epilogue.
// Roles: warp 0 = weight / far-operand ring producer (+ anti-dependency polls), warp 1 = MMA issuer, warps 2..17 =
// epilogue (TMEM lane quadrant q = warp % 4, 16-column group cc of each K-half) and halo import.  All CTAs are
// co-resident (cooperative launch, grid <= #SMs, 1 CTA/SM); every wait targets work of an earlier publish, so the
// dependency graph is acyclic.  Every spin is bounded: on expiry the CTA raises *err, sets its abort word and all
// further waits fall through, so the kernel always terminates and the host reports AST_ECUDA (ast_check).
//
// Segments and groups (clips with more tiles than SMs, time shards with a per-stack halo exchange).  One launch runs the
// layers [l_begin, l_end) on a GROUP of at most #SMs consecutive tiles of one pair (or on whole pairs).  l_begin > 0
// resumes from the fp32 stream H[l_begin] in global memory (written by the previous segment's launches, or received from
// the neighbour rank); l_end < NL leaves H[l_end] there.  A group edge that is not a clip end behaves like one (zero
// halo, stale far-tap rows): the error this makes travels sum(d) <= 1023 rows per 10-layer segment into the group, so
// the host gives every group ceil(sum(d) / 128) halo tiles per inner side and only tiles in [wr_t0, wr_t1) write masks,
// exports and extract copies (tc_plan_groups in kernels_tc.cu).
#pragma once
#include <cuda_fp16.h>

struct StackPeer {
    const float* src[2];             // the neighbour's buffer this segment resumes from (fp32 [rows][128]), or NULL
    int t0[2], t1[2];                // tiles with t0 in [t0[s], t1[s]) take their rows from side s
    int row_off[2];                  // neighbour row = in-window t + row_off[s]
    const unsigned* sig_wait[2];     // this rank's word the neighbour signals; wait until *sig_wait[s] >= *step * 8 + wait_seq
    unsigned* sig_out[2];            // the neighbours' words to signal (*step * 8 + signal_seq) when this launch (signal_seq != 0) has completed
    const unsigned* step;            // device word: evaluations of this sweep direction started so far (bumped by a one-thread kernel in
                                     // front of the sweep's first launch, so a captured graph counts on as it is replayed)
    unsigned wait_seq, signal_seq;   // position of the segment in its sweep (wait) / + 1 (signal)
};
// sys-scope flag and data accesses of the peer-mapped halo path
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ float4 ld_relaxed_sys_f4(const float* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
#define FS_PEER_WAIT_NS 30000000000ull            // a neighbour rank may lag by seconds (first step, host jitter): 30 s
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// Which neighbour (if any) serves the rows of the tile at in-window time t0; waits for its signal.
__device__ __forceinline__ int stack_peer_side(const StackPeer& pr, int t0, volatile unsigned* ab, unsigned* err);

struct alignas(64) FwdStackArgs {
    CUtensorMap tmA_hi, tmA_lo;      // bf16 relu(h) hi / lo [2][B][T][128] ping-pong, box {64,128,1,1}: far taps
    CUtensorMap tmW1, tmW2;          // forward weight operand rows (see tc_weights_fwd_kernel), box {64,128}
    const float* x;                  // [B][T] waveform
    const float* W0; const float* b0;            // start conv [3][128], [128]
    const float* bd; const float* br;            // [NL][128] each
    float* H;                        // fp32 residual stream [NL+1][R][128]: only exported layers are written
    uint32_t* zmask; uint32_t* hmask;            // [NL][R][4], [NL+1][R][4]
    __nv_bfloat16* a_hi; __nv_bfloat16* a_lo;    // the buffers behind tmA_* (published with plain stores)
    __nv_bfloat16* F;                // bf16 extract copies [n_style][R][128]
    __nv_bfloat16* F_lo;             // accurate tier: bf16(h - float(bf16(h))), same layout, or NULL
    __half* Ht;                      // transposed fp16 extract copies [R/32][n_tslots][128][32 rows] (channel-wise head, gram_ours_tma.cu) or NULL
    int n_tslots;
    uint4* ll;                       // near hand-off lines [2][ntiles][2 sides][2 K-halves][4 cc][2 parts][4][32 rows] x 16 B
    unsigned* flags;                 // [ntiles][2] (K-half), zeroed before the launch
    unsigned* epoch;                 // [0] launch epoch (tags), [1] CTAs done: the last CTA of a launch bumps the epoch
    unsigned* err;
    int T, R, NL, tiles_per_pair, ntiles;
    int tile0;                       // global index of this launch's first tile
    int l_begin, l_end;              // layers of this launch; l_begin == 0 starts from the waveform (start conv)
    int grp_t0, grp_t1;              // in-pair time range covered by this launch's tiles (hand-off edges)
    int wr_t0, wr_t1;                // tiles with t0 in this range write masks / exports / extract copies
    // Time shards with peer-mapped halos (ast_p2p_attach): a resumed launch reads the rows of its window that only the
    // neighbour rank computes exactly straight from the neighbour's buffer over NVLink, after the neighbour's previous
    // segment has signalled completion into this rank's memory; the last CTA of a segment's last launch signals both
    // neighbours.  side 0 = the rank that owns the samples before this window's, 1 = after.
    StackPeer peer;
    unsigned short need[34];         // need[i]: flag value once publish i (relu(h_i), i = 0: start conv) is complete
    signed char f_slot[32];          // style slot that receives bf16(H[l+1]), or -1
    signed char t_slot[32];          // style slot whose transposed copy Ht[slot] receives H[l+1], or -1
    unsigned char export_h[33];      // export_h[i] != 0: H[i] is written to global memory (i = 0: start-conv output)
};

#define F2_HALO 32
#define F2_WROWS (128 + 2 * F2_HALO)
#define F2_WBUF (F2_WROWS * 128)                  // one (K-half, part) window buffer: [192 rows][128 B]
#define F2_CENTRE (F2_HALO * 128)                 // byte offset of the own tile inside a window buffer
#define F2_RING (4 * F2_WBUF)
#define F2_NSLOT 4
#define F2_SLOT 32768                             // hi [128][64] | lo [128][64]
#define F2_BAR (F2_RING + F2_NSLOT * F2_SLOT)
#define F2_SMEM (F2_BAR + 1024 + 1024)
#define F2_THREADS 576
#define LL_ROWS 32                                // row slots per (tile, side, K-half, cc, part, line) of the near hand-off
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
__device__ __host__ __forceinline__ int stack_npub(int d) { return stack_near(d) ? 8 : 1; }   // flag increments per publish and K-half

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
// Only tiles of this launch's group [g0, g1) are waited for (rows outside it are not produced by this launch).
// jsel: 0 / 1 = only that K-half's flag, 2 = both.
__device__ __forceinline__ void stack_wait_rows2(const unsigned* flags, int tiles_per_pair, int g0, int g1, int b, int ta, int tb,
                                                 unsigned need, volatile unsigned* ab, unsigned* err, int jsel = 2) {
    const bool la = !(ta + 127 < g0 || ta >= g1), lb = !(tb + 127 < g0 || tb >= g1);
    if (!la && !lb) return;
    if (!la) ta = tb;
    if (!lb) tb = ta;
    const unsigned* base = flags + (size_t)b * tiles_per_pair * 2;
    const unsigned* f[4] = {base + 2 * ((ta < g0 ? g0 : ta) >> 7), base + 2 * ((ta + 127 >= g1 ? g1 - 1 : ta + 127) >> 7),
                            base + 2 * ((tb < g0 ? g0 : tb) >> 7), base + 2 * ((tb + 127 >= g1 ? g1 - 1 : tb + 127) >> 7)};
    unsigned spins = 0;
    for (;;) {
        bool ok = true;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            unsigned vx, vy;
            asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];" : "=r"(vx), "=r"(vy) : "l"(f[i]) : "memory");
            ok = ok && (jsel == 1 || vx >= need) && (jsel == 0 || vy >= need);
        }
        if (ok) break;
        if (++spins > FS_SPIN_LIMIT || stack_aborted(ab)) { stack_abort(ab, err, 1u); break; }
    }
    fence_acq_rel_gpu();
}

// a far tap is live when its 128 source rows touch the group (rows outside the clip are zero-filled by TMA, rows of the
// clip outside the group are stale: both only reach rows the host discards, or are the reference's padding)
__device__ __forceinline__ bool stack_tap_live(int t0, int k, int shift, int g0, int g1) {
    const int t = t0 + (k - 1) * shift;
    return t + 128 > g0 && t < g1;
}

__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

// ---- LL lines (near hand-off) ------------------------------------------------------------------------------------
__device__ __forceinline__ void ll_store(uint4* p, uint32_t d0, uint32_t d1, uint32_t tag) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(d0), "r"(tag), "r"(d1), "r"(tag) : "memory");
}
__device__ __forceinline__ uint4 ll_load(const uint4* p) {
    uint4 v;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_relaxed_gpu_add(unsigned* p, unsigned v) {
    asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// One 64-byte LL item = 4 lines, LL_ROWS lines apart (the row slot is the fastest index so that the lanes of a warp --
// consecutive rows -- store and load contiguous bytes).  ll_fetch issues the four loads; ll_ready tells whether all 8 tags
// match and extracts the 8 payload words; ll_wait_item re-polls until they do (false after an abort).
__device__ __forceinline__ void ll_fetch(const uint4* src, uint4 (&v)[4]) {
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = ll_load(src + i * LL_ROWS);
}
__device__ __forceinline__ bool ll_ready(const uint4 (&v)[4], uint32_t tag, uint32_t (&w)[8]) {
    bool ok = true;
#pragma unroll
    for (int i = 0; i < 4; ++i) { ok = ok && v[i].y == tag && v[i].w == tag; w[2 * i] = v[i].x; w[2 * i + 1] = v[i].z; }
    return ok;
}
__device__ __forceinline__ bool ll_wait_item(const uint4* src, uint4 (&v)[4], uint32_t tag, uint32_t (&w)[8], volatile unsigned* ab,
                                             unsigned* err) {
    unsigned spins = 0;
    while (!ll_ready(v, tag, w)) {
        if (++spins > FS_SPIN_LIMIT || stack_aborted(ab)) { stack_abort(ab, err, 3u); return false; }
        ll_fetch(src, v);
    }
    return true;
}

__device__ __forceinline__ int stack_peer_side(const StackPeer& pr, int t0, volatile unsigned* ab, unsigned* err) {
    int side = -1;
    if (pr.src[0] && t0 >= pr.t0[0] && t0 < pr.t1[0]) side = 0;
    else if (pr.src[1] && t0 >= pr.t0[1] && t0 < pr.t1[1]) side = 1;
    if (side < 0) return -1;
    const unsigned want = *pr.step * 8u + pr.wait_seq;
    if (ld_acquire_sys(pr.sig_wait[side]) >= want) return side;
    const unsigned long long t_start = global_timer_ns();
    while (ld_acquire_sys(pr.sig_wait[side]) < want) {
        if (stack_aborted(ab)) return side;
        if (global_timer_ns() - t_start > FS_PEER_WAIT_NS) { stack_abort(ab, err, 4u); return side; }
        __nanosleep(200);
    }
    return side;
}
// the last CTA of a launch that closes a segment tells both neighbours that this rank's rows are in place
__device__ __forceinline__ void stack_peer_signal(const StackPeer& pr) {
    if (!pr.signal_seq) return;
    __threadfence_system();
#pragma unroll
    for (int sd = 0; sd < 2; ++sd)
        if (pr.sig_out[sd]) st_release_sys(pr.sig_out[sd], *pr.step * 8u + pr.signal_seq);
}

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
    const int NL = p.NL, LB = p.l_begin, LE = p.l_end;
    const int g0 = p.grp_t0, g1 = p.grp_t1;
    const uint32_t epoch = *reinterpret_cast<const volatile unsigned*>(p.epoch);      // stable: bumped by the previous launch's last CTA
    if (threadIdx.x == 0) TL(0);

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.tmA_hi); tma_prefetch_desc(&p.tmA_lo); tma_prefetch_desc(&p.tmW1); tma_prefetch_desc(&p.tmW2);
        for (int i = 0; i < F2_NSLOT; ++i) { mbar_init(bar + F2B_FULL(i), 1); mbar_init(bar + F2B_EMPTY(i), 1); }
        for (int j = 0; j < 2; ++j) { mbar_init(bar + F2B_ACT(j), 16); mbar_init(bar + F2B_HALO(j), 16); mbar_init(bar + F2B_RT(j), 16); }
        mbar_init(bar + F2B_ACC1, 1); mbar_init(bar + F2B_ACC2, 1); mbar_init(bar + F2B_ANTI, 1);
        *ab = 0u;
        mbar_init_fence();
    }
    if (warp == 1) tmem_alloc<512>(bar + F2B_TMEM);
    if (warp >= 2) {
        // halo rows of the operand window start as zeros and stay zero where the pair ends (the 'SAME' padding)
        const int e = (warp - 2) * 32 + lane;                        // 0..511
        for (int i = e; i < 4 * 2 * F2_HALO * 8; i += 512) {         // (buffer, halo row, 16-byte chunk)
            const int buf = i / (2 * F2_HALO * 8), rem = i % (2 * F2_HALO * 8), hr = rem >> 3, ch = rem & 7;
            const int wr = hr < F2_HALO ? hr : 128 + hr;
            *reinterpret_cast<uint4*>(gen + buf * F2_WBUF + wr * 128 + ch * 16) = make_uint4(0u, 0u, 0u, 0u);
        }
        fence_proxy_async_smem();
    }
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
            for (int l = LB; l < LE; ++l) {
                const int d = stack_dil(l);
                const int w1row = l * 3 * 2 * AST_C, w2row = l * 2 * AST_C;
                auto load_w1 = [&](int k, int j) {
                    const uint32_t slot = acquire();
                    const uint32_t st = s_ring + slot * F2_SLOT, fb = bar + F2B_FULL(slot);
                    tma_load_2d(st, &p.tmW1, fb, j * 64, w1row + (k * 2 + 0) * 128);
                    tma_load_2d(st + TILE16, &p.tmW1, fb, j * 64, w1row + (k * 2 + 1) * 128);
                };
                auto load_a = [&](int k, int j) {
                    const uint32_t slot = acquire();
                    const uint32_t st = s_ring + slot * F2_SLOT, fb = bar + F2B_FULL(slot);
                    tma_load_4d(st, &p.tmA_hi, fb, j * 64, t0 + (k - 1) * d, b, l & 1);
                    tma_load_4d(st + TILE16, &p.tmA_lo, fb, j * 64, t0 + (k - 1) * d, b, l & 1);
                };
                TLD(0);
                if (stack_near(d)) {
                    // fill order = the MMA issuer's take order: both centre-tap K-halves first (they need no neighbour),
                    // then per K-half the -d and +d tap weights
                    load_w1(1, 0); load_w1(1, 1);
                    for (int j = 0; j < 2; ++j) { load_w1(0, j); load_w1(2, j); }
                } else {
                    // centre weights, then K-half by K-half: the live side taps' weights (they wait for no neighbour), a poll of
                    // THAT K-half's flags of the tiles the taps read, the shifted operand tiles.  The neighbours publish K-half 0
                    // ~4.5k cycles before K-half 1: its taps' MMAs run while K-half 1 is still on its way.
                    load_w1(1, 0);
                    load_w1(1, 1);
                    const bool live0 = stack_tap_live(t0, 0, d, g0, g1), live2 = stack_tap_live(t0, 2, d, g0, g1);
                    if (live0 || live2) {
                        for (int j = 0; j < 2; ++j) {
                            if (live0) load_w1(0, j);
                            if (live2) load_w1(2, j);
                            stack_wait_rows2(p.flags, p.tiles_per_pair, g0, g1, b, live0 ? t0 - d : t0 + d, live2 ? t0 + d : t0 - d,
                                             p.need[l], ab, p.err, j);
                            fence_proxy_async_all();
                            TLD(2 + j);
                            if (live0) load_a(0, j);
                            if (live2) load_a(2, j);
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
                // Anti-dependency of a FAR publish l+1 (whole tile -> ping-pong buffer (l+1)&1, which holds relu(h_{l-1}) if
                // publish l-1 was far too): the tiles whose +-d_{l-1} taps read this tile must be past layer l-1's conv, i.e.
                // have made publish l.  Every publish bumps the flags (near ones without release semantics, for this purpose
                // only).  Polled here, off the critical path.
                if (l - 1 >= LB && l + 1 < LE && !stack_near(stack_dil(l + 1)) && !stack_near(stack_dil(l - 1))) {
                    const int dp = stack_dil(l - 1);
                    stack_wait_rows2(p.flags, p.tiles_per_pair, g0, g1, b, t0 - dp, t0 + dp, p.need[l], ab, p.err);
                    mbar_arrive(bar + F2B_ANTI);
                }
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
            for (int l = LB; l < LE; ++l) {
                const uint32_t par = (uint32_t)(l - LB) & 1u;
                const int d = stack_dil(l);
                // centre tap, both K-halves: the own tile is resident, nothing to wait for but its staging
                for (int j = 0; j < 2; ++j) {
                    stack_mbar_wait(bar + F2B_ACT(j), par, ab, p.err);           // own relu(h_l), K-half j, staged
                    tc_fence_after();
                    TLD(8 + j);
                    const uint32_t sw = take();
                    tc_fence_after();
                    mma_slice(acc1, win(j, 0) + F2_CENTRE, win(j, 1) + F2_CENTRE, s_ring + sw * F2_SLOT, j > 0 ? 1u : 0u);
                    umma_commit(bar + F2B_EMPTY(sw));
                }
                if (stack_near(d)) {
                    for (int j = 0; j < 2; ++j) {
                        stack_mbar_wait(bar + F2B_HALO(j), n_near & 1u, ab, p.err);   // the neighbours' boundary rows are in the window
                        tc_fence_after();
                        TLD(10 + j);
#pragma unroll
                        for (int k = 0; k < 3; k += 2) {
                            const uint32_t sw = take();
                            tc_fence_after();
                            const uint32_t ro = (uint32_t)(F2_HALO + (k - 1) * d) * 128u;
                            mma_slice(acc1, win(j, 0) + ro, win(j, 1) + ro, s_ring + sw * F2_SLOT, 1u);
                            umma_commit(bar + F2B_EMPTY(sw));
                        }
                    }
                    ++n_near;
                } else {
                    const bool live0 = stack_tap_live(t0, 0, d, g0, g1), live2 = stack_tap_live(t0, 2, d, g0, g1);
                    if (live0 || live2) {                                   // producer fill order per K-half: W (W) A (A)
                        const int ntap = (int)live0 + (int)live2;
                        for (int j = 0; j < 2; ++j) {
                            uint32_t sw[2];
                            for (int k = 0; k < ntap; ++k) sw[k] = take();
                            for (int k = 0; k < ntap; ++k) {
                                const uint32_t sa = take();
                                tc_fence_after();
                                TLD(11 + 2 * k + j);
                                mma_slice(acc1, s_ring + sa * F2_SLOT, s_ring + sa * F2_SLOT + TILE16, s_ring + sw[k] * F2_SLOT, 1u);
                                umma_commit(bar + F2B_EMPTY(sa));
                                umma_commit(bar + F2B_EMPTY(sw[k]));
                            }
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
        const bool has_up = t0 > g0, has_dn = t0 + 128 < g1;
        const bool wr = t0 >= p.wr_t0 && t0 < p.wr_t1;               // this tile's rows are exact: it writes the global outputs
        uint32_t anti_n = 0;
        // LL item of (ping-pong pp, tile tl, side sd, K-half j, cc, part, row slot): lines [4][LL_ROWS] with the slot fastest
        auto ll_item = [&](int pp, int tl, int sd, int j, int slot, int c, int part) -> uint4* {
            return p.ll + ((((((size_t)pp * p.ntiles + tl) * 2 + sd) * 2 + j) * 8 + c * 2 + part) * 4) * LL_ROWS + slot;
        };

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
        // to the neighbours.  Nothing else: masks and exports follow in finish().
        auto emit = [&](int i, int j, const float (&v)[16]) {
            const int l = i - 1;                       // for TLD
            const int col = j * 64 + cc * 16;
            tmem_st16(hacc + lane_off + col, v);
            if (i >= LE) return;                       // h_{l_end} is not consumed by this launch (LE == NL: by no layer at all)
            uint32_t Hh[8], Ll[8];
            stage16(j, v, Hh, Ll);
            fence_proxy_async_smem();
            tc_fence_before();
            const int dn = stack_dil(i);               // dilation of the layer that consumes relu(h_i)
#ifdef AST_EXPERIMENTS
            const bool mute = g_dbg_skip_publish == tile + 1 && i >= 3;     // fault injection (tests/test_gpu_loud.py)
#else
            const bool mute = false;
#endif
            if (stack_near(dn)) {
                const uint32_t tag = (epoch << 6) | (uint32_t)(i + 1);
                if ((r < dn || r >= 128 - dn) && !mute) {
                    const int sd = r < dn ? 0 : 1, slot = r < dn ? r : r - (128 - dn);
                    uint4* d0 = ll_item(i & 1, tile, sd, j, slot, cc, 0);
#pragma unroll
                    for (int e = 0; e < 4; ++e) ll_store(d0 + e * LL_ROWS, Hh[2 * e], Hh[2 * e + 1], tag);
#pragma unroll
                    for (int e = 0; e < 4; ++e) ll_store(d0 + (4 + e) * LL_ROWS, Ll[2 * e], Ll[2 * e + 1], tag);
                }
                __syncwarp();
                if (lane == 0) {
                    if (q == 0 || q == 3) red_relaxed_gpu_add(my_flag + j, 1u);      // bookkeeping for the anti-dependency polls only
                    mbar_arrive(bar + F2B_ACT(j));
                }
            } else {
                // far publish: the four warps of a row quadrant copy full 128-byte lines of the staged K-half; after a CTA
                // barrier ONE thread releases the flag (one gpu-scope membar per K-half instead of sixteen)
                if (j == 0 && i - 2 >= LB && !stack_near(stack_dil(i - 2))) { stack_mbar_wait(bar + F2B_ANTI, anti_n & 1u, ab, p.err); ++anti_n; }
                named_bar_sync(1 + q, 128);
                const uint8_t* whi = gen + (j * 2 + 0) * F2_WBUF + F2_CENTRE;
                const uint8_t* wlo = gen + (j * 2 + 1) * F2_WBUF + F2_CENTRE;
                const size_t g0 = ((size_t)(i & 1) * p.R + row0) * AST_C + j * 64 + (lane & 7) * 8;
#pragma unroll
                for (int it = 0; it < 2; ++it) {
                    const int rr = q * 32 + cc * 8 + it * 4 + (lane >> 3);
                    const uint32_t so = sw128_off(rr, lane & 7);
                    const uint4 H = *reinterpret_cast<const uint4*>(whi + so);
                    const uint4 L = *reinterpret_cast<const uint4*>(wlo + so);
                    *reinterpret_cast<uint4*>(p.a_hi + g0 + (size_t)rr * AST_C) = H;
                    *reinterpret_cast<uint4*>(p.a_lo + g0 + (size_t)rr * AST_C) = L;
                }
                named_bar_sync(5, 512);
                if (warp == 2 && lane == 0 && !mute) red_release_gpu_add(my_flag + j, 1u);
                if (lane == 0) mbar_arrive(bar + F2B_ACT(j));
            }
            if (threadIdx.x == 64) TLD(36 + j);
        };
        // Everything of h_i that nobody waits for -- sign mask, fp32 export, bf16 extract copy -- re-read from TMEM once the
        // hand-off is out: these stores must not sit in front of a release (membar) or delay the second K-half.
        auto finish = [&](int i) {
            if (!wr) return;
            const int f = i >= 1 ? (int)p.f_slot[i - 1] : -1;
            const int ts = i >= 1 ? (int)p.t_slot[i - 1] : -1;
            const bool exp = p.export_h[i] != 0;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int col = j * 64 + cc * 16;
                float v[16];
                tmem_ld16_nowait(hacc + lane_off + col, v);
                tmem_wait_ld();
                uint32_t word = 0;
#pragma unroll
                for (int e = 0; e < 16; ++e) or_bit_if_pos(word, v[e], 1u << e);
                hmask16[((size_t)i * p.R + row) * 8 + j * 4 + cc] = (uint16_t)word;
                if (exp) {                             // 16 columns = 64 bytes = two full sectors per thread
                    float* out = p.H + ((size_t)i * p.R + row) * AST_C + col;
#pragma unroll
                    for (int e = 0; e < 2; ++e)
                        st_global_v8(out + 8 * e, __float_as_uint(v[8 * e]), __float_as_uint(v[8 * e + 1]), __float_as_uint(v[8 * e + 2]),
                                     __float_as_uint(v[8 * e + 3]), __float_as_uint(v[8 * e + 4]), __float_as_uint(v[8 * e + 5]),
                                     __float_as_uint(v[8 * e + 6]), __float_as_uint(v[8 * e + 7]));
                }
                if (f >= 0 && !p.F_lo)                 // bf16 extract copy: 16 columns = 32 bytes = one full sector per thread
                    st_global_v8(p.F + ((size_t)f * p.R + row) * AST_C + col,
                                 pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]),
                                 pack_bf16(v[8], v[9]), pack_bf16(v[10], v[11]), pack_bf16(v[12], v[13]), pack_bf16(v[14], v[15]));
                if (f >= 0 && p.F_lo) {                // accurate tier: hi and lo parts
                    uint32_t Fh[8], Fl[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) split_bf16x2(v[2 * e], v[2 * e + 1], Fh[e], Fl[e]);
                    const size_t fo = ((size_t)f * p.R + row) * AST_C + col;
                    st_global_v8(p.F + fo, Fh[0], Fh[1], Fh[2], Fh[3], Fh[4], Fh[5], Fh[6], Fh[7]);
                    st_global_v8(p.F_lo + fo, Fl[0], Fl[1], Fl[2], Fl[3], Fl[4], Fl[5], Fl[6], Fl[7]);
                }
                if (ts >= 0) {                         // transposed fp16 copy [row / 32][slot][channel][row % 32]: a warp's 32 rows of one channel = one 64-byte run
                    __half* ht = p.Ht + ((((size_t)(row >> 5) * p.n_tslots + ts) * AST_C + col) << 5) + (row & 31);
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        unsigned short hv;
                        asm("cvt.rn.satfinite.f16.f32 %0, %1;" : "=h"(hv) : "f"(v[e]));
                        reinterpret_cast<unsigned short*>(ht)[e * 32] = hv;
                    }
                }
            }
        };
        // Near hand-off, consumer side: the d boundary rows of both neighbours (publish i) -> window halo, K-half by K-half.
        // Item e of a K-half = (halo side hs, cc, part, row slot): 64 bytes of LL lines -> two 16-byte chunks.
        auto import_halo = [&](int i) {
            const int l = i - 1;                       // for TLD
            const int d = stack_dil(i);
            const uint32_t tag = (epoch << 6) | (uint32_t)(i + 1);
            const int e = (warp - 2) * 32 + lane;
            const int slot = e % d, rest = e / d, part = rest & 1, c = (rest >> 1) & 3, hs = rest >> 3;      // slot fastest: coalesced
            const bool active = hs < 2 && (hs == 0 ? has_up : has_dn);
            const uint4* src[2];
            uint4 v[2][4];
            if (active) {                              // both K-halves' lines are requested at once: one L2 round trip, not two
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    src[j] = ll_item(i & 1, hs == 0 ? tile - 1 : tile + 1, hs == 0 ? 1 : 0, j, slot, c, part);
                    ll_fetch(src[j], v[j]);
                }
            }
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                if (active) {
                    uint32_t w[8];
                    ll_wait_item(src[j], v[j], tag, w, ab, p.err);
                    const int wr = hs == 0 ? F2_HALO - d + slot : F2_HALO + 128 + slot;
                    uint8_t* wb = gen + (j * 2 + part) * F2_WBUF + wr * 128;
                    *reinterpret_cast<uint4*>(wb + (((2 * c) ^ (wr & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
                    *reinterpret_cast<uint4*>(wb + (((2 * c + 1) ^ (wr & 7)) << 4)) = make_uint4(w[4], w[5], w[6], w[7]);
                    fence_proxy_async_smem();
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(bar + F2B_HALO(j));
                if (threadIdx.x == 64) TLD(20 + j);
            }
        };

        if (LB > 0) {
            // ---- resume: h_{l_begin} of this row from the fp32 stream in global memory -- this rank's own, or, for the rows
            // of a time shard's window that only the neighbour computes exactly, the neighbour's over NVLink ----
            const int side = stack_peer_side(p.peer, t0, ab, p.err);
            const float* hrow = side < 0 ? p.H + ((size_t)LB * p.R + row) * AST_C
                                         : p.peer.src[side] + (size_t)(t0 + r + p.peer.row_off[side]) * AST_C;
            float4 hv[2][4];                            // all eight loads in flight at once: over NVLink a round trip is microseconds
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float* q4 = hrow + j * 64 + cc * 16 + 4 * e;
                    hv[j][e] = side < 0 ? *reinterpret_cast<const float4*>(q4) : ld_relaxed_sys_f4(q4);
                }
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                float v[16];
#pragma unroll
                for (int e = 0; e < 4; ++e) { v[4 * e] = hv[j][e].x; v[4 * e + 1] = hv[j][e].y; v[4 * e + 2] = hv[j][e].z; v[4 * e + 3] = hv[j][e].w; }
                emit(LB, j, v);
            }
            tmem_wait_st();
            if (stack_near(stack_dil(LB))) import_halo(LB);
        } else {
            // ---- pseudo layer: start conv (model.py:80-93), same arithmetic as startconv_fwd_kernel ----
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
            if (LE > 0) import_halo(0);
            finish(0);
        }
        for (int l = LB; l < LE; ++l) {
            const uint32_t par = (uint32_t)(l - LB) & 1u;
            // ---- epilogue 1: z = acc1 + bd -> sign mask, relu(z) hi/lo -> A operand of the 1x1 GEMM (window centre) ----
            stack_mbar_wait(bar + F2B_ACC1, par, ab, p.err);
            tc_fence_after();
            if (threadIdx.x == 64 && l < 30) TL(2 + l);
            if (threadIdx.x == 64) TLD(30);
            uint32_t zword = 0;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                if (j == 1) named_bar_sync(1 + q, 128);       // K-half 0 of the whole quadrant before anyone starts K-half 1
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
#pragma unroll
                for (int e = 0; e < 16; ++e) or_bit_if_pos(zword, v[e], 1u << (16 * j + e));
            }
            if (wr) {
                zmask16[((size_t)l * p.R + row) * 8 + cc] = (uint16_t)zword;
                zmask16[((size_t)l * p.R + row) * 8 + 4 + cc] = (uint16_t)(zword >> 16);
            }
            if (threadIdx.x == 64) TLD(32);
            // ---- epilogue 2: h' = h + (acc2 + br)  (same association as the reference, model.py:109-114) ----
            stack_mbar_wait(bar + F2B_ACC2, par, ab, p.err);
            tc_fence_after();
            if (threadIdx.x == 64) TLD(33);
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                if (j == 1) named_bar_sync(1 + q, 128);
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
            if (l + 1 < LE && stack_near(stack_dil(l + 1))) import_halo(l + 1);
            finish(l + 1);
            if (threadIdx.x == 64) TLD(41);
            if (threadIdx.x == 64 && l < 30) TL(32 + l);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) {
        TL(1);
        // the last CTA of this launch opens the next epoch (tags of this launch can then never validate again)
        __threadfence();
        if (atomicAdd(p.epoch + 1, 1u) == gridDim.x - 1) { p.epoch[1] = 0u; __threadfence(); atomicAdd(p.epoch, 1u); stack_peer_signal(p.peer); }
    }
    if (warp == 1) tmem_dealloc<512>(tmem);
}
