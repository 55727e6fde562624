// api.cu -- the C ABI (include/ast_b200.h): context, workspaces, step orchestration, CUDA graph.
#include "../../include/ast_b200.h"
#include "common.cuh"
#include "tc.cuh"

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <vector>

thread_local int ast_launch_counter = 0;
static thread_local char g_err[512] = "";

// A/B switches for profiling exist only in -DAST_EXPERIMENTS builds; the shipped library reads no environment variable.
static inline bool exp_switch(const char* name) {
#ifdef AST_EXPERIMENTS
    const char* v = getenv(name);
    return v && v[0] == '1';
#else
    (void)name;
    return false;
#endif
}

static int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(AST_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

static unsigned long long g_ctx_generation = 0;      // bumped by ast_create and ast_set_targets (lbfgs.cu graph cache)
struct ast_ctx {
    ast_config cfg;
    unsigned long long generation;
    int T, B, R, NL, L, n_cont, prec;
    int own_lo, own_hi;           // window-local owned rows
    int64_t total_T;
    int style_layer[AST_MAX_IDS]; // encoder layer (0..29) behind each style slot
    int cont_layer[AST_MAX_IDS];
    int head_buf[30];             // layer -> dF buffer index or -1
    int f_slot[30];               // layer -> style slot (Gatys bf16 extract copy) or -1
    int head_of_slot[AST_MAX_IDS];
    int n_head;
    int nmat, nmat_used, msize, n_side;
    int gram_chunk_rows, gram_nchunk, sse_nblk;
    bool gram_ours_tc;            // channel-wise Gram on tensor cores (gram_ours_tc.cu)
    bool ours_tma;                // ... fed by TMA from the transposed extract copies Ht the whole-stack forward writes (gram_ours_tma.cu)
    unsigned char* Ht;            // [R/32][L][128][32] fp16
    OursTma ours_tma_map;
    int64_t bytes;
    std::vector<void*> allocs;
    // weights (fp32 masters + transposes for the backward)
    float *W0, *b0, *Wd, *bd, *Wr, *br, *WdT, *WrT;
    // activations / masks
    float* H;          // [NL+1][R][128]
    uint32_t* hmask;   // [NL+1][R][4]
    uint32_t* zmask;   // [NL][R][4]
    float* g;          // [R][128]
    float* g_alt;      // [R][128]: second running-gradient buffer (backward segments ping-pong, stack_bwd.cuh)
    float* dz[2];      // fp32 mode: [R][128] fp32 each
    float* dF;         // [n_head][R][128]
    float* scratch3R;
    // heads
    float* gram_part;
    float* partials;   // [B*nmat*msize raw Gram | B content sse]
    int64_t n_partials;
    float *Gn, *Ssym, *ls_part, *sse_part;
    unsigned* heads_done;          // [B] arrival counters of the heads kernel (single-GPU step: last CTA writes the losses)
    float *Pc, *Ps;
    const float** d_Fptr;
    float** d_dFptr;
    OursPtrs ours_ptrs;            // the same tables by value for the tensor-core channel-wise head
    bool split;                    // AST_PREC_TC3: convs on tcgen05 with hi / lo operands both ways, heads in fp32 on CUDA cores
    unsigned* p2p_sig;             // [2 sides][2 directions] completion words the neighbour ranks signal (halo_exchange)
    std::vector<void*> ipc_open;   // neighbours' buffers mapped by ast_p2p_attach (closed in ast_destroy)
    const float** d_Hcptr;
    float lambd;
    // gamma term (stft_reg.cu): frames [reg_f_lo, reg_f_lo + reg_nf) of the clip are evaluated here, the ones
    // in [reg_own_f_lo, reg_own_f_hi) count towards this context's partial sum
    float gamma;
    int reg_f_lo, reg_nf, reg_own_f_lo, reg_own_f_hi, reg_f_total;
    float2* reg_tw; signed char* reg_sgn; float* reg_fpart; float* reg_dadx;
    bool have_w, have_t;
    TcState* tc;       // tensor-core path state (AST_PREC_BF16), else NULL
    // host-buffer step (ast_step_host)
    cudaStream_t hstream;
    cudaGraphExec_t gexec;
    cudaGraphExec_t gdev;                       // ast_step_graph: device-pointer step
    const float* gdev_x; float* gdev_losses; float* gdev_grad;
    cudaGraphExec_t gphase[3];                  // ast_step_timed: encoder fwd | gram + heads | backward
    float *d_x, *d_losses, *d_grad;
    float *p_x, *p_out;            // pinned; p_out = [B*4 losses | R gradient | 2 hand-off error words]
    int launches_per_step;         // kernels of the last forward + the last backward
    int launches_fwd, launches_bwd;
    bool use_stack;                // whole-stack forward kernel (one launch) instead of one kernel per layer
    bool use_stack_bwd;            // likewise for the backward sweep
    bool fuse_dots, fuse_heads;    // experiment switches (AST_B200_NO_DOTS / AST_B200_NO_FOLD)
    unsigned char h_kept[33];      // H[i] of the last forward is valid in global memory
    cudaEvent_t ev[5];
};

template <typename Tp>
static int dalloc(ast_ctx* c, Tp** p, size_t n) {
    void* q = nullptr;
    const size_t bytes = std::max<size_t>(n * sizeof(Tp), 256);
    cudaError_t e = cudaMalloc(&q, bytes);
    if (e != cudaSuccess) return fail(AST_ECUDA, "cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
    c->allocs.push_back(q);
    c->bytes += (int64_t)bytes;
    *p = reinterpret_cast<Tp*>(q);
    return AST_OK;
}
#define DA(ptr, n)                                   \
    do {                                             \
        int r_ = dalloc(c, &(ptr), (size_t)(n));     \
        if (r_ != AST_OK) { ast_destroy(c); return r_; } \
    } while (0)

static inline int dil(int l) { return 1 << (l % 10); }   // model.py:98
static inline float* Hl(const ast_ctx* c, int l) { return c->H + (size_t)l * c->R * AST_C; }
static inline uint32_t* hm(const ast_ctx* c, int l) { return c->hmask + (size_t)l * c->R * AST_MW; }
static inline uint32_t* zm(const ast_ctx* c, int l) { return c->zmask + (size_t)l * c->R * AST_MW; }
static inline float* dFb(const ast_ctx* c, int i) { return c->dF + (size_t)i * c->R * AST_C; }

extern "C" const char* ast_last_error(void) { return g_err; }
unsigned long long ast_ctx_generation(const ast_ctx* c) { return c ? c->generation : 0; }
int ast_set_error(int code, const char* msg) { return fail(code, "%s", msg); }   // lbfgs.cu
extern "C" const char* ast_version(void) { return "ast_b200 0.1 sm_100a"; }
extern "C" int64_t ast_workspace_bytes(const ast_ctx* c) { return c ? c->bytes : 0; }
extern "C" int ast_launches_per_step(const ast_ctx* c) { return c ? c->launches_per_step : 0; }

static void drop_graphs(ast_ctx* c) {
    if (c->gexec) { cudaGraphExecDestroy(c->gexec); c->gexec = nullptr; }
    if (c->gdev) { cudaGraphExecDestroy(c->gdev); c->gdev = nullptr; }
    for (auto& g : c->gphase) if (g) { cudaGraphExecDestroy(g); g = nullptr; }
}

extern "C" int ast_destroy(ast_ctx* c) {
    if (!c) return AST_OK;
    cudaSetDevice(c->cfg.device);
    if (c->tc) tc_destroy(c->tc);
    for (void* q : c->ipc_open) cudaIpcCloseMemHandle(q);
    if (c->gexec) cudaGraphExecDestroy(c->gexec);
    if (c->gdev) cudaGraphExecDestroy(c->gdev);
    for (auto& g : c->gphase) if (g) cudaGraphExecDestroy(g);
    if (c->hstream) cudaStreamDestroy(c->hstream);
    for (auto& e : c->ev) if (e) cudaEventDestroy(e);
    if (c->p_x) cudaFreeHost(c->p_x);
    if (c->p_out) cudaFreeHost(c->p_out);
    for (void* p : c->allocs) cudaFree(p);
    delete c;
    return AST_OK;
}

extern "C" int ast_create(const ast_config* cfg, ast_ctx** out) {
    if (!cfg || !out) return fail(AST_EINVAL, "null argument");
    *out = nullptr;
    if (cfg->T <= 0 || cfg->T % 128 != 0) return fail(AST_EINVAL, "T=%d must be a positive multiple of 128", cfg->T);
    if (cfg->B <= 0) return fail(AST_EINVAL, "B=%d must be positive", cfg->B);
    if ((int64_t)cfg->B * cfg->T > (int64_t)1 << 30) return fail(AST_EINVAL, "B*T too large");
    if (cfg->n_layers < 1 || cfg->n_layers > 30) return fail(AST_EINVAL, "n_layers=%d outside 1..30", cfg->n_layers);
    if (cfg->n_cont < 1 || cfg->n_cont > AST_MAX_IDS) return fail(AST_EINVAL, "n_cont=%d outside 1..%d", cfg->n_cont, AST_MAX_IDS);
    if (cfg->n_style < 1 || cfg->n_style > AST_MAX_IDS) return fail(AST_EINVAL, "n_style=%d outside 1..%d", cfg->n_style, AST_MAX_IDS);
    if (cfg->nb_channels < 1 || cfg->nb_channels > 128) return fail(AST_EINVAL, "nb_channels outside 1..128");
    if (cfg->cnt_channels < 1 || cfg->cnt_channels > 128) return fail(AST_EINVAL, "cnt_channels outside 1..128");
    if (cfg->precision != AST_PREC_FP32 && cfg->precision != AST_PREC_BF16 && cfg->precision != AST_PREC_TC3)
        return fail(AST_EINVAL, "unknown precision %d", cfg->precision);
    if (cfg->kernel_path < AST_PATH_AUTO || cfg->kernel_path > AST_PATH_PER_LAYER_BWD) return fail(AST_EINVAL, "unknown kernel_path %d", cfg->kernel_path);

    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(AST_ECUDA, "no CUDA device: this library has no CPU fallback");
    if (cfg->device < 0 || cfg->device >= ndev) return fail(AST_EINVAL, "device %d outside 0..%d", cfg->device, ndev - 1);
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major != 10) return fail(AST_ECUDA, "device %d is sm_%d%d; this library is built for sm_100a only", cfg->device, prop.major, prop.minor);
    CK(cudaSetDevice(cfg->device));

    ast_ctx* c = new ast_ctx();
    memset(static_cast<void*>(&c->cfg), 0, sizeof(ast_config));
    c->cfg = *cfg;
    c->T = cfg->T; c->B = cfg->B; c->R = cfg->B * cfg->T; c->NL = cfg->n_layers; c->L = cfg->n_style;
    c->n_cont = cfg->n_cont; c->prec = cfg->precision; c->bytes = 0;
    c->generation = ++g_ctx_generation;
    c->p2p_sig = nullptr;
    c->ours_tma = false; c->Ht = nullptr;
    c->tc = nullptr; c->hstream = nullptr; c->gexec = nullptr; c->p_x = c->p_out = nullptr;
    c->gdev = nullptr; c->gdev_x = nullptr; c->gdev_losses = c->gdev_grad = nullptr;
    for (auto& g : c->gphase) g = nullptr;
    for (auto& e : c->ev) e = nullptr;
    c->have_w = c->have_t = false; c->launches_per_step = c->launches_fwd = c->launches_bwd = 0; c->lambd = 0.f;
    c->gamma = 0.f; c->reg_nf = 0; c->reg_tw = nullptr; c->reg_sgn = nullptr; c->reg_fpart = nullptr; c->reg_dadx = nullptr;

    // time-sharding window (all zero = unsharded)
    c->total_T = cfg->total_T > 0 ? cfg->total_T : cfg->T;
    int64_t ob = cfg->own_begin, oe = cfg->own_end;
    if (cfg->total_T <= 0) { ob = 0; oe = cfg->T; }
    c->own_lo = (int)(ob - cfg->win_begin); c->own_hi = (int)(oe - cfg->win_begin);
    if (c->own_lo < 0 || c->own_hi > c->T || c->own_lo >= c->own_hi) { delete c; return fail(AST_EINVAL, "owned rows outside the window"); }
    if (c->own_lo % 128 || c->own_hi % 128) { delete c; return fail(AST_EINVAL, "owned range must be aligned to 128 samples"); }

    // resolve heads.  extracts[30] aliases extracts[29] (model.py:118-119); [31] (bottleneck) is out of scope.
    for (int i = 0; i < 30; ++i) { c->head_buf[i] = -1; c->f_slot[i] = -1; }
    c->n_head = 0;
    bool seen[30] = {false};
    for (int i = 0; i < c->L; ++i) {
        int id = cfg->style_ids[i];
        if (id < 0 || id > 30) { delete c; return fail(AST_EINVAL, "style id %d outside 0..30", id); }
        id = std::min(id, 29);
        if (id >= c->NL) { delete c; return fail(AST_EINVAL, "style id %d needs more than n_layers=%d layers", id, c->NL); }
        if (seen[id]) { delete c; return fail(AST_EUNSUPPORTED, "encoder layer %d appears twice in the style set", id); }
        seen[id] = true;
        c->style_layer[i] = id;
        c->f_slot[id] = i;
        if (c->head_buf[id] < 0) c->head_buf[id] = c->n_head++;
        c->head_of_slot[i] = c->head_buf[id];
    }
    for (int i = 0; i < c->n_cont; ++i) {
        int id = cfg->cont_ids[i];
        if (id < 0 || id > 30) { delete c; return fail(AST_EINVAL, "content id %d outside 0..30", id); }
        id = std::min(id, 29);
        if (id >= c->NL) { delete c; return fail(AST_EINVAL, "content id %d needs more than n_layers=%d layers", id, c->NL); }
        c->cont_layer[i] = id;
        if (c->head_buf[id] < 0) c->head_buf[id] = c->n_head++;
    }
    if (c->head_buf[c->NL - 1] < 0) { delete c; return fail(AST_EINVAL, "n_layers=%d but no head reads extracts[%d]", c->NL, c->NL - 1); }

    if (cfg->gatys) { c->nmat = c->L; c->nmat_used = c->L; c->n_side = AST_C; }
    else { c->nmat = AST_C; c->nmat_used = cfg->nb_channels; c->n_side = c->L; }
    c->msize = c->n_side * c->n_side;

    {   // Gram split-K geometry: few enough chunks to keep the partial buffer small, enough to fill the SMs
        int target = cfg->gatys ? 16 : 128;      // ours: 128 chunks x 4 channel quads = 512 CTAs of the latency-bound kernel
        if (!cfg->gatys) target = 64;            // 64 chunks x 4 channel quads = 256 CTAs, all resident at 2 per SM; halves the partials
        c->gram_ours_tc = !cfg->gatys && cfg->precision == AST_PREC_BF16 && gram_ours_tc_supported(c->L) && !exp_switch("AST_B200_GRAM_OURS_CUDA");
        if (c->gram_ours_tc) target = 37;        // one CTA per SM (164 KB of smem): 37 chunks x 4 quads = 148 CTAs, one wave
        c->ours_tma = !cfg->gatys && cfg->precision == AST_PREC_BF16 && ours_tma_supported(c->L) && cfg->kernel_path != AST_PATH_PER_LAYER &&
                      !exp_switch("AST_B200_OURS_NO_TMA");
        if (c->ours_tma) target = 37;            // one CTA per SM: 37 chunks x 4 channel quads = 148 CTAs, one wave
        // (the split does NOT depend on B: pair b of a batch then sums in the same order as the same pair alone, so a batch
        // equals its pairs evaluated one by one bit for bit -- tests/test_gpu_configs.py)
        int rows = (c->T + target - 1) / target;
        rows = std::max(128, ((rows + 127) / 128) * 128);
        if (c->ours_tma) rows = ours_tma_chunk_rows(c->T, target);     // multiples of 32 rows: 37 chunks at T = 16384, not 32
        if (cfg->gatys && cfg->precision != AST_PREC_FP32) {
            // tensor-core Gram: one wave of CTAs, each accumulating its tiles in TMEM
            const int ntiles = c->T / 128;
            int nch = std::max(1, std::min(ntiles, 148 / std::max(1, c->L)));
            const int ct = (ntiles + nch - 1) / nch;
            rows = ct * 128;
        }
        c->gram_chunk_rows = rows;
        c->gram_nchunk = (c->T + rows - 1) / rows;
        c->sse_nblk = 296;                     // 2 CTAs per SM
    }

    const size_t RC = (size_t)c->R * AST_C;
    DA(c->W0, 3 * AST_C); DA(c->b0, AST_C);
    DA(c->Wd, (size_t)c->NL * 3 * AST_C * AST_C); DA(c->bd, (size_t)c->NL * AST_C);
    DA(c->Wr, (size_t)c->NL * AST_C * AST_C);     DA(c->br, (size_t)c->NL * AST_C);
    DA(c->WdT, (size_t)c->NL * 3 * AST_C * AST_C); DA(c->WrT, (size_t)c->NL * AST_C * AST_C);
    DA(c->H, (size_t)(c->NL + 1) * RC);
    DA(c->hmask, (size_t)(c->NL + 1) * c->R * AST_MW);
    DA(c->zmask, (size_t)c->NL * c->R * AST_MW);
    DA(c->g, RC);
    if (c->prec != AST_PREC_FP32) { DA(c->g_alt, RC); } else c->g_alt = nullptr;   // backward segments ping-pong between g and g_alt
    if (c->prec == AST_PREC_FP32) { DA(c->dz[0], RC); DA(c->dz[1], RC); } else { c->dz[0] = c->dz[1] = nullptr; }
    DA(c->dF, (size_t)c->n_head * RC);
    DA(c->scratch3R, (size_t)3 * c->R);
    const size_t gsz = (size_t)c->B * c->nmat * c->msize;
    DA(c->gram_part, gsz * c->gram_nchunk);
    c->n_partials = (int64_t)gsz + 2 * c->B;           // raw Gram | content sse | raw STFT-L1 sum (0 when gamma = 0)
    DA(c->partials, (size_t)c->n_partials);
    CK(cudaMemset(c->partials, 0, (size_t)c->n_partials * sizeof(float)));
    DA(c->Gn, gsz); DA(c->Ssym, gsz); DA(c->ls_part, (size_t)c->B * c->nmat);
    DA(c->sse_part, (size_t)2 * c->B * c->sse_nblk);   // doubles
    DA(c->heads_done, (size_t)c->B);
    CK(cudaMemset(c->heads_done, 0, (size_t)c->B * sizeof(unsigned)));
    DA(c->Pc, (size_t)c->R * c->n_cont * cfg->cnt_channels);
    DA(c->Ps, (size_t)c->B * c->nmat_used * c->msize);
    DA(c->d_Fptr, AST_MAX_IDS); DA(c->d_dFptr, AST_MAX_IDS); DA(c->d_Hcptr, AST_MAX_IDS);
    DA(c->d_x, (size_t)c->R); DA(c->d_losses, (size_t)c->B * 4); DA(c->d_grad, (size_t)c->R);
    {
        const float* fp[AST_MAX_IDS] = {nullptr}; float* dfp[AST_MAX_IDS] = {nullptr}; const float* hc[AST_MAX_IDS] = {nullptr};
        for (int i = 0; i < c->L; ++i) { fp[i] = Hl(c, c->style_layer[i] + 1); dfp[i] = dFb(c, c->head_buf[c->style_layer[i]]); }
        for (int i = 0; i < c->n_cont; ++i) hc[i] = Hl(c, c->cont_layer[i] + 1);
        memset(&c->ours_ptrs, 0, sizeof(c->ours_ptrs));
        for (int i = 0; i < c->L && i < 32; ++i) { c->ours_ptrs.F[i] = fp[i]; c->ours_ptrs.dF[i] = dfp[i]; }
        cudaMemcpy(c->d_Fptr, fp, sizeof(fp), cudaMemcpyHostToDevice);
        cudaMemcpy(c->d_dFptr, dfp, sizeof(dfp), cudaMemcpyHostToDevice);
        cudaMemcpy(c->d_Hcptr, hc, sizeof(hc), cudaMemcpyHostToDevice);
    }
    if (cudaMallocHost(&c->p_x, (size_t)c->R * sizeof(float)) != cudaSuccess ||
        cudaMallocHost(&c->p_out, ((size_t)c->R + 4 * c->B + 2) * sizeof(float)) != cudaSuccess) {
        ast_destroy(c); return fail(AST_ECUDA, "cudaMallocHost failed");
    }
    if (cudaStreamCreateWithFlags(&c->hstream, cudaStreamNonBlocking) != cudaSuccess) { ast_destroy(c); return fail(AST_ECUDA, "stream create failed"); }
    for (auto& e : c->ev) if (cudaEventCreate(&e) != cudaSuccess) { ast_destroy(c); return fail(AST_ECUDA, "event create failed"); }
    if (setup_f32_kernels() != 0) { ast_destroy(c); return fail(AST_ECUDA, "cudaFuncSetAttribute failed (is the GPU sm_100?)"); }
    c->fuse_dots = !exp_switch("AST_B200_NO_DOTS"); c->fuse_heads = !exp_switch("AST_B200_NO_FOLD");
    c->split = c->prec == AST_PREC_TC3;
    if (c->prec != AST_PREC_FP32) {
        const char* why = nullptr;
        // accurate tier: hi / lo backward operands; its Gatys heads multiply hi / lo bf16 pairs of the extracts and of Ssym on
        // tcgen05 (gram_tc.cu: *_split_kernel), its channel-wise heads read the fp32 extracts on CUDA cores
        c->tc = tc_create(c->T, c->B, c->NL, c->H, c->g, c->dF, c->n_head, cfg->gatys ? c->L : 0, c->gram_part,
                          (int64_t)c->B * c->nmat * c->gram_nchunk, c->split, &c->bytes, &why);
        if (!c->tc) { ast_destroy(c); return fail(AST_ECUDA, "tensor-core path setup failed: %s", why ? why : "?"); }
        // cfg->kernel_path: AST_PATH_PER_LAYER = one kernel per layer both ways, AST_PATH_PER_LAYER_BWD = whole-stack forward only
        c->use_stack = tc_stack_ok(c->tc) && cfg->kernel_path != AST_PATH_PER_LAYER;
        c->use_stack_bwd = c->use_stack && cfg->kernel_path != AST_PATH_PER_LAYER_BWD;
        if (c->split && !c->use_stack_bwd) { ast_destroy(c); return fail(AST_EUNSUPPORTED, "AST_PREC_TC3 runs on the whole-stack kernels only (kernel_path auto)"); }
        if (c->ours_tma && !c->use_stack) c->ours_tma = false;         // only the whole-stack forward writes the transposed copies
        if (c->ours_tma) {
            DA(c->Ht, ours_tma_bytes(c->L, c->R));
            CK(cudaMemset(c->Ht, 0, ours_tma_bytes(c->L, c->R)));            // rows no launch writes (window edges of a time shard) stay finite
            if (ours_tma_make_map(&c->ours_tma_map, c->Ht, c->L, c->R) != 0) { ast_destroy(c); return fail(AST_ECUDA, "tensor map over the transposed extracts failed"); }
            tc_set_transposed_extracts(c->tc, c->Ht, c->L);
        }
        if (cfg->halo_exchange) {
            if (!c->use_stack_bwd) { ast_destroy(c); return fail(AST_EUNSUPPORTED, "halo_exchange needs the whole-stack kernels (kernel_path auto)"); }
            tc_plan_segments(c->tc, true);
            DA(c->p2p_sig, 8);                             // 4 completion words + 2 evaluation counters (tc_set_peer_geometry)
            CK(cudaMemset(c->p2p_sig, 0, 8 * sizeof(unsigned)));
            tc_set_peer_geometry(c->tc, c->own_lo, c->own_hi, c->p2p_sig);
        }
    } else if (cfg->halo_exchange) {
        ast_destroy(c); return fail(AST_EUNSUPPORTED, "halo_exchange is a tensor-core tier (AST_PREC_BF16) feature");
    }
    if (cudaDeviceSynchronize() != cudaSuccess || cudaGetLastError() != cudaSuccess) { ast_destroy(c); return fail(AST_ECUDA, "context setup failed"); }
    *out = c;
    return AST_OK;
}

// ---- weights --------------------------------------------------------------------------------
// out[m][j][i] = in[m][i][j] for 128x128 matrices (backward operands Wd[k]^T, Wr^T)
__global__ void transpose128_kernel(const float* __restrict__ in, float* __restrict__ out) {
    __shared__ float tile[32][33];
    const size_t base = (size_t)blockIdx.z * AST_C * AST_C;
    const int x = blockIdx.x * 32 + threadIdx.x, y0 = blockIdx.y * 32;
    for (int j = threadIdx.y; j < 32; j += 8) tile[j][threadIdx.x] = in[base + (size_t)(y0 + j) * AST_C + x];
    __syncthreads();
    const int xo = blockIdx.y * 32 + threadIdx.x, yo0 = blockIdx.x * 32;
    for (int j = threadIdx.y; j < 32; j += 8) out[base + (size_t)(yo0 + j) * AST_C + xo] = tile[threadIdx.x][j];
}

extern "C" int ast_set_weights(ast_ctx* c, const float* W0, const float* b0, const float* Wd, const float* bd,
                               const float* Wr, const float* br, void* stream) {
    if (!c || !W0 || !b0 || !Wd || !bd || !Wr || !br) return fail(AST_EINVAL, "null argument");
    cudaStream_t s = (cudaStream_t)stream;
    CK(cudaSetDevice(c->cfg.device));
    const size_t f = sizeof(float);
    CK(cudaMemcpyAsync(c->W0, W0, 3 * AST_C * f, cudaMemcpyDefault, s));
    CK(cudaMemcpyAsync(c->b0, b0, AST_C * f, cudaMemcpyDefault, s));
    CK(cudaMemcpyAsync(c->Wd, Wd, (size_t)c->NL * 3 * AST_C * AST_C * f, cudaMemcpyDefault, s));
    CK(cudaMemcpyAsync(c->bd, bd, (size_t)c->NL * AST_C * f, cudaMemcpyDefault, s));
    CK(cudaMemcpyAsync(c->Wr, Wr, (size_t)c->NL * AST_C * AST_C * f, cudaMemcpyDefault, s));
    CK(cudaMemcpyAsync(c->br, br, (size_t)c->NL * AST_C * f, cudaMemcpyDefault, s));
    transpose128_kernel<<<dim3(4, 4, c->NL * 3), dim3(32, 8), 0, s>>>(c->Wd, c->WdT);
    transpose128_kernel<<<dim3(4, 4, c->NL), dim3(32, 8), 0, s>>>(c->Wr, c->WrT);
    if (c->tc && tc_set_weights(c->tc, c->Wd, c->Wr, s) != 0) return fail(AST_ECUDA, "tc_set_weights failed");
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(s));     // the caller's host buffers may be reused after return
    c->have_w = true;
    return AST_OK;
}

static int setup_regulariser(ast_ctx* c);
extern "C" int ast_set_targets(ast_ctx* c, const float* Pc, const float* Ps, double lambd, double gamma, void* stream) {
    if (!c || !Pc || !Ps) return fail(AST_EINVAL, "null argument");
    cudaStream_t s = (cudaStream_t)stream;
    CK(cudaSetDevice(c->cfg.device));
    if (gamma != 0.0) {
        int r = setup_regulariser(c);
        if (r != AST_OK) return r;
    }
    c->gamma = (float)gamma;
    CK(cudaMemcpyAsync(c->Pc, Pc, (size_t)c->R * c->n_cont * c->cfg.cnt_channels * sizeof(float), cudaMemcpyDefault, s));
    CK(cudaMemcpyAsync(c->Ps, Ps, (size_t)c->B * c->nmat_used * c->msize * sizeof(float), cudaMemcpyDefault, s));
    CK(cudaStreamSynchronize(s));
    c->lambd = (float)lambd;
    c->have_t = true;
    c->generation = ++g_ctx_generation;
    drop_graphs(c);                                                         // loss weights are baked in the graphs
    return AST_OK;
}

// ---- gamma term: which STFT frames this context evaluates (methods.py:122: frame 1024, hop 512, no end padding)
static int setup_regulariser(ast_ctx* c) {
    const int64_t Ttot = c->total_T, wb = c->cfg.total_T > 0 ? c->cfg.win_begin : 0;
    if (Ttot < 1024) return fail(AST_EINVAL, "gamma != 0 needs at least one 1024-sample STFT frame (T = %lld)", (long long)Ttot);
    const int64_t F = 1 + (Ttot - 1024) / 512;
    const int64_t ob = wb + c->own_lo, oe = wb + c->own_hi;
    // frames that touch an owned sample: 512 f <= n < 512 f + 1024 for some n in [ob, oe)
    int64_t f_lo = ob >= 1023 ? (ob - 1023 + 511) / 512 : 0, f_hi = std::min<int64_t>(F - 1, (oe - 1) / 512);
    if (f_hi < f_lo) { f_lo = 0; f_hi = -1; }
    if (f_hi >= f_lo && (512 * f_lo < wb || 512 * f_hi + 1024 > wb + c->T))
        return fail(AST_EINVAL, "gamma != 0: the window [%lld, %lld) does not hold every STFT frame that touches the owned samples",
                    (long long)wb, (long long)(wb + c->T));
    c->reg_f_lo = (int)f_lo; c->reg_nf = (int)(f_hi - f_lo + 1); c->reg_f_total = (int)F;
    c->reg_own_f_lo = (int)((ob + 511) / 512); c->reg_own_f_hi = (int)std::min<int64_t>(F, (oe + 511) / 512);
    if (c->reg_tw) return AST_OK;
    std::vector<float2> tw(1024);
    for (int j = 0; j < 1024; ++j) {
        // exact at the quadrant points, so that Im of bins 0 and 512 is exactly 0 as in an FFT
        const int q = j & 255, quad = j >> 8;
        const double ang = 2.0 * 3.14159265358979323846 * q / 1024.0;
        const double cq = cos(ang), sq = q ? sin(ang) : 0.0;
        double cj, sj;
        switch (quad) { case 0: cj = cq; sj = sq; break; case 1: cj = -sq; sj = cq; break;
                        case 2: cj = -cq; sj = -sq; break; default: cj = sq; sj = -cq; }
        tw[j] = make_float2((float)cj, (float)sj);
    }
    int r_;
    if ((r_ = dalloc(c, &c->reg_tw, 1024)) != AST_OK) return r_;
    CK(cudaMemcpy(c->reg_tw, tw.data(), sizeof(float2) * 1024, cudaMemcpyHostToDevice));
    // sized for any window position: at most T/512 + 2 frames touch T samples
    const size_t nfmax = (size_t)c->T / 512 + 2;
    if ((r_ = dalloc(c, &c->reg_sgn, (size_t)c->B * nfmax * 513 * 2)) != AST_OK) return r_;
    if ((r_ = dalloc(c, &c->reg_fpart, (size_t)c->B * nfmax * 4)) != AST_OK) return r_;
    if ((r_ = dalloc(c, &c->reg_dadx, (size_t)c->R)) != AST_OK) return r_;
    return AST_OK;
}

// ---- forward --------------------------------------------------------------------------------
// keep_all: every H[i] is written to global memory (ast_embeds / ast_get_extract); the optimisation step keeps only
// what its heads read -- the whole-stack kernel holds the residual stream in TMEM.
static int n_segments(const ast_ctx* c) { return (c->tc && c->use_stack_bwd) ? tc_segments(c->tc) : 1; }
static int run_encoder(ast_ctx* c, const float* d_x, cudaStream_t s, bool keep_all, int seg = -1) {
    if (c->tc && c->use_stack) {
        unsigned char keep[33];
        for (int i = 0; i <= c->NL; ++i) keep[i] = keep_all ? 1 : 0;
        for (int i = 0; i < c->n_cont; ++i) keep[c->cont_layer[i] + 1] = 1;
        if (!c->cfg.gatys && !c->ours_tma)      // the channel-wise heads of the fp32-extract kernels (accurate tier, per-layer path)
            for (int i = 0; i < c->L; ++i) keep[c->style_layer[i] + 1] = 1;
        if (c->ours_tma) keep[c->NL] = 1;       // the head reads the transposed copies; the last extract stays for ast_get_extract
        memcpy(c->h_kept, keep, sizeof(keep));
        if (tc_forward_stack(c->tc, d_x, c->W0, c->b0, c->bd, c->br, c->H, c->zmask, c->hmask, c->f_slot, keep, seg, s) != 0)
            return fail(AST_ECUDA, "whole-stack forward launch failed: %s", cudaGetErrorString(cudaGetLastError()));
        return AST_OK;
    }
    memset(c->h_kept, 1, sizeof(c->h_kept));
    launch_startconv_fwd(d_x, c->W0, c->b0, Hl(c, 0), hm(c, 0), c->T, c->R, s);
    if (c->tc) tc_prepare_input(c->tc, Hl(c, 0), s);
    for (int l = 0; l < c->NL; ++l) {
        if (c->tc) {
            tc_layer_fwd(c->tc, l, c->bd + (size_t)l * AST_C, c->br + (size_t)l * AST_C, zm(c, l), hm(c, l + 1),
                         c->f_slot[l], s);
            continue;
        }
        LayerArgsF32 a;
        memset(&a, 0, sizeof(a));
        a.A1 = Hl(c, l); a.W1 = c->Wd + (size_t)l * 3 * AST_C * AST_C; a.bias1 = c->bd + (size_t)l * AST_C;
        a.W2 = c->Wr + (size_t)l * AST_C * AST_C; a.bias2 = c->br + (size_t)l * AST_C;
        a.Hin = Hl(c, l); a.Hout = Hl(c, l + 1); a.zmask_out = zm(c, l); a.hmask_out = hm(c, l + 1);
        a.T = c->T; a.R = c->R; a.shift = dil(l);
        launch_layer_f32(a, false, s);
    }
    return AST_OK;
}
// fused (single-GPU ast_step*): nothing all-reduces the partials between the two phases, so the heads kernel sums the
// Gatys split-K chunks and the content partials itself and its last CTA writes the losses (3 launches fewer).
static inline bool fold_gram_reduce(const ast_ctx* c, bool fused) { return fused && c->cfg.gatys && c->gram_nchunk < 32; }
static void run_gram(ast_ctx* c, cudaStream_t s, bool fused = false) {
    if (c->cfg.gatys && c->tc)
        tc_gram(c->tc, c->L, c->gram_nchunk, c->gram_chunk_rows / 128, c->own_lo / 128, c->own_hi / 128, s);
    else if (c->cfg.gatys)
        launch_gram_gatys(c->d_Fptr, c->L, c->T, c->B, c->own_lo, c->own_hi, c->gram_part, c->gram_nchunk, c->gram_chunk_rows, s);
    else
        if (c->ours_tma) launch_gram_ours_tma(c->ours_tma_map, c->L, c->T, c->B, c->own_lo, c->own_hi, c->gram_part, c->gram_nchunk, c->gram_chunk_rows, s);
        else if (c->gram_ours_tc) launch_gram_ours_tc(c->ours_ptrs, c->L, c->T, c->B, c->own_lo, c->own_hi, c->gram_part, c->gram_nchunk, c->gram_chunk_rows, s);
        else launch_gram_ours(c->d_Fptr, c->L, c->T, c->B, c->own_lo, c->own_hi, c->gram_part, c->gram_nchunk, c->gram_chunk_rows, s);
    if (!fold_gram_reduce(c, fused))
        launch_gram_reduce(c->gram_part, c->partials, c->cfg.gatys, c->L, c->B, c->gram_nchunk, s);
}

// everything between the encoder and the (optional) all-reduce: raw Gram, content squared error, raw STFT-L1 sum
static void run_partials(ast_ctx* c, const float* d_x, cudaStream_t s, bool fused) {
    // (Forking the content SSE onto a second stream next to the Gram kernels was measured: no gain within run-to-run noise.)
    run_gram(c, s, fused);
    float* tail = c->partials + (size_t)c->B * c->nmat * c->msize;
    launch_content_sse(c->d_Hcptr, c->n_cont, c->cfg.cnt_channels, c->Pc, c->T, c->B, c->own_lo, c->own_hi,
                       c->sse_part, c->sse_nblk, fused ? nullptr : tail, s);
    if (c->gamma != 0.f && c->reg_nf > 0)
        launch_stft_l1_fwd(d_x, c->reg_tw, c->T, c->B, (int)(c->cfg.total_T > 0 ? c->cfg.win_begin : 0), c->reg_f_lo, c->reg_nf,
                           c->reg_own_f_lo, c->reg_own_f_hi, c->reg_sgn, c->reg_fpart, c->reg_dadx, tail + c->B, s);
}

// seg < 0: the whole forward.  seg >= 0 (time shards that exchange halos): encoder segment `seg` only; the last segment
// is followed by the partial sums.
static int step_forward_impl(ast_ctx* c, const float* d_x, void* stream, bool fused, int seg = -1) {
    if (!c || !d_x) return fail(AST_EINVAL, "null argument");
    if (!c->have_w || !c->have_t) return fail(AST_ESTATE, "ast_set_weights and ast_set_targets must precede a step");
    const int nseg = n_segments(c);
    if (seg >= nseg) return fail(AST_EINVAL, "segment %d outside 0..%d", seg, nseg - 1);
    cudaStream_t s = (cudaStream_t)stream;
    const int before = ast_launch_counter;
    const int re = run_encoder(c, d_x, s, false, nseg > 1 ? seg : -1);
    if (re != AST_OK) return re;
    if (seg < 0 || seg == nseg - 1) run_partials(c, d_x, s, fused);
    if (seg > 0) { c->launches_fwd += ast_launch_counter - before; c->launches_per_step = c->launches_fwd + c->launches_bwd; CK(cudaGetLastError()); return AST_OK; }
    c->launches_fwd = ast_launch_counter - before;
    c->launches_per_step = c->launches_fwd + c->launches_bwd;
    CK(cudaGetLastError());
    return AST_OK;
}

extern "C" int ast_step_forward(ast_ctx* c, const float* d_x, void* stream) { return step_forward_impl(c, d_x, stream, false); }

extern "C" int ast_partials(ast_ctx* c, float** d_buf, int64_t* count) {
    if (!c || !d_buf || !count) return fail(AST_EINVAL, "null argument");
    *d_buf = c->partials;
    *count = c->n_partials;
    return AST_OK;
}

static HeadArgs head_args(ast_ctx* c, bool with_targets, float* losses, bool fused = false) {
    HeadArgs h;
    memset(&h, 0, sizeof(h));
    const double Ns = (double)c->nmat_used * c->msize;
    const double Nc = (double)c->total_T * c->n_cont * c->cfg.cnt_channels;
    h.G = c->partials; h.Ps = c->Ps; h.Gn = c->Gn; h.Ssym = with_targets ? c->Ssym : nullptr;
    h.Ssym_bf16 = (with_targets && c->cfg.gatys && c->tc) ? tc_ssym_bf16(c->tc) : nullptr;
    h.Ssym_bf16_lo = (with_targets && c->cfg.gatys && c->tc && c->split) ? tc_ssym_bf16_lo(c->tc) : nullptr;
    h.ls_part = c->ls_part; h.sse = c->partials + (size_t)c->B * c->nmat * c->msize; h.losses = losses;
    h.nmat = c->nmat; h.nmat_used = with_targets ? c->nmat_used : 0; h.n = c->n_side; h.B = c->B;
    if (fused) {
        if (fold_gram_reduce(c, true)) { h.gram_part = c->gram_part; h.nchunk = c->gram_nchunk; }
        h.sse_part = reinterpret_cast<const double*>(c->sse_part); h.sse_nblk = c->sse_nblk;
        h.done = c->heads_done;
    }
    h.reg_raw = h.sse + c->B; h.gamma = c->gamma;
    h.reg_scale = c->gamma != 0.f ? (float)(1.0 / ((double)c->reg_f_total * 513.0)) : 0.f;
    h.coef_dGn = (float)((double)c->lambd * 1e3 * 2.0 / Ns);
    h.ls_scale = (float)(1e3 / Ns); h.lc_scale = (float)(10.0 / Nc); h.lambd = c->lambd;
    return h;
}

// ---- backward -------------------------------------------------------------------------------
// seg < 0: the whole backward.  seg >= 0: the heads + head gradients go with the TOP segment (seg == nseg-1, called
// first), the start-conv adjoint and the gamma term with segment 0 (called last).
static int step_backward_impl(ast_ctx* c, float* d_losses, float* d_grad, void* stream, bool fused, int seg = -1) {
    if (!c || !d_losses || !d_grad) return fail(AST_EINVAL, "null argument");
    if (!c->have_w || !c->have_t) return fail(AST_ESTATE, "ast_set_weights and ast_set_targets must precede a step");
    const int nseg = n_segments(c);
    if (seg >= nseg) return fail(AST_EINVAL, "segment %d outside 0..%d", seg, nseg - 1);
    if (nseg == 1) seg = -1;
    const bool do_top = seg < 0 || seg == nseg - 1, do_bottom = seg <= 0;
    cudaStream_t s = (cudaStream_t)stream;
    const int bwd_before = ast_launch_counter;
    if (do_top) launch_heads(head_args(c, true, d_losses, fused), s);
    // Gatys + tensor cores: the style gradient of an extract is fused into the backward layer kernel unless the
    // same extract also carries a content term (then the standalone kernel fills its dF buffer first)
    bool is_cont[30];
    for (int l = 0; l < 30; ++l) is_cont[l] = false;
    for (int i = 0; i < c->n_cont; ++i) is_cont[c->cont_layer[i]] = true;
    const bool fuse_sg = c->cfg.gatys && c->tc && !c->split;
    if (!do_top) { /* head gradients were formed with the top segment */ }
    else if (fuse_sg) {
        int hos[AST_MAX_IDS];
        for (int i = 0; i < c->L; ++i) hos[i] = is_cont[c->style_layer[i]] ? c->head_of_slot[i] : -1;
        tc_style_grad(c->tc, c->L, hos, s);
    }
    else if (c->cfg.gatys && c->tc && c->split) {       // accurate tier: every slot through the hi / lo tcgen05 kernel; the backward reads dF
        int hos[AST_MAX_IDS];
        for (int i = 0; i < c->L; ++i) hos[i] = c->head_of_slot[i];
        tc_style_grad(c->tc, c->L, hos, s);
    }
    else if (c->cfg.gatys) launch_style_grad_gatys(c->d_Fptr, c->d_dFptr, c->Ssym, c->L, c->T, c->B, s);
    else if (c->ours_tma) launch_style_grad_ours_tma(c->ours_tma_map, c->ours_ptrs, c->Ssym, c->L, c->T, c->B, s);
    else if (c->gram_ours_tc && !exp_switch("AST_B200_SGO_CUDA")) launch_style_grad_ours_tc(c->ours_ptrs, c->Ssym, c->L, c->T, c->B, s);
    else              launch_style_grad_ours(c->d_Fptr, c->d_dFptr, c->Ssym, c->L, c->T, c->B, s);
    if (do_top) {
        bool init[30];
        for (int l = 0; l < 30; ++l) init[l] = false;
        for (int i = 0; i < c->L; ++i) init[c->style_layer[i]] = true;
        const double Nc = (double)c->total_T * c->n_cont * c->cfg.cnt_channels;
        for (int i = 0; i < c->n_cont; ++i) {
            const int l = c->cont_layer[i];
            launch_content_grad(Hl(c, l + 1), c->Pc, dFb(c, c->head_buf[l]), i, c->n_cont, c->cfg.cnt_channels,
                                (float)(20.0 / Nc), init[l] ? 1 : 0, c->R, s);
            init[l] = true;
        }
    }
    if (c->tc && c->use_stack_bwd) {
        int df_idx[33], sg_slot[33];
        for (int l = 0; l <= c->NL; ++l) {
            df_idx[l] = (l >= 1) ? c->head_buf[l - 1] : -1;
            sg_slot[l] = -1;
            if (fuse_sg && l >= 1 && c->f_slot[l - 1] >= 0 && !is_cont[l - 1]) { sg_slot[l] = c->f_slot[l - 1]; df_idx[l] = -1; }
        }
        // the start-conv adjoint's channel dots are taken from the final running gradient while it is still in registers
        if (tc_backward_stack(c->tc, c->hmask, c->zmask, df_idx, sg_slot, c->L, c->g, c->g_alt, c->W0, c->fuse_dots ? c->scratch3R : nullptr, seg, s) != 0)
            return fail(AST_ECUDA, "whole-stack backward launch failed: %s", cudaGetErrorString(cudaGetLastError()));
    }
    int cur = 0;
    for (int l = c->NL; l >= 0 && !(c->tc && c->use_stack_bwd); --l) {
        const float* dF = (l >= 1 && c->head_buf[l - 1] >= 0) ? dFb(c, c->head_buf[l - 1]) : nullptr;
        if (c->tc) {
            int df_idx = (l >= 1) ? c->head_buf[l - 1] : -1, sg_slot = -1;
            if (fuse_sg && l >= 1 && c->f_slot[l - 1] >= 0 && !is_cont[l - 1]) { sg_slot = c->f_slot[l - 1]; df_idx = -1; }
            tc_layer_bwd(c->tc, l, l < c->NL ? hm(c, l) : nullptr, df_idx, sg_slot, c->L, l >= 1 ? zm(c, l - 1) : nullptr, s);
            continue;
        }
        LayerArgsF32 a;
        memset(&a, 0, sizeof(a));
        a.T = c->T; a.R = c->R; a.g = c->g; a.dF = dF;
        if (l < c->NL) {
            a.A1 = c->dz[cur]; a.W1 = c->WdT + (size_t)l * 3 * AST_C * AST_C; a.shift = -dil(l); a.hmask_in = hm(c, l);
        } else {
            a.g_init = 1;
        }
        if (l >= 1) {
            a.W2 = c->WrT + (size_t)(l - 1) * AST_C * AST_C; a.zmask_in = zm(c, l - 1); a.dz_out = c->dz[cur ^ 1];
        }
        launch_layer_f32(a, true, s);
        cur ^= 1;
    }
    if (!do_bottom) {
        c->launches_bwd = (do_top ? 0 : c->launches_bwd) + ast_launch_counter - bwd_before;
        c->launches_per_step = c->launches_fwd + c->launches_bwd;
        CK(cudaGetLastError());
        return AST_OK;
    }
    if (c->tc && c->use_stack_bwd && c->fuse_dots) launch_startconv_bwd_combine(c->scratch3R, d_grad, c->T, c->R, s);
    else launch_startconv_bwd(c->g, c->W0, c->scratch3R, d_grad, c->T, c->R, s);
    if (c->gamma != 0.f && c->reg_nf > 0)                 // + gamma * d reg / d x (methods.py:125)
        launch_stft_l1_grad(c->reg_dadx, c->reg_tw, c->reg_sgn, c->T, c->B, (int)(c->cfg.total_T > 0 ? c->cfg.win_begin : 0),
                            c->reg_f_lo, c->reg_nf, (float)((double)c->gamma / ((double)c->reg_f_total * 513.0)), d_grad, s);
    c->launches_bwd = (do_top ? 0 : c->launches_bwd) + ast_launch_counter - bwd_before;
    c->launches_per_step = c->launches_fwd + c->launches_bwd;
    CK(cudaGetLastError());
    return AST_OK;
}

// ---- per-stack segments for time shards that exchange halo rows (include/ast_b200.h) ----------------------------
extern "C" int ast_segments(const ast_ctx* c) { return c ? n_segments(c) : 0; }
extern "C" int ast_step_forward_seg(ast_ctx* c, const float* d_x, int32_t seg, void* stream) {
    if (seg < 0) return fail(AST_EINVAL, "segment %d outside 0..", seg);
    return step_forward_impl(c, d_x, stream, false, seg);
}
extern "C" int ast_step_backward_seg(ast_ctx* c, int32_t seg, float* d_losses, float* d_grad, void* stream) {
    if (seg < 0) return fail(AST_EINVAL, "segment %d outside 0..", seg);
    return step_backward_impl(c, d_losses, d_grad, stream, false, seg);
}
extern "C" int ast_segment_buffer(ast_ctx* c, int32_t backward, int32_t seg, float** d_buf) {
    if (!c || !d_buf) return fail(AST_EINVAL, "null argument");
    const int nseg = n_segments(c);
    if (seg < 0 || seg >= nseg) return fail(AST_EINVAL, "segment %d outside 0..%d", seg, nseg - 1);
    if (backward) { *d_buf = (seg & 1) ? c->g_alt : c->g; return AST_OK; }
    if (seg == nseg - 1) return fail(AST_EINVAL, "the last forward segment leaves nothing to exchange");
    *d_buf = Hl(c, tc_segment_end(c->tc, seg));
    return AST_OK;
}

// ---- peer-mapped halos: the neighbour ranks' buffers inside this process (include/ast_b200.h) ----------------------
static int p2p_ready(ast_ctx* c, int side) {
    if (!c) return fail(AST_EINVAL, "null argument");
    if (!c->p2p_sig) return fail(AST_ESTATE, "ast_p2p_*: the context was not created with halo_exchange = 1");
    if (side < 0 || side > 1) return fail(AST_EINVAL, "side %d: 0 = the rank before this window, 1 = after", side);
    if (c->B != 1) return fail(AST_EUNSUPPORTED, "peer-mapped halos are built for B = 1");
    return AST_OK;
}
static int p2p_set(ast_ctx* c, int side, const float* H, const float* g, const float* g_alt, unsigned* sig, int64_t peer_wb, int peer_T) {
    const int64_t off = c->cfg.win_begin - peer_wb;              // neighbour row = in-window t + off
    if (off < -(int64_t)1 << 30 || off > (int64_t)1 << 30) return fail(AST_EINVAL, "window offsets too far apart");
    // the rows this rank reads must lie inside the neighbour's window
    const int lo = c->own_lo, hi = c->own_hi;
    const int64_t a = side == 0 ? lo - 2 * AST_HALO_ROWS + off : hi + off, b = side == 0 ? lo + off : hi + 2 * AST_HALO_ROWS + off;
    if (a < 0 || b > peer_T) return fail(AST_EINVAL, "the neighbour's window [%lld, +%d) does not hold the halo rows of side %d", (long long)peer_wb, peer_T, side);
    tc_set_peer(c->tc, side, H, g, g_alt, sig, peer_T, (int)off);
    return AST_OK;
}
extern "C" int ast_p2p_export(ast_ctx* c, void* h_handles) {
    int r = p2p_ready(c, 0);
    if (r != AST_OK) return r;
    if (!h_handles) return fail(AST_EINVAL, "null argument");
    cudaIpcMemHandle_t* h = static_cast<cudaIpcMemHandle_t*>(h_handles);
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaIpcGetMemHandle(&h[0], c->H));
    CK(cudaIpcGetMemHandle(&h[1], c->g));
    CK(cudaIpcGetMemHandle(&h[2], c->g_alt));
    CK(cudaIpcGetMemHandle(&h[3], c->p2p_sig));
    return AST_OK;
}
extern "C" int ast_p2p_attach(ast_ctx* c, int32_t side, const void* h_handles, int64_t peer_win_begin, int32_t peer_T) {
    int r = p2p_ready(c, side);
    if (r != AST_OK) return r;
    if (!h_handles) return fail(AST_EINVAL, "null argument");
    const cudaIpcMemHandle_t* h = static_cast<const cudaIpcMemHandle_t*>(h_handles);
    CK(cudaSetDevice(c->cfg.device));
    void* q[4] = {nullptr, nullptr, nullptr, nullptr};
    for (int i = 0; i < 4; ++i) {
        cudaError_t e = cudaIpcOpenMemHandle(&q[i], h[i], cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) return fail(AST_ECUDA, "cudaIpcOpenMemHandle failed: %s (is peer access between the two GPUs possible?)", cudaGetErrorString(e));
        c->ipc_open.push_back(q[i]);
    }
    return p2p_set(c, side, (const float*)q[0], (const float*)q[1], (const float*)q[2], (unsigned*)q[3], peer_win_begin, peer_T);
}
// Back to exchanging the halo rows by copies (ast_segment_buffer): forget the neighbours (their mappings stay open until
// ast_destroy).  Every rank of the job must do the same.
extern "C" int ast_p2p_detach(ast_ctx* c) {
    int r = p2p_ready(c, 0);
    if (r != AST_OK) return r;
    tc_clear_peers(c->tc);
    return AST_OK;
}
// Same process (several contexts of one process: tests, one-process multi-GPU with peer access enabled by the caller).
extern "C" int ast_p2p_attach_local(ast_ctx* c, int32_t side, ast_ctx* peer) {
    int r = p2p_ready(c, side);
    if (r != AST_OK) return r;
    if ((r = p2p_ready(peer, side)) != AST_OK) return r;
    return p2p_set(c, side, peer->H, peer->g, peer->g_alt, peer->p2p_sig, peer->cfg.win_begin, peer->T);
}

extern "C" int ast_step_backward(ast_ctx* c, float* d_losses, float* d_grad, void* stream) {
    return step_backward_impl(c, d_losses, d_grad, stream, false);
}

extern "C" int ast_step(ast_ctx* c, const float* d_x, float* d_losses, float* d_grad, void* stream) {
    int r = step_forward_impl(c, d_x, stream, c->fuse_heads);
    if (r != AST_OK) return r;
    return step_backward_impl(c, d_losses, d_grad, stream, c->fuse_heads);
}

// ---- get_embeds (methods.py:86-95) -------------------------------------------------------------
extern "C" int ast_embeds(ast_ctx* c, const float* d_x, float* d_Ec, float* d_Gn, void* stream) {
    if (!c || !d_x) return fail(AST_EINVAL, "null argument");
    if (!c->have_w) return fail(AST_ESTATE, "ast_set_weights must precede ast_embeds");
    cudaStream_t s = (cudaStream_t)stream;
    const int re = run_encoder(c, d_x, s, true);
    if (re != AST_OK) return re;
    if (d_Ec) launch_content_gather(c->d_Hcptr, c->n_cont, c->cfg.cnt_channels, d_Ec, c->R, s);
    if (d_Gn) {
        run_gram(c, s);
        launch_heads(head_args(c, false, nullptr), s);
        const size_t used = (size_t)c->nmat_used * c->msize * sizeof(float), full = (size_t)c->nmat * c->msize * sizeof(float);
        CK(cudaMemcpy2DAsync(d_Gn, used, c->Gn, full, used, c->B, cudaMemcpyDeviceToDevice, s));
    }
    CK(cudaGetLastError());
    return AST_OK;
}

extern "C" int ast_get_extract(ast_ctx* c, int32_t layer, float* d_out, void* stream) {
    if (!c || !d_out) return fail(AST_EINVAL, "null argument");
    if (layer < -1 || layer > 30) return fail(AST_EINVAL, "layer %d outside -1..30", layer);
    const int l = std::min(layer, 29);        // -1 = start-conv output
    if (l >= c->NL) return fail(AST_EINVAL, "layer %d was not evaluated (n_layers=%d)", layer, c->NL);
    if (!c->h_kept[l + 1])
        return fail(AST_ESTATE, "layer %d was not retained by the last step (only the configured content/style extracts "
                    "are; ast_embeds retains every layer)", layer);
    CK(cudaMemcpyAsync(d_out, Hl(c, l + 1), (size_t)c->R * AST_C * sizeof(float), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return AST_OK;
}

extern "C" int ast_get_masks(ast_ctx* c, uint32_t* d_hmask, uint32_t* d_zmask, void* stream) {
    if (!c || !d_hmask || !d_zmask) return fail(AST_EINVAL, "null argument");
    const size_t n = (size_t)c->NL * c->R * AST_MW * sizeof(uint32_t);
    CK(cudaMemcpyAsync(d_hmask, c->hmask, n, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    CK(cudaMemcpyAsync(d_zmask, c->zmask, n, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return AST_OK;
}

// ---- CUDA graphs ---------------------------------------------------------------------------------
// Captures whatever `body` enqueues on the context's private stream into an executable graph.
template <typename F>
static int capture_graph(ast_ctx* c, cudaGraphExec_t* out, F body) {
    cudaGraph_t graph = nullptr;
    CK(cudaStreamBeginCapture(c->hstream, cudaStreamCaptureModeThreadLocal));
    const int r = body(c->hstream);
    const cudaError_t e = cudaStreamEndCapture(c->hstream, &graph);
    if (r != AST_OK) { if (graph) cudaGraphDestroy(graph); return r; }
    if (e != cudaSuccess) return fail(AST_ECUDA, "graph capture failed: %s", cudaGetErrorString(e));
    const cudaError_t e2 = cudaGraphInstantiate(out, graph, 0);
    cudaGraphDestroy(graph);
    if (e2 != cudaSuccess) { *out = nullptr; return fail(AST_ECUDA, "graph instantiate failed: %s", cudaGetErrorString(e2)); }
    return AST_OK;
}

// A bounded wait of the whole-stack kernels expired (stack_fwd.cuh: 1 = a neighbour's hand-off flag, 2 = an mbarrier):
// the results of that evaluation are invalid.  The words stay set (sticky) until ast_check clears them.
static int stack_fault(ast_ctx* c, unsigned efwd, unsigned ebwd) {
    (void)c;
    return fail(AST_ECUDA, "whole-stack kernel hand-off timed out (forward %u, backward %u; 1 = neighbour flag, 2 = mbarrier): "
                "the results of this evaluation are invalid", efwd, ebwd);
}
int ast_stack_fault_words(ast_ctx* c, unsigned* e2) {                 // lbfgs.cu (status read-back of ast_minimize)
    e2[0] = e2[1] = 0;
    if (!c || !c->tc) return AST_OK;
    CK(cudaMemcpy(e2, tc_stack_err(c->tc), 2 * sizeof(unsigned), cudaMemcpyDeviceToHost));
    return AST_OK;
}
extern "C" int ast_check(ast_ctx* c, void* stream) {
    if (!c) return fail(AST_EINVAL, "null argument");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaStreamSynchronize((cudaStream_t)stream));
    CK(cudaGetLastError());
    unsigned e[2];
    const int r = ast_stack_fault_words(c, e);
    if (r != AST_OK) return r;
    if (e[0] | e[1]) {
        cudaMemset(tc_stack_err(c->tc), 0, 2 * sizeof(unsigned));
        return stack_fault(c, e[0], e[1]);
    }
    return AST_OK;
}
extern "C" int ast_uses_stack_kernels(const ast_ctx* c) { return c && c->use_stack ? 1 : 0; }

// the evaluation as SciPy sees it: host float64 in, host float64 out
extern "C" int ast_step_host(ast_ctx* c, const double* h_x, double* h_losses, double* h_grad) {
    if (!c || !h_x || !h_losses || !h_grad) return fail(AST_EINVAL, "null argument");
    if (!c->have_w || !c->have_t) return fail(AST_ESTATE, "ast_set_weights and ast_set_targets must precede a step");
    CK(cudaSetDevice(c->cfg.device));
    if (!c->gexec) {
        const int r = capture_graph(c, &c->gexec, [&](cudaStream_t s) {
            // (Writing losses and gradient straight into the pinned host buffer from the last kernels -- no D2H copy
            // nodes -- was measured: no change, 1492.7 vs 1493.7 steps/s end to end.)
            cudaMemcpyAsync(c->d_x, c->p_x, (size_t)c->R * sizeof(float), cudaMemcpyHostToDevice, s);
            const int rr = ast_step(c, c->d_x, c->d_losses, c->d_grad, s);
            cudaMemcpyAsync(c->p_out, c->d_losses, (size_t)c->B * 4 * sizeof(float), cudaMemcpyDeviceToHost, s);
            cudaMemcpyAsync(c->p_out + 4 * c->B, c->d_grad, (size_t)c->R * sizeof(float), cudaMemcpyDeviceToHost, s);
            if (c->tc) cudaMemcpyAsync(c->p_out + 4 * c->B + c->R, tc_stack_err(c->tc), 2 * sizeof(unsigned), cudaMemcpyDeviceToHost, s);
            return rr;
        });
        if (r != AST_OK) return r;
    }
    for (int i = 0; i < c->R; ++i) c->p_x[i] = (float)h_x[i];
    if (c->tc) { c->p_out[4 * c->B + c->R] = 0.f; c->p_out[4 * c->B + c->R + 1] = 0.f; }
    CK(cudaGraphLaunch(c->gexec, c->hstream));
    CK(cudaStreamSynchronize(c->hstream));
    if (c->tc) {
        const unsigned* ew = reinterpret_cast<const unsigned*>(c->p_out + 4 * c->B + c->R);
        if (ew[0] | ew[1]) return stack_fault(c, ew[0], ew[1]);
    }
    for (int i = 0; i < 4 * c->B; ++i) h_losses[i] = (double)c->p_out[i];
    const float* g = c->p_out + 4 * c->B;
    for (int i = 0; i < c->R; ++i) h_grad[i] = (double)g[i];
    return AST_OK;
}

// device pointers in, device pointers out: one cudaGraphLaunch per evaluation
extern "C" int ast_step_graph(ast_ctx* c, const float* d_x, float* d_losses, float* d_grad, void* stream) {
    if (!c || !d_x || !d_losses || !d_grad) return fail(AST_EINVAL, "null argument");
    if (!c->have_w || !c->have_t) return fail(AST_ESTATE, "ast_set_weights and ast_set_targets must precede a step");
    if (c->gdev && (c->gdev_x != d_x || c->gdev_losses != d_losses || c->gdev_grad != d_grad)) {
        cudaGraphExecDestroy(c->gdev); c->gdev = nullptr;          // pointers are baked into the graph
    }
    if (!c->gdev) {
        CK(cudaStreamSynchronize((cudaStream_t)stream));
        const int r = capture_graph(c, &c->gdev, [&](cudaStream_t s) { return ast_step(c, d_x, d_losses, d_grad, s); });
        if (r != AST_OK) return r;
        c->gdev_x = d_x; c->gdev_losses = d_losses; c->gdev_grad = d_grad;
    }
    CK(cudaGraphLaunch(c->gdev, (cudaStream_t)stream));
    return AST_OK;
}

// ---- phase timing with CUDA events on the launching stream (bench.py roofline) ------------------
// Each phase is its own captured graph (same launches as ast_step), replayed with events between.
extern "C" int ast_step_timed(ast_ctx* c, const float* d_x, float* d_losses, float* d_grad, void* stream, float* ms_out) {
    if (!c || !d_x || !d_losses || !d_grad || !ms_out) return fail(AST_EINVAL, "null argument");
    if (!c->have_w || !c->have_t) return fail(AST_ESTATE, "ast_set_weights and ast_set_targets must precede a step");
    cudaStream_t s = (cudaStream_t)stream;
    if (!c->gphase[0]) {
        CK(cudaStreamSynchronize(s));
        int r = capture_graph(c, &c->gphase[0], [&](cudaStream_t cs) { return run_encoder(c, c->d_x, cs, false); });
        if (r != AST_OK) return r;
        r = capture_graph(c, &c->gphase[1], [&](cudaStream_t cs) {
            run_partials(c, c->d_x, cs, c->fuse_heads);
            return AST_OK;
        });
        if (r != AST_OK) return r;
        r = capture_graph(c, &c->gphase[2], [&](cudaStream_t cs) { return step_backward_impl(c, c->d_losses, c->d_grad, cs, c->fuse_heads); });
        if (r != AST_OK) return r;
    }
    CK(cudaMemcpyAsync(c->d_x, d_x, (size_t)c->R * sizeof(float), cudaMemcpyDeviceToDevice, s));
    CK(cudaEventRecord(c->ev[0], s));
    CK(cudaGraphLaunch(c->gphase[0], s));
    CK(cudaEventRecord(c->ev[1], s));
    CK(cudaGraphLaunch(c->gphase[1], s));
    CK(cudaEventRecord(c->ev[2], s));
    CK(cudaGraphLaunch(c->gphase[2], s));
    CK(cudaEventRecord(c->ev[3], s));
    CK(cudaMemcpyAsync(d_losses, c->d_losses, (size_t)c->B * 4 * sizeof(float), cudaMemcpyDeviceToDevice, s));
    CK(cudaMemcpyAsync(d_grad, c->d_grad, (size_t)c->R * sizeof(float), cudaMemcpyDeviceToDevice, s));
    CK(cudaEventSynchronize(c->ev[3]));
    float enc = 0, gram = 0, bwd = 0, tot = 0;
    CK(cudaEventElapsedTime(&enc, c->ev[0], c->ev[1]));
    CK(cudaEventElapsedTime(&gram, c->ev[1], c->ev[2]));
    CK(cudaEventElapsedTime(&bwd, c->ev[2], c->ev[3]));
    CK(cudaEventElapsedTime(&tot, c->ev[0], c->ev[3]));
    ms_out[0] = enc; ms_out[1] = bwd; ms_out[2] = gram; ms_out[3] = tot;
    return AST_OK;
}

#ifdef AST_EXPERIMENTS
// profiling / fault-injection aids (scripts/timeline_stack.py, tests/test_gpu_loud.py): -DAST_EXPERIMENTS builds only
void tc_set_timeline(unsigned long long* d_buf);
void tc_set_tl_layer(int l);
void tc_set_dbg_skip_publish(int v);
extern "C" int ast_debug_timeline(void* d_buf) { tc_set_timeline((unsigned long long*)d_buf); return AST_OK; }
extern "C" int ast_debug_tl_layer(int l) { tc_set_tl_layer(l); return AST_OK; }
extern "C" int ast_debug_skip_publish(int tile_plus_1) { tc_set_dbg_skip_publish(tile_plus_1); return AST_OK; }
void tc_set_dbg_group_tiles(int v);
extern "C" int ast_debug_group_tiles(int n) { tc_set_dbg_group_tiles(n); return AST_OK; }
#endif
