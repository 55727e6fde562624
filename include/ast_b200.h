/* ast_b200.h -- C ABI of the B200-native audio-style-transfer step.
 *
 * The reference (winlp4ever/audio_style_transfer) has no FFI: its step is an anonymous TensorFlow
 * sub-graph driven from Python.  The boundary this library honours is therefore the Python-level
 * contract of `GatysNet` (methods.py:19-216); each entry point below names the reference lines it
 * replaces.  The ctypes binding a maintainer would add is shown in INTEGRATION.md.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes, no C++ / torch types.
 *   - every function returns 0 on success or a negative AST_E* code; ast_last_error() gives a
 *     thread-local message.  No exception crosses the ABI.
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).
 *   - pointers named d_* are DEVICE pointers owned by the caller; pointers named h_* are HOST
 *     pointers; pointers named p_* may be either (copied with cudaMemcpyDefault).
 *   - an ast_ctx is bound to one device and is not thread-safe; distinct contexts are independent.
 *   - there is NO CPU fallback: if no sm_100 device is present ast_create fails with AST_ECUDA.
 *
 * Layouts (row-major, innermost last)
 *   x, grad      [B, T]            float32, mu-law units (methods.py:49-54)
 *   extracts     [B, T, 128]       float32 (model.py:116, TF's [1, T, 128])
 *   Pc           [B, T, n_cont*cnt_channels]                         (methods.py:58)
 *   Ps / Gn      [B, nb_channels, L, L]  channel-wise ("ours", methods.py:69,73-76)
 *                [B, L, 128, 128]        --gatys                     (methods.py:71,73)
 *   weights      W0 [3,1,128]  b0 [128]  Wd [NL,3,128,128] (k,cin,cout)  bd [NL,128]
 *                Wr [NL,128,128] (cin,cout)  br [NL,128]   -- TF's [1,K,Cin,Cout] order
 *                (nsynth/wavenet/masked.py:136), so a checkpoint reader is a plain copy.
 */
#ifndef AST_B200_H
#define AST_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AST_OK            0
#define AST_EINVAL       -1   /* bad argument / unsupported shape                       */
#define AST_ECUDA        -2   /* CUDA runtime/driver error, or no sm_100 device          */
#define AST_EUNSUPPORTED -3   /* valid in the reference but outside this path (reserved)   */
#define AST_ESTATE       -4   /* call order violated (weights/targets not set)          */

#define AST_PREC_FP32     0   /* CUDA-core fp32 everywhere.  Gradient vs the fp64 oracle's backward evaluated with    */
                              /* this tier's ReLU sign masks: <= 1.8e-5 at every tested size (the <= 1e-4 tier); the  */
                              /* raw distance at 26..30 layers is 3e-4 .. 6e-4, like an fp32 CPU run's: a handful of  */
                              /* sign bits that two roundings of the same forward disagree on   */
#define AST_PREC_BF16     1   /* tcgen05 bf16 tensor cores, fp32 accumulate in TMEM: the <=1e-2 tier.  */
                              /* forward operands are 2-term bf16 splits (hi+lo), backward plain bf16; */
                              /* channel-wise head with 13..32 style layers: mma.sync TF32 (Gram 3xTF32) */

/* ast_config.kernel_path (tensor-core tier): AUTO = the cooperative whole-stack kernels: one launch per direction when a
 * pair has at most one 128-row tile per SM, else one launch per 10-layer stack and group of <= #SMs tiles (a group
 * carries 8 halo tiles per inner side that are computed and discarded).  The other two force the one-kernel-per-layer
 * path (parity tests compare the paths; they compute the same step). */
#define AST_PATH_AUTO          0
#define AST_PATH_PER_LAYER     1   /* one kernel per layer, forward and backward                  */
#define AST_PATH_PER_LAYER_BWD 2   /* whole-stack forward, one kernel per layer in the backward   */

#define AST_PREC_TC3      2   /* accurate tensor-core tier: the dilated / 1x1 convs and their transposes on tcgen05 with  */
                              /* 2-term bf16 operand splits BOTH ways (3 MMAs per product, fp32 accumulate in TMEM); Gatys   */
                              /* Gram and style gradient likewise on tcgen05 from hi / lo bf16 copies of the extracts;      */
                              /* channel-wise Gram and style gradient, normalisation and losses in fp32 on CUDA cores        */

#define AST_MAX_IDS      32
#define AST_CHANNELS    128   /* ae_width, model.py:73 */

typedef struct ast_ctx ast_ctx;

typedef struct ast_config {
    int32_t T;                       /* samples per pair = --batch_size (methods.py:250); T % 128 == 0    */
    int32_t B;                       /* independent content/style pairs (reference: 1, methods.py:50)    */
    int32_t n_layers;                /* NL = 1 + highest extract used (TF prunes the rest)              */
    int32_t n_cont;                  /* --cont_lyrs (methods.py:254)                                     */
    int32_t cont_ids[AST_MAX_IDS];
    int32_t n_style;                 /* resolved style ids (methods.py:60-66); an encoder layer may appear only once
                                      * (29 and 30 are the same layer): AST_EUNSUPPORTED otherwise, see INTEGRATION.md */
    int32_t style_ids[AST_MAX_IDS];
    int32_t gatys;                   /* --gatys (methods.py:261)                                         */
    int32_t nb_channels;             /* --channels (methods.py:258), ours only                           */
    int32_t cnt_channels;            /* --cnt_channels (methods.py:259)                                  */
    int32_t precision;               /* AST_PREC_*                                                       */
    /* Time sharding (new capability, DESIGN.md "multi-GPU"): this context holds the window
     * [win_begin, win_begin + T) of a clip of total_T samples and OWNS rows
     * [own_begin, own_end) of it (global sample indices).  Gram / content sums run over owned rows
     * only; loss means are normalised with total_T.  Unsharded: win_begin = 0, own = [0, T),
     * total_T = T (or all zero, which means the same). */
    int64_t total_T;
    int64_t win_begin;
    int64_t own_begin;
    int64_t own_end;
    int32_t device;                  /* CUDA device ordinal                                              */
    int32_t kernel_path;             /* AST_PATH_* (0 = auto)                                            */
    /* Time shards that EXCHANGE halo rows with their neighbour ranks (BASELINE.json north_star; the reference is
     * single-device, methods.py:48-54,143).  1: the step runs stack by stack (ast_segments / ast_step_forward_seg /
     * ast_step_backward_seg) and between two segments the caller moves AST_HALO_ROWS rows per side of the fp32 stream
     * (forward) / running gradient (backward) between neighbours (ast_segment_buffer); the window then needs only
     * 2 * AST_HALO_ROWS rows beyond the owned range per interior side instead of 2 x the whole receptive field.
     * Tensor-core tier only. */
    int32_t halo_exchange;
} ast_config;

#define AST_HALO_ROWS 1024   /* >= 1023 = sum of the dilations of one 10-layer stack (model.py:98) */

/* GatysNet.__init__/build (methods.py:20-77): allocate workspaces for a fixed (T, B, layer set). */
int  ast_create(const ast_config* cfg, ast_ctx** out);
int  ast_destroy(ast_ctx* ctx);
const char* ast_last_error(void);
/* Library identification: "ast_b200 <version> sm_100a". */
const char* ast_version(void);
/* Bytes of device memory the context holds. */
int64_t ast_workspace_bytes(const ast_ctx* ctx);

/* GatysNet.load_model (methods.py:79-84).  Copies; host or device pointers. */
int  ast_set_weights(ast_ctx* ctx, const float* p_W0, const float* p_b0,
                     const float* p_Wd, const float* p_bd,
                     const float* p_Wr, const float* p_br, void* stream);

/* define_loss (methods.py:113-125): targets and loss weights baked for the following steps.
 * gamma != 0 turns on the STFT regulariser (methods.py:121-123; needs total_T >= 1024, and when time-sharded a
 * window that holds every 1024-sample frame touching an owned sample -> else AST_EINVAL).  With gamma == 0 it is
 * not evaluated and losses[3] is 0. */
int  ast_set_targets(ast_ctx* ctx, const float* p_Pc, const float* p_Ps,
                     double lambd, double gamma, void* stream);

/* get_embeds (methods.py:86-95), forward only.  Either output may be NULL. */
int  ast_embeds(ast_ctx* ctx, const float* d_x, float* d_Ec, float* d_Gn, void* stream);

/* One function evaluation of the L-BFGS-B loop (methods.py:133-137,167):
 * d_losses [B,4] = {total, content, style, regulariser}, d_grad [B,T] = d total / d x.
 * Asynchronous on `stream`, no host sync, CUDA-graph capturable.
 * A context holds ONE evaluation's activations, masks and hand-off flags: evaluations of the same context (ast_step*,
 * ast_embeds, ast_minimize; ast_step_host runs on the context's private stream) must not overlap on the GPU --
 * order them on one stream or synchronise between streams; overlapping ones fail loudly (AST_ECUDA from ast_check /
 * ast_step_host: hand-off timed out).  Different contexts are independent. */
int  ast_step(ast_ctx* ctx, const float* d_x, float* d_losses, float* d_grad, void* stream);

/* ast_step replayed from a captured CUDA graph (north_star (e)): captured on first use for this
 * (d_x, d_losses, d_grad) pointer triple, re-captured when the pointers change; one
 * cudaGraphLaunch on `stream` per call. */
int  ast_step_graph(ast_ctx* ctx, const float* d_x, float* d_losses, float* d_grad, void* stream);

/* The same evaluation through HOST buffers, as SciPy's fun(x)->(f,g) sees it (float64 each way,
 * methods.py:167 / ScipyOptimizerInterface): H2D of x, a replay of the captured CUDA graph of
 * ast_step, D2H of losses and gradient, stream synchronised before return. */
int  ast_step_host(ast_ctx* ctx, const double* h_x, double* h_losses, double* h_grad);

/* Two-phase step for time-sharded clips.  Phase 1 runs the forward and leaves the raw (un-normalised)
 * Gram partials, content squared-error partials and (gamma != 0) the raw STFT-L1 sum of the owned frames in one contiguous fp32 device
 * buffer (ast_partials) that the caller all-reduces (sum) across ranks; phase 2 normalises, forms
 * the losses and runs the backward.  ast_step == forward + backward with no reduction between. */
int  ast_step_forward(ast_ctx* ctx, const float* d_x, void* stream);
int  ast_partials(ast_ctx* ctx, float** d_buf, int64_t* count);
int  ast_step_backward(ast_ctx* ctx, float* d_losses, float* d_grad, void* stream);

/* The same two phases stack by stack, for contexts created with halo_exchange = 1 (n = ast_segments(ctx), 3 for
 * 21..30 layers).  Forward: seg = 0 .. n-1; after segment seg < n-1 the fp32 stream at the segment's end is exact on
 * [own_begin - AST_HALO_ROWS, own_end + AST_HALO_ROWS) and the caller overwrites the AST_HALO_ROWS rows outside that range on
 * each interior side of the window with the neighbour rank's rows (ast_segment_buffer(ctx, 0, seg): [B][T][128] fp32);
 * the last segment also forms the partial sums (then: all-reduce ast_partials).  Backward: seg = n-1 .. 0; segment n-1
 * starts with the heads; after segment seg > 0 the running gradient (ast_segment_buffer(ctx, 1, seg)) is exact on the
 * owned rows and the caller overwrites the AST_HALO_ROWS rows on each interior side of the owned range with the neighbour's;
 * segment 0 ends with the gradient in d_grad.  Unsharded contexts report their own segment count (1 unless the clip
 * has more tiles than SMs) and may be driven the same way without any exchange. */
int  ast_segments(const ast_ctx* ctx);
int  ast_step_forward_seg(ast_ctx* ctx, const float* d_x, int32_t seg, void* stream);
int  ast_step_backward_seg(ast_ctx* ctx, int32_t seg, float* d_losses, float* d_grad, void* stream);
int  ast_segment_buffer(ast_ctx* ctx, int32_t backward, int32_t seg, float** d_buf);

/* Peer-mapped halos: instead of moving the halo rows between two segments (ast_segment_buffer + send / receive), a rank
 * maps its neighbours' buffers into its own address space once, and the stack kernels do the exchange themselves: the
 * kernel that resumes a segment reads the neighbour's rows over NVLink (ld.relaxed.sys on the mapped pointer) after the
 * neighbour's previous segment has signalled completion into this rank's memory (st.release.sys from the last CTA of
 * that kernel / ld.acquire.sys polls, bounded: 30 s).  With every interior side attached, ast_step_forward /
 * ast_step_backward (or the _seg calls without any copy between them) evaluate the shard; the all-reduce of
 * ast_partials stays the caller's.  All ranks must start the same evaluations in the same order.
 *   ast_p2p_export : AST_P2P_HANDLE_BYTES bytes (four cudaIpcMemHandle_t) to hand to the neighbour processes;
 *   ast_p2p_attach : side 0 = the rank that owns the samples before this window, 1 = after; peer_win_begin / peer_T as
 *                    in the neighbour's ast_config;  ast_p2p_attach_local: the neighbour is a context of this process;
 *   ast_p2p_detach : back to copies (every rank of the job, or none).
 * No reference counterpart (single device, methods.py:48-54,143).  B = 1 only. */
#define AST_P2P_HANDLE_BYTES 256
int  ast_p2p_export(ast_ctx* ctx, void* h_handles);
int  ast_p2p_attach(ast_ctx* ctx, int32_t side, const void* h_handles, int64_t peer_win_begin, int32_t peer_T);
int  ast_p2p_attach_local(ast_ctx* ctx, int32_t side, ast_ctx* peer);
int  ast_p2p_detach(ast_ctx* ctx);   /* forget the neighbours again: halo rows are exchanged by copies as before */

/* cfg.extracts[layer] of the last forward (model.py:116), [B,T,128] fp32 -> d_out.  Debug/parity.
 * After ast_embeds every layer is available.  After ast_step* in the tensor-core tier only the layers a head reads
 * are: the whole-stack kernels keep the residual stream on chip (TMEM) and write the content extracts (and, for the
 * channel-wise Gram, the style extracts) only; any other layer returns AST_ESTATE ("not retained"). */
int  ast_get_extract(ast_ctx* ctx, int32_t layer, float* d_out, void* stream);

/* The ReLU sign masks the last forward left for the backward (TF's ReluGrad: mask = output > 0; reference
 * model.py:99,107 + tf.gradients, methods.py:133-137), one bit per element, bit (c % 32) of word c / 32:
 *   d_hmask [n_layers][B*T][4] uint32: h_l > 0 (input of layer l's first ReLU, l = 0 .. n_layers-1),
 *   d_zmask [n_layers][B*T][4] uint32: z_l > 0 (output of layer l's dilated conv).
 * Debug/parity: lets a test impose the device's masks on the oracle's backward, so that the gradient comparison is
 * not dominated by the handful of sign bits that legitimately differ between two roundings of the same forward. */
int  ast_get_masks(ast_ctx* ctx, uint32_t* d_hmask, uint32_t* d_zmask, void* stream);

/* Asynchronous error check.  ast_step / ast_step_graph / ast_step_forward / ast_step_backward only enqueue work; the
 * whole-stack kernels hand rows between SMs through flags with BOUNDED waits, and a wait that expires (profiler, debugger,
 * heavy preemption) invalidates that evaluation.  ast_check synchronises `stream` and returns AST_ECUDA (sticky until this
 * call, which clears it) if any evaluation since the last check timed out, or if CUDA reported an error.
 * ast_step_host and ast_minimize run the same check themselves before they return. */
int  ast_check(ast_ctx* ctx, void* stream);
/* 1 if the context runs the cooperative whole-stack kernels (one launch per direction), 0 if one kernel per layer. */
int  ast_uses_stack_kernels(const ast_ctx* ctx);

/* Number of kernels one ast_step launches (for bench.py's gpu_launches). */
int  ast_launches_per_step(const ast_ctx* ctx);

/* One step as three captured graphs (encoder forward | Gram partials + content SSE | heads + backward)
 * replayed on `stream` with CUDA events between them.
 * ms_out[0] = encoder forward, [1] = heads + backward, [2] = Gram partials, [3] = total. */
int  ast_step_timed(ast_ctx* ctx, const float* d_x, float* d_losses, float* d_grad, void* stream,
                    float* ms_out);

/* The gamma term alone -- reg = mean(abs(Re) + abs(Im)) of STFT(inv_mu_law(x), 1024, 512) and d reg / d x -- for ONE clip,
 * with NO allocation of any kind.  This entry point has no counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid it).  It calls no cudaMalloc / cudaFree / allocating library routine: the scratch the context-based
 * path keeps in its workspace lives in one file-scope device array of 404 KB that every GPU loading the library holds
 * until exit, which is why T is limited to AST_NOALLOC_MAX_T samples (larger -> AST_EINVAL; T % 128 == 0, T >= 1024).
 * d_x [T], d_grad [T] (overwritten), d_reg [1] are device pointers on the CURRENT device; asynchronous on `stream`.
 * The same numbers come out of ast_set_targets(gamma != 0) + ast_step as losses[3] and (grad(gamma) - grad(0)) / gamma.
 * All calls on a device share the scratch: calls that overlap on the GPU (different streams) race -- serialise them.
 * The first call on a device does a one-time synchronous set-up (8 KB table) and must be outside stream capture
 * (AST_ESTATE otherwise); every later call only enqueues one memset and four kernels and is capturable. */
#define AST_NOALLOC_MAX_T 65536
int  ast_stft_reg_noalloc(const float* d_x, int32_t T, float* d_reg, float* d_grad, void* stream);

/* ---- the optimiser around the step, resident on the device ---------------------------------
 * Reference: `tf.contrib.opt.ScipyOptimizerInterface(loss, method='L-BFGS-B', options={'maxiter': 100})`
 * (methods.py:133-137), run once per epoch from the variable's current value, stopping after an epoch with fewer than
 * 50 evaluations (methods.py:164-181).  SciPy's L-BFGS-B (no bounds: limited-memory BFGS direction + the MINPACK-2
 * dcsrch line search; defaults m 10, factr 1e7, pgtol 1e-5, maxls 20) is restated as a state machine that lives in
 * device memory: one `ast_opt_advance` consumes the (losses, gradient) evaluated at the point `ast_opt_buffers`
 * names and leaves the next evaluation point there.  No host read-back is needed between evaluations; B pairs advance
 * independently (one thread-block cluster each).  The SciPy-driven `ast_step_host` path stays the parity reference. */
typedef struct ast_opt ast_opt;

typedef struct ast_opt_config {
    int32_t T;             /* variables per pair (= samples)                                                */
    int32_t B;             /* independent problems                                                          */
    int32_t m;             /* history pairs (SciPy maxcor), 0 = 10, <= 10                                   */
    int32_t maxiter;       /* iterations per epoch (methods.py:137), 0 = 100                                */
    int32_t epochs;        /* fresh runs from the current iterate (methods.py:164), 0 = 1                   */
    int32_t min_evals;     /* stop after an epoch with fewer evaluations (methods.py:180-181: 50)           */
    int32_t maxls;         /* line-search evaluations before the memory is dropped (SciPy maxls), 0 = 20    */
    int32_t max_log;       /* rows of the per-evaluation losses log [B][max_log][4]                         */
    int32_t keep_epoch_x;  /* keep the iterate at the end of every epoch (the reference writes ep-N.wav)    */
    int32_t device;
    double  factr;         /* 0 = 1e7 (ftol = factr * eps)                                                  */
    double  pgtol;         /* < 0 = 1e-5                                                                    */
} ast_opt_config;

int  ast_opt_create(const ast_opt_config* cfg, ast_opt** out);
int  ast_opt_destroy(ast_opt* opt);
int64_t ast_opt_workspace_bytes(const ast_opt* opt);
/* Start from x0 [B,T] float64 (host or device pointer).  Asynchronous on `stream`. */
int  ast_opt_reset(ast_opt* opt, const double* p_x0, void* stream);
/* Device buffers of the machine: evaluate the step at *d_x [B,T] and put its results in *d_losses [B,4],
 * *d_grad [B,T] (any pointer may be NULL). */
int  ast_opt_buffers(ast_opt* opt, float** d_x, float** d_losses, float** d_grad);
/* One transition (one kernel launch, asynchronous, CUDA-graph capturable). */
int  ast_opt_advance(ast_opt* opt, void* stream);
/* Synchronising read-back: h_status [B][8] = {phase (0 epoch start, 1 line search, 2 done), epochs finished,
 * iteration, evaluations this epoch, evaluations in total, why the last epoch ended (1 maxiter, 2 pgtol, 3 ftol,
 * 4 abnormal line search), history pairs held, line-search evaluations}; h_f [B] = loss at the iterate (may be NULL). */
int  ast_opt_status(ast_opt* opt, int32_t* h_status, double* h_f, void* stream);
/* The iterate x_k [B,T] as float64 (synchronises). */
int  ast_opt_result(ast_opt* opt, double* h_x, void* stream);
/* Per-epoch records [B][epochs]: evaluations used, end reason, and (keep_epoch_x) the iterate [B][epochs][T] fp32. */
int  ast_opt_epochs(ast_opt* opt, int32_t* h_nfev, int32_t* h_end, float* h_x, void* stream);
/* The losses {total, content, style, reg} of every evaluation, [B][max_log][4] (methods.py:147-157 prints them). */
int  ast_opt_log(ast_opt* opt, float* h_log, void* stream);
/* Whole optimisation: replays [ast_step at the machine's point, ast_opt_advance] as one CUDA graph until every pair
 * has stopped or max_evals (0 = no cap) evaluations were enqueued; status is read back every `chunk` (0 = 4)
 * evaluations, one chunk behind, so the device never waits for the host.  ctx must have B and T of opt, weights and
 * targets set.  Synchronises before returning. */
int  ast_minimize(ast_ctx* ctx, ast_opt* opt, int64_t max_evals, int32_t chunk, int64_t* evals_enqueued);

#ifdef __cplusplus
}
#endif
#endif /* AST_B200_H */
