// common.cuh -- shared declarations of the B200-native style-transfer step (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

#define AST_C 128           // ae_width (model.py:73 of the reference)
#define AST_MW 4            // 32-bit words of one row's 128-bit sign mask

// kernels launched since the counter was last reset (bench.py gpu_launches); bumped by every launcher
extern thread_local int ast_launch_counter;

// ---- kernel argument blocks ---------------------------------------------------------------
// One "layer" launch = two chained GEMMs over a tile of rows (see DESIGN.md "kernels"):
//   forward  (model.py:99-114):  z = conv3_d(relu(h)) + bd ; h' = h + relu(z) Wr + br
//   backward :  g += conv3_d^T(dz_l) * [h_l > 0] (+ head terms) ; dz_{l-1} = (g Wr_{l-1}^T) * [z_{l-1} > 0]
struct LayerArgsF32 {
    const float* A1;        // fwd: H[l] (ReLU applied on load); bwd: dz_l.  [R,128].  NULL: skip GEMM1
    const float* W1;        // [3][128 k][128 n] fp32 (fwd: Wd[l]; bwd: Wd[l]^T per tap)
    const float* bias1;     // fwd: bd[l]; bwd: NULL
    const float* W2;        // [128 k][128 n] (fwd: Wr[l]; bwd: Wr[l-1]^T).  NULL: skip GEMM2
    const float* bias2;     // fwd: br[l]
    const float* Hin;       // fwd: H[l] (residual)
    float*       Hout;      // fwd: H[l+1]
    uint32_t*    zmask_out; // fwd: [R,4]
    uint32_t*    hmask_out; // fwd: [R,4]  sign mask of H[l+1]
    float*       g;         // bwd: running d h, in place [R,128]
    const uint32_t* hmask_in;  // bwd: [h_l > 0]
    const float* dF;        // bwd: head gradient for extracts[l-1] = H[l], or NULL
    const uint32_t* zmask_in;  // bwd: [z_{l-1} > 0]
    float*       dz_out;    // bwd: dz_{l-1}
    int T;                  // rows per pair (shifts never cross a pair boundary)
    int R;                  // total rows = B*T
    int shift;              // tap k reads row t + (k-1)*shift  (fwd: +d, bwd: -d)
    int g_init;             // bwd: 1 = treat incoming g as zero (top of the stack)
};

// ---- launchers implemented in kernels_f32.cu ------------------------------------------------
void launch_startconv_fwd(const float* x, const float* W0, const float* b0, float* H0, uint32_t* hmask0,
                          int T, int R, cudaStream_t s);
void launch_startconv_bwd(const float* g, const float* W0, float* scratch3R, float* dx, int T, int R, cudaStream_t s);
// second half only: the dots p[3][R] are already in scratch3R (written by the whole-stack backward kernel)
void launch_startconv_bwd_combine(const float* scratch3R, float* dx, int T, int R, cudaStream_t s);
void launch_layer_f32(const LayerArgsF32& a, bool bwd, cudaStream_t s);
int  setup_f32_kernels();

// Gram partials.  ours: part[b][chunk][c][l][m]; gatys: part[b][l][chunk][128][128].  Rows outside
// [own_lo, own_hi) (row index within the pair) do not contribute (time sharding).
void launch_gram_ours(const float* const* d_F, int L, int T, int B, int own_lo, int own_hi,
                      float* part, int nchunk, int chunk_rows, cudaStream_t s);
int  gram_ours_max_small_L();     // L up to which launch_gram_ours uses the one-stack kernel
// tensor-core tier, 13 <= L <= 32 (gram_ours_tc.cu): per-channel 32 x 32 Gram by mma.sync with 3xTF32 operands
int  gram_ours_tc_supported(int L);
int  style_grad_ours_tc_supported(int L);
// Per-layer base pointers of the style extracts / their gradient buffers, passed BY VALUE (kernel parameter space ->
// uniform constant loads): indexing a pointer array in global memory inside the streaming loops put a dependent global
// load in front of every address (ncu, round 2: 26 % of the style-gradient kernel's stall samples).
struct OursPtrs { const float* F[32]; float* dF[32]; };
void launch_style_grad_ours_tc(const OursPtrs& ptrs, const float* Ssym, int L, int T, int B, cudaStream_t s);
void launch_gram_ours_tc(const OursPtrs& ptrs, int L, int T, int B, int own_lo, int own_hi,
                         float* part, int nchunk, int chunk_rows, cudaStream_t s);
// The same two fed by TMA from transposed fp16 extract copies Ht [R/32][L][128][32 rows] that the whole-stack forward kernel
// writes (gram_ours_tma.cu).  OursTma::map holds the CUtensorMap over Ht; chunk_rows / own_lo / own_hi / T multiples of 32.
struct OursTma { alignas(64) unsigned char map[128]; };
int  ours_tma_supported(int L);
size_t ours_tma_bytes(int L, int R);                              // size of Ht
int  ours_tma_make_map(OursTma* o, void* Ht, int L, int R);       // 0 = ok
int  ours_tma_chunk_rows(int T, int target);
void launch_gram_ours_tma(const OursTma& o, int L, int T, int B, int own_lo, int own_hi, float* part, int nchunk,
                          int chunk_rows, cudaStream_t s);
void launch_style_grad_ours_tma(const OursTma& o, const OursPtrs& ptrs, const float* Ssym, int L, int T, int B, cudaStream_t s);
void launch_gram_gatys(const float* const* d_F, int L, int T, int B, int own_lo, int own_hi,
                       float* part, int nchunk, int chunk_rows, cudaStream_t s);
// deterministic fixed-order reduction of partials -> G (fp32).  nmat matrices of msize elements;
// partial p of matrix m is at part[m_outer(m) ... ] -- see kernels_f32.cu for the two layouts.
void launch_gram_reduce(const float* part, float* G, int gatys, int L, int B, int nchunk, cudaStream_t s);

// content squared error over owned rows -> sse[b] (double accumulate, fp32 out)
// sse == NULL: partials only (the heads kernel sums them)
void launch_content_sse(const float* const* d_Hc, int n_cont, int cnt_ch, const float* Pc,
                        int T, int B, int own_lo, int own_hi, float* part, int nblk, float* sse, cudaStream_t s);

// heads: normalise (methods.py:74), style MSE (methods.py:118-119), dG, Ssym = dG + dG^T, losses.
struct HeadArgs {
    const float* G;       // raw Gram [B][nmat][msize]
    const float* Ps;      // targets  [B][nmat_used][msize]
    float* Gn;            // out, may be NULL: normalised Gram [B][nmat][msize]
    float* Ssym;          // out, may be NULL: [B][nmat][msize]
    __nv_bfloat16* Ssym_bf16;   // out, may be NULL: bf16 image of Ssym (B operand of the tensor-core style gradient)
    __nv_bfloat16* Ssym_bf16_lo; // out, may be NULL: bf16(Ssym - float(Ssym_bf16)) (accurate tier: hi / lo operand pairs)
    float* ls_part;       // [B][nmat] per-matrix sum of squared differences
    const float* sse;     // [B] content squared error (already all-reduced when sharded)
    float* losses;        // out [B][4]
    int nmat, nmat_used, n;   // n = matrix side (L for ours, 128 for gatys)
    int B;
    float coef_dGn;       // lambd * 1e3 * 2 / N_s
    float ls_scale;       // 1e3 / N_s
    float lc_scale;       // 10 / N_c
    float lambd;
    // single-GPU step only (all NULL in the two-phase, all-reducible path):
    const float* gram_part;  // Gatys split-K partials [B][nmat][nchunk][n*n]: summed here instead of by gram_reduce_kernel
    int nchunk;
    const double* sse_part;  // [B][sse_nblk] content squared-error partials: summed here instead of by content_sse_final
    int sse_nblk;
    unsigned* done;          // [B] arrival counters (zero between launches): the last CTA of a pair writes the losses
    const float* reg_raw; // [B] raw STFT-L1 sum (stft_reg.cu; all-reduced when sharded); read only when gamma != 0
    float reg_scale;      // 1 / (frames * 513)
    float gamma;
};
void launch_heads(const HeadArgs& a, cudaStream_t s);

// head gradients -> dF buffers
void launch_style_grad_ours(const float* const* d_F, float* const* d_dF, const float* Ssym, int L, int T, int B,
                            cudaStream_t s);
void launch_style_grad_gatys(const float* const* d_F, float* const* d_dF, const float* Ssym, int L, int T, int B,
                             cudaStream_t s);
// dF[t, j] (+)= coef * (H[t, j] - Pc[t, slot*cnt + j]) for j < cnt ; accumulate=0 also zero-fills j >= cnt
void launch_content_grad(const float* H, const float* Pc, float* dF, int slot, int n_cont, int cnt_ch, float coef,
                         int accumulate, int R, cudaStream_t s);
void launch_content_gather(const float* const* d_Hc, int n_cont, int cnt_ch, float* Ec, int R, cudaStream_t s);

// Launch with programmatic stream serialization (PDL): the kernel MUST call pdl_wait() (ptx.cuh) before it
// touches anything an earlier kernel in the stream wrote or reads.
template <typename Kern, typename Args>
static inline cudaError_t launch_pdl(Kern kernel, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const Args& args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, args);
}

// ---- gamma term (stft_reg.cu): reg = mean(abs(Re) + abs(Im)) of STFT(inv_mu_law(x)), methods.py:121-123 -------
// forward: frames [f_lo, f_lo + nf) -> sign(Re), sign(Im) per bin (int8), d a / d x per sample, and
// raw[b] = sum over the frames in [own_f_lo, own_f_hi).  grad: grad += coef * d(raw sum)/dx over the evaluated frames.
void launch_stft_l1_fwd(const float* x, const float2* tw, int T, int B, int win_begin, int f_lo, int nf,
                        int own_f_lo, int own_f_hi, signed char* sgn, float* fpart, float* dadx, float* raw, cudaStream_t s);
void launch_stft_l1_grad(const float* dadx, const float2* tw, const signed char* sgn, int T, int B, int win_begin,
                         int f_lo, int nf, float coef, float* grad, cudaStream_t s);

#ifdef __CUDACC__
// ---- double-float arithmetic on the fp32 pipe ------------------------------------------------------------------
struct df { float hi, lo; };      // an unevaluated sum hi + lo of two floats
// Measured on B200 (scripts/opt_profile.py stamps + ncu of lbfgs_advance_kernel): F2F.F64.F32 -- converting a loaded
// float to double -- runs at ~2 conversions per clock per SM, so a pass that converts every element is bound by the
// conversions.  Reductions that want more than fp32 therefore keep every sum as an unevaluated (hi, lo) pair of floats
// (error-free product via fma, TwoSum for the add: ~2^-46 relative) and convert ONCE per reduced value.
__device__ __forceinline__ void df_acc_prod(df& a, float x, float y) {       // a += x * y
    const float p = __fmul_rn(x, y);
    const float e = __fmaf_rn(x, y, -p);
    const float t = __fadd_rn(a.hi, p);
    const float bb = __fsub_rn(t, a.hi);
    const float err = __fadd_rn(__fsub_rn(a.hi, __fsub_rn(t, bb)), __fsub_rn(p, bb));
    a.hi = t;
    a.lo = __fadd_rn(a.lo, __fadd_rn(err, e));
}
__device__ __forceinline__ df df_add(df a, df b) {
    const float t = __fadd_rn(a.hi, b.hi);
    const float bb = __fsub_rn(t, a.hi);
    const float err = __fadd_rn(__fsub_rn(a.hi, __fsub_rn(t, bb)), __fsub_rn(b.hi, bb));
    df r;
    r.hi = t;
    r.lo = __fadd_rn(__fadd_rn(a.lo, b.lo), err);
    return r;
}
__device__ __forceinline__ double df_to_double(df a) { return (double)a.hi + (double)a.lo; }
#endif
