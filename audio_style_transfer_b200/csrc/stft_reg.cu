// stft_reg.cu -- the gamma term of define_loss (reference methods.py:121-123, utils.py:92-104), sm_100a.
//
//   a   = inv_mu_law(x)                                   utils.py:99-104 (with the reference's own abs / sign)
//   Y   = stft(a, frame 1024, hop 512, periodic hann)     methods.py:122 (tf.contrib.signal.stft defaults)
//   reg = mean(abs(Re Y) + abs(Im Y))                     methods.py:123, abs(v) = max(v,1e-12) + max(0,-v)
//
// 31 frames x 513 bins at T = 16384: 16 M multiply-adds each way -- a side computation next to the encoder's
// 130 GFLOP, so it is a direct DFT against a 1024-entry twiddle table in shared memory (no FFT butterflies, no
// tensor cores): fp32 CUDA cores, one thread per bin (forward) or per sample (adjoint).  The adjoint needs only
// sign(Re), sign(Im) per bin, which the forward leaves behind as int8.
//
// Time sharding: a context computes the frames that touch its owned samples; the raw sum counts the frames that
// START in the owned range (partials buffer, all-reduced with the Gram partials), the gradient is complete on the
// owned samples.
#include "common.cuh"
#include <math.h>

#define REG_FRAME 1024
#define REG_HOP 512
#define REG_BINS 513
#define REG_EPS 1e-12f
#define REG_BCHUNK 4                 // CTAs per frame in the forward
#define REG_BPC 129                  // bins per CTA (4 * 129 >= 513)

__device__ __forceinline__ float ref_abs(float v) { return fmaxf(v, REG_EPS) + fmaxf(0.f, -v); }          // utils.py:92-93
__device__ __forceinline__ float ref_abs_grad(float v) { return (v >= REG_EPS ? 1.f : 0.f) - (v < 0.f ? 1.f : 0.f); }

// utils.py:99-104.  Returns a; *dadx = d a / d x as TF's autodiff gives it through the custom abs / sign.
__device__ __forceinline__ float inv_mu_law_dev(float x, float* dadx) {
    const float v = (x + 0.5f) * 2.0f / 256.0f;
    const float A = ref_abs(v), dA = ref_abs_grad(v);
    const bool tiny = fabsf(v) <= REG_EPS;
    const float s = (tiny ? 0.f : v) / A;
    const float ds = ((tiny ? 0.f : 1.f) - s * dA) / A;
    const float LN256 = 5.545177444479562f;
    const float Pm1 = expm1f(A * LN256);                               // 256^A - 1
    float a = s * (1.0f / 255.0f) * Pm1;
    float d = (ds * Pm1 + s * (Pm1 + 1.0f) * LN256 * dA) * (1.0f / 255.0f) * (2.0f / 256.0f);
    if (x == 0.f) { a = x; d = 1.f; }                                  // utils.py:103 where(equal(x, 0), x, out)
    if (dadx) *dadx = d;
    return a;
}

// tw[j] = (cos, sin)(2 pi j / 1024), from the host in double precision
__global__ void __launch_bounds__(160) stft_l1_fwd_kernel(const float* __restrict__ x, const float2* __restrict__ tw,
                                                          int T, int win_begin, int f_lo, int nf,
                                                          signed char* __restrict__ sgn, float* __restrict__ fpart,
                                                          float* __restrict__ dadx_out) {
    __shared__ float aw[REG_FRAME];
    __shared__ float2 tws[REG_FRAME];
    __shared__ float red[5];
    const int fi = blockIdx.x, chunk = blockIdx.y, b = blockIdx.z;
    const int f = f_lo + fi;
    const float* xb = x + (size_t)b * T + ((size_t)f * REG_HOP - win_begin);
    // d a / d x of the frame's first hop (the last evaluated frame: of all its samples) for the adjoint kernel
    float* dd = (chunk == 0) ? dadx_out + (size_t)b * T + ((size_t)f * REG_HOP - win_begin) : nullptr;
    const int dd_n = (fi == nf - 1) ? REG_FRAME : REG_HOP;
    for (int m = threadIdx.x; m < REG_FRAME; m += blockDim.x) {
        const float2 t = tw[m];
        tws[m] = t;
        float d;
        aw[m] = inv_mu_law_dev(xb[m], &d) * (0.5f - 0.5f * t.x);   // periodic hann: 0.5 - 0.5 cos(2 pi m / N)
        if (dd && m < dd_n) dd[m] = d;
    }
    __syncthreads();
    const int k = chunk * REG_BPC + threadIdx.x;
    float v = 0.f;
    if (threadIdx.x < REG_BPC && k < REG_BINS) {
        // lane-skewed walk over m so that neighbouring bins do not hit the same twiddle bank at every step
        const int lane = threadIdx.x & 31;
        int m = lane, j = (k * lane) & (REG_FRAME - 1);
        float re0 = 0.f, re1 = 0.f, im0 = 0.f, im1 = 0.f;
#pragma unroll 4
        for (int it = 0; it < REG_FRAME; it += 2) {
            const float a0 = aw[m];
            const float2 t0 = tws[j];
            re0 = fmaf(a0, t0.x, re0); im0 = fmaf(-a0, t0.y, im0);
            m = (m + 1) & (REG_FRAME - 1); j = (j + k) & (REG_FRAME - 1);
            const float a1 = aw[m];
            const float2 t1 = tws[j];
            re1 = fmaf(a1, t1.x, re1); im1 = fmaf(-a1, t1.y, im1);
            m = (m + 1) & (REG_FRAME - 1); j = (j + k) & (REG_FRAME - 1);
        }
        const float re = re0 + re1, im = im0 + im1;
        v = ref_abs(re) + ref_abs(im);
        signed char* sg = sgn + (((size_t)b * nf + fi) * REG_BINS + k) * 2;
        sg[0] = (signed char)ref_abs_grad(re);
        sg[1] = (signed char)ref_abs_grad(im);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0)
        fpart[((size_t)b * nf + fi) * REG_BCHUNK + chunk] = ((red[0] + red[1]) + (red[2] + red[3])) + red[4];
}

// raw[b] = sum over the frames that START in the owned range, fixed order, double accumulate
__global__ void stft_l1_reduce_kernel(const float* __restrict__ fpart, int nf, int f_lo, int own_f_lo, int own_f_hi,
                                      float* __restrict__ raw) {
    const int b = blockIdx.x;
    double s = 0.0;
    for (int i = threadIdx.x; i < nf * REG_BCHUNK; i += 32) {
        const int f = f_lo + i / REG_BCHUNK;
        if (f >= own_f_lo && f < own_f_hi) s += (double)fpart[(size_t)b * nf * REG_BCHUNK + i];
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (threadIdx.x == 0) raw[b] = (float)s;
}

// grad[n] += coef * dadx(x[n]) * sum over the (at most two) frames covering n of
//            w[m] * sum_k ( sRe[k] cos(2 pi k m / N) - sIm[k] sin(2 pi k m / N) ),  m = n - 512 f
__global__ void __launch_bounds__(128) stft_l1_grad_kernel(const float* __restrict__ dadx, const float2* __restrict__ tw,
                                                           const signed char* __restrict__ sgn, int T, int win_begin,
                                                           int f_lo, int nf, float coef, float* __restrict__ grad) {
    __shared__ float2 tws[REG_FRAME];
    __shared__ float2 sg[2][REG_BINS + 1];
    const int b = blockIdx.y;
    const int n_loc = blockIdx.x * 128 + threadIdx.x;
    const int n = win_begin + n_loc;
    const int f1 = n >> 9, f0 = f1 - 1;                   // the same for the whole CTA (128 | 512)
    const bool have[2] = {f0 >= f_lo && f0 < f_lo + nf, f1 >= f_lo && f1 < f_lo + nf};
    if (!have[0] && !have[1]) return;
    for (int m = threadIdx.x; m < REG_FRAME; m += 128) tws[m] = tw[m];
    for (int q = 0; q < 2; ++q) {
        const int f = q ? f1 : f0;
        for (int k = threadIdx.x; k < REG_BINS; k += 128) {
            float2 v = make_float2(0.f, 0.f);
            if (have[q]) {
                const signed char* p = sgn + (((size_t)b * nf + (f - f_lo)) * REG_BINS + k) * 2;
                v = make_float2((float)p[0], (float)p[1]);
            }
            sg[q][k] = v;
        }
    }
    __syncthreads();
    const int m1 = n & (REG_HOP - 1), m0 = m1 + REG_HOP;
    // lane-skewed walk over k (see the forward); k' wraps at 513
    int k = threadIdx.x & 31;
    int j1 = (k * m1) & (REG_FRAME - 1);
    float acc0 = 0.f, acc1 = 0.f;
#pragma unroll 4
    for (int it = 0; it < REG_BINS; ++it) {
        const int j0 = j1 ^ ((k & 1) << 9);               // k * (m1 + 512) = k m1 + 512 k (mod 1024)
        const float2 t1 = tws[j1], t0 = tws[j0];
        const float2 s1 = sg[1][k], s0 = sg[0][k];
        acc1 = fmaf(s1.x, t1.x, fmaf(-s1.y, t1.y, acc1));
        acc0 = fmaf(s0.x, t0.x, fmaf(-s0.y, t0.y, acc0));
        ++k; j1 = (j1 + m1) & (REG_FRAME - 1);
        if (k == REG_BINS) { k = 0; j1 = 0; }
    }
    const float w1 = 0.5f - 0.5f * tws[m1].x, w0 = 0.5f - 0.5f * tws[m0].x;
    grad[(size_t)b * T + n_loc] += coef * dadx[(size_t)b * T + n_loc] * (w1 * acc1 + w0 * acc0);
}

void launch_stft_l1_fwd(const float* x, const float2* tw, int T, int B, int win_begin, int f_lo, int nf,
                        int own_f_lo, int own_f_hi, signed char* sgn, float* fpart, float* dadx, float* raw, cudaStream_t s) {
    stft_l1_fwd_kernel<<<dim3(nf, REG_BCHUNK, B), 160, 0, s>>>(x, tw, T, win_begin, f_lo, nf, sgn, fpart, dadx);
    stft_l1_reduce_kernel<<<B, 32, 0, s>>>(fpart, nf, f_lo, own_f_lo, own_f_hi, raw);
    ast_launch_counter += 2;
}
void launch_stft_l1_grad(const float* dadx, const float2* tw, const signed char* sgn, int T, int B, int win_begin,
                         int f_lo, int nf, float coef, float* grad, cudaStream_t s) {
    stft_l1_grad_kernel<<<dim3(T / 128, B), 128, 0, s>>>(dadx, tw, sgn, T, win_begin, f_lo, nf, coef, grad);
    ast_launch_counter += 1;
}

// =====================================================================================================================
// ast_stft_reg_noalloc -- the gamma term (value and gradient) of ONE clip with NO allocation of any kind.
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  The scratch that launch_stft_l1_fwd /
// launch_stft_l1_grad take from the context's workspace (twiddles, sign(Re)/sign(Im) per bin, d a / d x per sample,
// per-CTA partials) lives here in ONE file-scope device array of fixed size: every GPU that loads the library holds
// these 404 KB until exit.  All calls on a device share the array, it keeps the last call's contents (nothing in it
// has to start at zero: every word a call reads is written by that call first; the caller's d_grad IS cleared here
// because the adjoint kernel accumulates), and two calls that overlap on the GPU race -- serialise them on one stream.
// =====================================================================================================================
#include "../../include/ast_b200.h"
#include <cstdio>
#include <mutex>
#include <vector>

int ast_set_error(int code, const char* msg);          // api.cu: routes into ast_last_error()

#define NA_MAX_T AST_NOALLOC_MAX_T                                      // 65536 samples
#define NA_MAX_NF (NA_MAX_T / REG_HOP)                                   // 128 >= 1 + (T - 1024) / 512 frames
#define NA_OFF_TW 0                                                      // float2[1024]
#define NA_OFF_DADX (NA_OFF_TW + REG_FRAME * 8)                          // float[NA_MAX_T]
#define NA_OFF_FPART (NA_OFF_DADX + NA_MAX_T * 4)                        // float[NA_MAX_NF * REG_BCHUNK]
#define NA_OFF_RAW (NA_OFF_FPART + NA_MAX_NF * REG_BCHUNK * 4)           // float[1] (padded to 256 B)
#define NA_OFF_SGN (NA_OFF_RAW + 256)                                    // int8[NA_MAX_NF * 513 * 2]
#define NA_BYTES (((NA_OFF_SGN + NA_MAX_NF * REG_BINS * 2) + 255) / 256 * 256)
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_BYTES];

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the device address is asked for once
// per device (with that device current) and remembered here.
#define NA_MAX_DEVICES 64
static unsigned char* g_na_ptr[NA_MAX_DEVICES] = {nullptr};
static std::mutex g_na_mutex;

__global__ void stft_l1_finish_kernel(const float* __restrict__ raw, double scale, float* __restrict__ reg) {
    reg[0] = (float)(scale * (double)raw[0]);                            // as losses_final_kernel forms losses[3]
}

static int noalloc_setup(int dev, cudaStream_t s, unsigned char** out) {
    std::lock_guard<std::mutex> lock(g_na_mutex);
    if (g_na_ptr[dev]) { *out = g_na_ptr[dev]; return AST_OK; }
    // one-time, synchronous: not allowed while the caller's stream is being captured
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &cs) != cudaSuccess) return ast_set_error(AST_ECUDA, "ast_stft_reg_noalloc: cudaStreamIsCapturing failed");
    if (cs != cudaStreamCaptureStatusNone)
        return ast_set_error(AST_ESTATE, "ast_stft_reg_noalloc: the first call on a device sets up its scratch (one synchronous 8 KB "
                                         "copy) and must be made outside stream capture");
    unsigned char* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || !ptr) return ast_set_error(AST_ECUDA, "ast_stft_reg_noalloc: cudaGetSymbolAddress(scratch) failed");
    // the twiddle table of setup_regulariser (api.cu): exact at the quadrant points
    float2 tw[REG_FRAME];
    for (int j = 0; j < REG_FRAME; ++j) {
        const int q = j & 255, quad = j >> 8;
        const double ang = 2.0 * 3.14159265358979323846 * q / 1024.0;
        const double cq = cos(ang), sq = q ? sin(ang) : 0.0;
        double cj, sj;
        switch (quad) { case 0: cj = cq; sj = sq; break; case 1: cj = -sq; sj = cq; break;
                        case 2: cj = -cq; sj = -sq; break; default: cj = sq; sj = -cq; }
        tw[j] = make_float2((float)cj, (float)sj);
    }
    if (cudaMemcpy(ptr + NA_OFF_TW, tw, sizeof(tw), cudaMemcpyHostToDevice) != cudaSuccess)
        return ast_set_error(AST_ECUDA, "ast_stft_reg_noalloc: copying the twiddle table failed");
    g_na_ptr[dev] = ptr;
    *out = ptr;
    return AST_OK;
}

extern "C" int ast_stft_reg_noalloc(const float* d_x, int32_t T, float* d_reg, float* d_grad, void* stream) {
    if (!d_x || !d_reg || !d_grad) return ast_set_error(AST_EINVAL, "ast_stft_reg_noalloc: null argument");
    if (T < REG_FRAME || T % 128 != 0) return ast_set_error(AST_EINVAL, "ast_stft_reg_noalloc: T must be a multiple of 128 and at least 1024 (one STFT frame)");
    if (T > NA_MAX_T) {
        char msg[160];
        snprintf(msg, sizeof(msg), "ast_stft_reg_noalloc: T = %d exceeds AST_NOALLOC_MAX_T = %d (the fixed scratch array); use "
                 "ast_set_targets(gamma) + ast_step for longer clips", (int)T, (int)NA_MAX_T);
        return ast_set_error(AST_EINVAL, msg);
    }
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= NA_MAX_DEVICES) return ast_set_error(AST_ECUDA, "ast_stft_reg_noalloc: no current CUDA device");
    cudaStream_t s = (cudaStream_t)stream;
    unsigned char* p = nullptr;
    const int r = noalloc_setup(dev, s, &p);
    if (r != AST_OK) return r;
    const float2* tw = reinterpret_cast<const float2*>(p + NA_OFF_TW);
    float* dadx = reinterpret_cast<float*>(p + NA_OFF_DADX);
    float* fpart = reinterpret_cast<float*>(p + NA_OFF_FPART);
    float* raw = reinterpret_cast<float*>(p + NA_OFF_RAW);
    signed char* sgn = reinterpret_cast<signed char*>(p + NA_OFF_SGN);
    const int nf = 1 + (T - REG_FRAME) / REG_HOP;                        // methods.py:122: no end padding
    const double scale = 1.0 / ((double)nf * (double)REG_BINS);
    if (cudaMemsetAsync(d_grad, 0, (size_t)T * sizeof(float), s) != cudaSuccess) return ast_set_error(AST_ECUDA, "ast_stft_reg_noalloc: cudaMemsetAsync failed");
    launch_stft_l1_fwd(d_x, tw, T, 1, 0, 0, nf, 0, nf, sgn, fpart, dadx, raw, s);
    stft_l1_finish_kernel<<<1, 1, 0, s>>>(raw, scale, d_reg);
    ast_launch_counter += 1;
    launch_stft_l1_grad(dadx, tw, sgn, T, 1, 0, 0, nf, (float)scale, d_grad, s);
    if (cudaGetLastError() != cudaSuccess) return ast_set_error(AST_ECUDA, "ast_stft_reg_noalloc: kernel launch failed");
    return AST_OK;
}
