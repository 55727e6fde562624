// tc.cuh -- interface of the tcgen05 / TMEM / TMA layer kernels (kernels_tc.cu), AST_PREC_BF16.
#pragma once
#include "common.cuh"

struct TcState;

// Allocates the bf16 operand buffers and encodes the TMA tensor maps over the caller's fp32 buffers:
// H [NL+1][R][128] residual stream, g [R][128] running gradient, dF [n_head][R][128] head gradients.
TcState* tc_create(int T, int B, int NL, float* H, float* g, float* dF, int n_head, int64_t* bytes, const char** why);
void tc_destroy(TcState* s);
// bf16 operand copies of the fp32 master weights (device pointers, layouts of ast_set_weights).
int tc_set_weights(TcState* s, const float* Wd, const float* Wr, cudaStream_t st);
// split relu(H[0]) into the hi/lo bf16 A operands of layer 0
void tc_prepare_input(TcState* s, const float* H0, cudaStream_t st);
// forward layer l (model.py:99-114): reads relu(H[l]) as bf16 hi/lo + H[l] fp32, writes H[l+1],
// the bf16 hi/lo of relu(H[l+1]), zmask[l], hmask[l+1].
void tc_layer_fwd(TcState* s, int l, const float* bd, const float* br, uint32_t* zmask_out, uint32_t* hmask_out,
                  cudaStream_t st);
// backward boundary l in NL..0: g += conv^T(dz_l) * hmask + dF[df_idx] (df_idx < 0: none);
// dz_{l-1} = (g Wr_{l-1}^T) * zmask.  bf16 dz buffers owned by the TcState.  l == NL: no conv (top);
// l == 0: no 1x1 (bottom).
void tc_layer_bwd(TcState* s, int l, const uint32_t* hmask_in, int df_idx, const uint32_t* zmask_in, cudaStream_t st);
