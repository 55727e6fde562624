// tc.cuh -- interface of the tcgen05 / TMEM / TMA kernels (kernels_tc.cu, gram_tc.cu), AST_PREC_BF16.
#pragma once
#include "common.cuh"

#include <cuda.h>

struct TcState;

// Allocates the bf16 operand buffers and encodes the TMA tensor maps over the caller's fp32 buffers:
// H [NL+1][R][128] residual stream, g [R][128] running gradient, dF [n_head][R][128] head gradients.
// n_fslots > 0 (Gatys head): also a bf16 copy F [n_fslots][R][128] of the style extracts, a bf16 image
// of Ssym [B*n_fslots][128][128], and a map over the caller's Gram partial buffer
// gram_part [gram_part_mats][128][128] fp32.
// split: the accurate tier (AST_PREC_TC3): the backward sweep keeps hi / lo bf16 pairs of every operand (stack_bwd_split.cuh)
TcState* tc_create(int T, int B, int NL, float* H, float* g, float* dF, int n_head, int n_fslots,
                   float* gram_part, int64_t gram_part_mats, bool split, int64_t* bytes, const char** why);
void tc_destroy(TcState* s);
// bf16 operand copies of the fp32 master weights (device pointers, layouts of ast_set_weights).
int tc_set_weights(TcState* s, const float* Wd, const float* Wr, cudaStream_t st);
// split relu(H[0]) into the hi/lo bf16 A operands of layer 0
void tc_prepare_input(TcState* s, const float* H0, cudaStream_t st);
// forward layer l (model.py:99-114): reads relu(H[l]) as bf16 hi/lo + H[l] fp32, writes H[l+1],
// the bf16 hi/lo of relu(H[l+1]), zmask[l], hmask[l+1] and, if f_slot >= 0, bf16(H[l+1]) -> F[f_slot].
void tc_layer_fwd(TcState* s, int l, const float* bd, const float* br, uint32_t* zmask_out, uint32_t* hmask_out,
                  int f_slot, cudaStream_t st);
// Whole encoder forward (start conv + all layers) in one cooperative launch; available when the clip has at most
// one 128-row tile per SM.  Writes zmask[l], hmask[l], the bf16 extract copies, and H[i] only where export_h[i] != 0
// (i = 0: start-conv output, i = l+1: output of layer l).  *tc_stack_err() != 0 afterwards = a bounded spin expired.
// Clips with more tiles than SMs run as layer SEGMENTS (one per 10-layer stack) x tile GROUPS (stack_fwd.cuh); seg < 0 runs
// every segment, seg >= 0 only that one (time shards exchange halo rows of H[tc_segment_end(seg)] / g between segments).
bool tc_stack_ok(const TcState* s);
// The whole-stack forward also writes H[l+1] of every style layer transposed, in fp16, into Ht [R/32][n_slots][128][32 rows]
// (slot = f_slot[l]): the operand layout of the TMA-fed channel-wise head (gram_ours_tma.cu).  NULL switches it off.
void tc_set_transposed_extracts(TcState* s, void* Ht, int n_slots);
unsigned* tc_stack_err(TcState* s);
void tc_plan_segments(TcState* s, bool per_stack);
// peer-mapped halos (time shards whose neighbours' buffers are mapped into this process: ast_p2p_attach)
void tc_set_peer_geometry(TcState* s, int own_lo, int own_hi, unsigned* sig_local);
void tc_clear_peers(TcState* s);
void tc_set_peer(TcState* s, int side, const float* H, const float* g, const float* g_alt, unsigned* sig, int peer_R, int row_off);
int tc_segments(const TcState* s);
int tc_segment_end(const TcState* s, int k);
int tc_forward_stack(TcState* s, const float* x, const float* W0, const float* b0, const float* bd, const float* br,
                     float* H, uint32_t* zmask, uint32_t* hmask, const int* f_slot, const unsigned char* export_h,
                     int seg, cudaStream_t st);
// Whole backward sweep (boundaries NL..0) in one cooperative launch (same availability as tc_forward_stack).
// df_idx[l] / sg_slot[l] as for tc_layer_bwd, per boundary l; the final running gradient is written to g [R][128].
int tc_backward_stack(TcState* s, const uint32_t* hmask, const uint32_t* zmask, const int* df_idx, const int* sg_slot,
                      int sg_L, float* g, float* g_alt, const float* W0, float* pdots, int seg, cudaStream_t st);
// backward boundary l in NL..0: g += conv^T(dz_l) * hmask + dF[df_idx] (df_idx < 0: none);
// dz_{l-1} = (g Wr_{l-1}^T) * zmask.  bf16 dz buffers owned by the TcState.  l == NL: no conv (top);
// l == 0: no 1x1 (bottom).
// sg_slot >= 0 (and df_idx < 0): the Gatys style gradient F[sg_slot] . Ssym[sg_slot] is computed in the same kernel.
void tc_layer_bwd(TcState* s, int l, const uint32_t* hmask_in, int df_idx, int sg_slot, int sg_L, const uint32_t* zmask_in,
                  cudaStream_t st);

// Gatys Gram partials of the owned tiles [tile_lo, tile_hi) of every pair: part[(b*L+l)][chunk][128][128]
void tc_gram(TcState* s, int L, int nchunk, int chunk_tiles, int tile_lo, int tile_hi, cudaStream_t st);
// bf16 image of Ssym that the heads kernel fills ([B*L][128][128]), or NULL; its lo part (accurate tier), or NULL
__nv_bfloat16* tc_ssym_bf16(TcState* s);
__nv_bfloat16* tc_ssym_bf16_lo(TcState* s);
// dF[head_of_slot[i]] = F[i] . Ssym[i] for every style slot i with head_of_slot[i] >= 0
void tc_style_grad(TcState* s, int L, const int* head_of_slot, cudaStream_t st);

// ---- gram_tc.cu ------------------------------------------------------------------------------------
struct alignas(64) GramTcArgs {
    CUtensorMap tmF;                 // bf16 [L][B][T][128], box {64,128,1,1}
    CUtensorMap tmFlo;               // accurate tier: the lo parts bf16(h - float(bf16(h))), same layout
    CUtensorMap tmP;                 // fp32 partials as rows [mats*128][128], box {32,32}
    int L, chunk_tiles, tile_lo, tile_hi;
};
struct alignas(64) SGradTcArgs {
    CUtensorMap tmF;                 // bf16 [L][B][T][128], box {64,128,1,1}
    CUtensorMap tmS;                 // bf16 Ssym rows [B*L*128][128], box {64,128}
    CUtensorMap tmDF;                // fp32 [n_head][B][T][128], store box {32,32,1,1}
    CUtensorMap tmFlo, tmSlo;        // accurate tier: lo parts of F and Ssym
    int T, L;
    int n_active;                    // grid.y: only slots that need the standalone kernel are launched
    int slot_of[32];                 // blockIdx.y -> style slot
    int head_of_slot[32];            // style slot -> dF buffer
};
void launch_gram_tc(const GramTcArgs& a, int nchunk, int L, int B, bool split, cudaStream_t s);
void launch_sgrad_tc(const SGradTcArgs& a, int ntiles, bool split, cudaStream_t s);
int setup_gram_tc_kernels();
