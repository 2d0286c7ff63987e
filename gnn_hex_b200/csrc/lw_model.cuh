// modern_two_headed on the layer-at-a-time path: workspace / stash layouts and launchers (lw_model.cu, lw_bwd.cu).
#pragma once
#include "lw.cuh"

namespace hexgnn {

// partial-sum slots per statistics record: the producing kernels use 128-row (lw_sage) or 64-row (layer 0, norm) blocks
inline int64_t lw_stat_parts(int64_t N) { return (N + 63) / 64 > 0 ? (N + 63) / 64 : 1; }

struct LwModelWs {   // byte offsets into the workspace
  int64_t degf, zero, bias, img, imgT, stats, act, total;
  int64_t parts;
};
LwModelWs lw_model_ws(const Dims& d, int64_t N, bool backward);

// Training stash of the lw path (floats):  h_s [Lt][N][Hp] (output of layer s after [norm +] ReLU)
//   norm only:  | pre_s [Lt][N][Hp] (layer outputs before the norm) | after_embed_norm output [N][Hp]
//               | statistics partial sums, (2 Lt + 2) records of lw_stat_parts(N) x (sum, sum of squares) doubles
int64_t lw_stash_floats(const Dims& d, int64_t N, bool norm);

// norm parameter block (floats): per LayerNorm gamma[H] then beta[H]; order: body layers 0..L-1, after_embed_norm,
// maker head layers 0..Lh-1, breaker head layers 0..Lh-1   ->  (L + 1 + 2 Lh) * 2H floats
inline int64_t lw_norm_floats(const Dims& d) { return (int64_t)(d.L + 1 + 2 * d.Lh) * 2 * d.H; }

int launch_forward_lw(const Dims& d, const ParamMap& pm, const float* params, const float* normp, const float* x, int64_t N,
                      int64_t E, int64_t B, const void* csr, int flags, float* out_nodes, float* out_value, float* stash,
                      void* ws, cudaStream_t st);

// backward (lw_bwd.cu)
int64_t lw_bwd_ws_bytes(const Dims& d, int64_t N, int64_t B, bool norm);
int launch_backward_lw(const Dims& d, const ParamMap& pm, const float* params, const float* normp, const float* x, int64_t N,
                       int64_t E, int64_t B, const void* csr, int flags, const float* d_nodes, const float* d_value,
                       const float* stash, float* grad_params, float* grad_norm, float* d_embeds, void* ws, cudaStream_t st);

// weight-gradient GEMM over a table of (input rows X [N][Hp], [dpre | dZl] rows Y [N][2Hp]) pairs, one per H -> H conv:
// records rec [nitems][lw_dw_groups][(Hp + 1)][2Hp] (row k < Hp: d W^T rows, row Hp: the bias row; columns [0, Hp) belong to
// lin_r / the bias, [Hp, 2Hp) to lin_l); first-layer partials part [lw_dw0_blocks][2 Cin + 1][Hp] (rows: lin_l.W[:, k], bias,
// lin_r.W[:, k])
int lw_dw_groups(int nitems);
int64_t lw_dw_rec_floats(int Hp, int nitems);
int launch_lw_dw_tab(const float* const* xs, const float* const* ys, int nitems, int64_t N, int Hp, float* rec, cudaStream_t st);
int lw_dw0_blocks(int64_t N);
int launch_lw_dw0(const float* x, int ldx, int Cin, int64_t N, const int* rowptr, const int* col, const float* degf,
                  const float* dpre, int ldd, int Hp, float* part, cudaStream_t st);

}  // namespace hexgnn
