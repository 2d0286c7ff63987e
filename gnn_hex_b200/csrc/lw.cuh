// Layer-at-a-time tensor-core path ("lw"): one launch per SAGE layer over ALL node rows of the batch, activations in
// global memory (L2-resident between launches).  It exists for what the fused one-tile-per-CTA kernels (sage_fwd_tc.cu /
// sage_bwd_tc.cu) cannot hold on chip:
//   * graphs that do not fit one 128-row tile (12x12 .. 19x19 boards of the board-size-transfer evaluation,
//     GN0/RainbowDQN/evaluate_elo.py:228-251, GN0/util/plotting.py:61-80) -- forward AND training,
//   * `--norm=True` (PyG LayerNorm in graph mode WITHOUT a batch vector: statistics over all nodes and channels of the batch,
//     GN0/models.py:283-287,481-482,550-551,935,945) -- a grid-wide reduction between every two layers,
//   * the HexAra `SAGE_torch_script` network (GN0/torch_script_models.py:286-379): in_channels 3, no activation after the last
//     layer of a stack, a policy stack that ends in a 1-channel SAGE layer, both stacks evaluated per call.
// One kernel serves every H -> H' SAGE layer in both directions:
//     out[i] = epi( [ x_i | aggr_i(x) ] * [ W_self ; W_agg ] )           (128 rows per CTA, tcgen05, A in tensor memory)
//   forward   aggr_i = mean_{j in N(i)} x_j,            W_self = lin_r.W^T, W_agg = lin_l.W^T, epi = + bias [, ReLU] [, statistics]
//   backward  aggr_j = sum_{i in N(j)} dpre_i / deg_i,  W_self = lin_r.W,   W_agg = lin_l.W,   epi = * (h > 0)  (dH -> dpre of the
//             layer below); the graph is symmetric, so the adjoint of the mean-gather is a gather over the same CSR rows.
// i.e. aggregate first, exactly the reference's order (torch_script_models.py:52-73: lin_l(mean_j x_j) + lin_r(x_i)).
// The contraction uses the all-fp16 3-product scheme of sage_tc.cuh (fp32 parity, two accumulators).
#pragma once
#include "common.cuh"
#include "tc_common.cuh"

namespace hexgnn {

constexpr int kLwRows = 128;                 // node rows per CTA (MMA M)
constexpr int kLwWorkWarps = 16;
constexpr int kLwWorkThreads = kLwWorkWarps * 32;
constexpr int kLwThreads = kLwWorkThreads + 32;   // + one control warp (weight copies, TMEM allocation, MMA issue)
constexpr int kLwMaxK = 112;                 // padded channel counts supported (multiples of 16)
constexpr int kLwMaxPack = 40;               // weight-image entries per pack launch

inline int lw_pad16(int c) { return (c + 15) / 16 * 16; }
// bytes of the weight image of one layer direction: 2 Kp / 16 stages of Np * 64 bytes
inline int64_t lw_img_bytes(int Kp, int Np) { return (int64_t)(2 * Kp / 16) * Np * 64; }

struct LwSageArgs {
  const float* x; int ldx;          // [N][ldx] input rows; columns [0, Kp) are read (pad columns hold zeros)
  int64_t N;
  int Kp, Np, Nout;                 // padded in / out channels (multiples of 16, <= 112), real out channels
  const int* rowptr; const int* col;
  const float* degf;                // [N] max(deg, 1) as float
  const float* rdeg;                // [N] 1 / max(deg, 1)
  int adjoint;                      // 0: aggr_i = (sum_j x_j) / deg_i   1: aggr_j = sum_i x_i / deg_i
  const uint8_t* wimg;              // lw_img_bytes(Kp, Np)
  const float* bias;                // [Np] or nullptr
  float* out; int ldo;              // [N][ldo], columns [0, Np) written (pad columns: zeros)
  int relu;
  const float* mask_src; int ldm;   // out *= (mask_src[row][c] > 0), or nullptr
  const unsigned* in_absmax;        // bits of max |x| over the whole input: power-of-two operand scale (gradients), or nullptr
  unsigned* out_absmax;             // atomicMax of the bits of max |out|, or nullptr
  double* stats;                    // [gridDim.x][2]: sum, sum of squares of `out` over live rows and real channels; or nullptr
  float* agg_out; int lda;          // the aggregated half of the operand rows ([N][lda], columns [0, Kp)), or nullptr
  float* raw_out; int ldr;          // a copy of `out` BEFORE the mask (the embedding-gradient tap), or nullptr
  int accum;                        // 1: out += result (after mask / ReLU; each element has one writer: deterministic)
  int* status;                      // HEXGNN_ST_* bits
  int rows;                         // set by the launcher: live rows per block (lw_sage_rows)
};
int launch_lw_sage(const LwSageArgs& a, cudaStream_t st);
int debug_read_lw_phases(long long* host, int n);   // HEXGNN_PHASES builds: summed per-phase cycles of lw_sage_kernel
int lw_sage_rows(int64_t N);        // rows per block of lw_sage_kernel for N node rows
int lw_sage_blocks(int64_t N);      // its grid size == number of statistics partial sums it writes

// Weight image of one layer direction.  value(c, k) of the [Np x 2Kp] operand, c = output column of the GEMM, k = contraction
// index: k < Kp -> W_self, else W_agg (k - Kp); transposed == 0: W[c * ld + k] (forward: lin.weight[out][in]);
// transposed == 1: W[k * ld + c] (backward: dH = dZ * W).  Offsets are floats from `params`; [2] = maker / breaker variant
// (chosen on device from x[2] == 1, models.py:556; give the same offset twice for shared layers).
struct LwPackEntry {
  int64_t w_self[2], w_agg[2], bias[2];   // bias < 0: none
  int ld, transposed;
  int K, Kp, Nn, Np;                      // real / padded contraction and output sizes
  int64_t img_off;                        // bytes into the image buffer
  int64_t bias_off;                       // floats into the bias buffer ([Np] written when bias >= 0)
};
struct LwPackArgs {
  LwPackEntry e[kLwMaxPack];
  int n;
  const float* params;
  const float* x;                         // head selector x[2]; nullptr: variant 0
  uint8_t* img;
  float* bias_out;
};
int launch_lw_pack(const LwPackArgs& a, cudaStream_t st);

// degf[i] = max(deg_i, 1), degf[N + i] = 1 / degf[i]; also zeroes `zero_words` 32-bit words at `zero` (absmax slots, status scratch)
int launch_lw_prepare(const int* rowptr, int64_t N, float* degf, unsigned* zero, int zero_words, cudaStream_t st);

// First layer (C_in <= 4 input channels read from x[:, xoff : xoff + C_in], row pitch ldx) on the CUDA cores:
//   out = [relu]( lin_l.W mean_j(x_j) + b + lin_r.W x_i )
struct LwLayer0Args {
  const float* x; int ldx, xoff, Cin;
  int64_t N;
  const int* rowptr; const int* col; const float* degf;
  const float* wl; const float* bl; const float* wr;   // [H][Cin], [H], [H][Cin]
  int H, Hp;
  float* out; int ldo;
  int relu;
  double* stats;                          // [gridDim.x][2] or nullptr
};
int lw_layer0_blocks(int64_t N);
int launch_lw_layer0(const LwLayer0Args& a, cudaStream_t st);

// Graph-mode LayerNorm (+ optional ReLU) as an element-wise pass: y = ((x - mu) / (sigma + eps)) * gamma + beta, statistics
// (mu, sigma over all N * H elements) from the `nparts` partial sums the producing kernel left in stats_in.
struct LwNormArgs {
  const float* in; int ldi;
  float* out; int ldo;
  int64_t N; int H, Hp;
  const double* stats_in; int nparts;
  const float* gamma[2]; const float* beta[2];   // maker / breaker variant (same pointer twice for shared norms)
  const float* x;                                // head selector
  int relu;
  double* stats_out;                             // [gridDim.x][2] partial sums of the OUTPUT, or nullptr
  float eps;
};
int lw_norm_blocks(int64_t N);
int launch_lw_norm(const LwNormArgs& a, cudaStream_t st);

}  // namespace hexgnn
