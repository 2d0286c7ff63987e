// Shared definitions of the tcgen05 (tensor-core) fused kernels.
//
// Per SAGE layer the node-linear contraction is ONE tensor-core GEMM per 128-row tile,
//     Z[128, 2Hp] = h[128, Hp] * Wcat[Hp, 2Hp],   Wcat = [lin_l.W^T | lin_r.W^T]
// (transform first, aggregate after: mean_j(h_j) W_l^T == mean_j(h_j W_l^T)), computed as 3xTF32
// (hi*hi + lo*hi + hi*lo, fp32 accumulate) so the result keeps fp32 parity with the reference
// (GN0/torch_script_models.py:52-73 restates the SAGEConv math).  A = h lives in TENSOR MEMORY
// (lane = node row), B = weight images stream through a shared-memory ring from L2, D = Z accumulates
// in tensor memory; the epilogue threads (one per node row) dump Zl to shared memory, gather it over the
// CSR, add Zr + bias, ReLU, and write the next layer's h straight back into tensor memory.
#pragma once
#include "common.cuh"
#include "tc_common.cuh"

namespace hexgnn {

constexpr int kTcRows = 128;          // MMA M: node rows per tile
constexpr int kTcKC = 16;             // k values per streamed weight chunk (2 MMA k-steps)
constexpr int kTcStages = 4;          // ring depth
constexpr int kTcEpiWarps = 16;       // epilogue warps: 4 TMEM lane quarters x 4 column slices
constexpr int kTcSlices = kTcEpiWarps / 4;
constexpr int kTcEpiThreads = kTcEpiWarps * 32;
constexpr int kTcThreads = kTcEpiThreads + 64;  // + producer warp + MMA warp

// Weight images for the forward, one per SAGE layer s = 1..Lt-1 of the taken head:
//   img[s-1][q][half][...]  q = k chunk (Hp/16 of them), half 0 = tf32 hi, 1 = tf32 lo
//   inside a half: K-major SWIZZLE_NONE canonical B operand, B[n][k], n in [0,2Hp), k in [16q,16q+16):
//   byte = ((k%16)/4) * (2Hp*16) + (n/8)*128 + (n%8)*16 + (k%4)*4
struct TcPackView {
  float* wimg;    // forward images   (Lt-1) * Hp * 2Hp * 2 floats
  float* wimgT;   // backward images  (same size): B[n=(which,k)][kk=c] = W_which[c][k]
  float* bias;    // [Lt][Hp]
};
inline int64_t tc_img_floats(const Dims& d) { return (int64_t)(d.Lt - 1) * d.Hp * 2 * d.Hp * 2; }
inline int64_t tc_pack_floats(const Dims& d) { return 2 * tc_img_floats(d) + (int64_t)d.Lt * d.Hp; }
inline TcPackView tc_pack_view(void* ws, const Dims& d) {
  TcPackView p;
  p.wimg = (float*)ws;
  p.wimgT = p.wimg + tc_img_floats(d);
  p.bias = p.wimgT + tc_img_floats(d);
  return p;
}

int launch_pack_tc(const Dims& d, const ParamMap& pm, const float* params, const float* x, TcPackView pv,
                   bool with_transpose, cudaStream_t st);
int launch_forward_tc(const Dims& d, const ParamMap& pm, const float* params, const float* x, int64_t N, int64_t E,
                      int64_t B, const void* csr, int flags, float* out_nodes, float* out_value, float* stash,
                      TcPackView pv, cudaStream_t st);

int debug_read_fwd_phases(long long* host, int n);

}  // namespace hexgnn
