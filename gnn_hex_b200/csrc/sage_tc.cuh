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
  float* wimgT;   // backward images  (same size), layout above tc_imgB_floats
  float* bias;    // [Lt][Hp]
};
inline int64_t tc_img_floats(const Dims& d) { return (int64_t)(d.Lt - 1) * d.Hp * 2 * d.Hp * 2; }
inline int64_t tc_pack_floats(const Dims& d);
inline int64_t tc_pack_floats(const Dims& d) { return 2 * tc_img_floats(d) + (int64_t)d.Lt * d.Hp; }
inline TcPackView tc_pack_view(void* ws, const Dims& d) {
  TcPackView p;
  p.wimg = (float*)ws;
  p.wimgT = p.wimg + tc_img_floats(d);   // backward images (tc_imgB_floats == tc_img_floats)
  p.bias = p.wimgT + tc_img_floats(d);
  return p;
}

// ---- training stashes of the tcgen05 path ------------------------------------------------------------
// Both stashes are stored per (layer, tile) as "RM32" images: the row-major [row][channel] tile cut into
// 32-channel (128-byte) blocks, rows grouped by 4 with the 32-byte chunks XOR-swizzled by row%4 -- exactly
// the MN-major SWIZZLE_128B_BASE32B operand form, so the weight-gradient GEMM reads them from shared
// memory with no transposition, and 32 rows of a block are one contiguous 4 KB bulk copy.
//   byte(ch,row) = (ch/32)*16384 + (row/4)*512 + (row%4)*128 + ((((ch%32)/8) ^ (row%4))*32) + (ch%8)*4
// Activations: channels [0,Hp) = h, channel Hp = 1 for live rows (its row of h^T dZ is the bias gradient);
// rows [n, ceil32(n)) are written as zeros so whole 32-row k-chunks can be contracted.
constexpr int kImgBlockBytes = 16384;                       // one 32-channel block of a 128-row tile
__host__ __device__ inline int tc_act_blocks(int Hp) { return (Hp + 1 + 31) / 32; }
__host__ __device__ inline int tc_dz_blocks(int Hp) { return (2 * Hp) / 32; }
__device__ __forceinline__ uint32_t rm32_granule_off(int g, int row) {   // 8-channel granule g, 32 bytes
  return (uint32_t)(g >> 2) * kImgBlockBytes + (uint32_t)(row >> 2) * 512u + (uint32_t)(row & 3) * 128u +
         ((uint32_t)((g & 3) ^ (row & 3)) << 5);
}
// forward stash: [Lt][B tiles][act_blocks*16KB] images + row-major embedding copy [N][Hp] (final_conv_acts)
inline int64_t tc_stash_bytes(const Dims& d, int64_t N, int64_t B) {
  return (int64_t)d.Lt * B * tc_act_blocks(d.Hp) * kImgBlockBytes + N * d.Hp * 4;
}
inline int64_t tc_dz_bytes(const Dims& d, int64_t B) { return (int64_t)(d.Lt - 1) * B * tc_dz_blocks(d.Hp) * kImgBlockBytes; }
inline bool tc_bwd_supported(const Dims& d) { return d.Hp + 1 <= 128; }   // the ones-channel must fit M = 128

// Backward weight images, per layer s>=1, per which (0: lin_r, 1: lin_l), per 16-wide chunk q of the OUT channel c
// (the contraction index of dH = dZ * W), hi half then lo half; inside a half the K-major SWIZZLE_NONE form
//   B[n = k (in channel)][kk = c]:  float = ((c%16)/4)*(Hp*4) + (k/8)*32 + (k%8)*4 + (c%4)
inline int64_t tc_imgB_floats(const Dims& d) { return (int64_t)(d.Lt - 1) * 2 * d.Hp * d.Hp * 2; }

int launch_pack_tc(const Dims& d, const ParamMap& pm, const float* params, const float* x, TcPackView pv,
                   bool with_transpose, cudaStream_t st);
int launch_forward_tc(const Dims& d, const ParamMap& pm, const float* params, const float* x, int64_t N, int64_t E,
                      int64_t B, const void* csr, int flags, float* out_nodes, float* out_value, float* stash,
                      TcPackView pv, cudaStream_t st);

int debug_read_fwd_phases(long long* host, int n);
int debug_read_bwd_phases(long long* host, int n);
int tc_dw_groups(const Dims& d);
int launch_backward_tc_chain(const Dims& d, const ParamMap& pm, const float* params, const float* x, int64_t N, int64_t E,
                             int64_t B, const void* csr, int flags, const float* d_nodes, const float* d_value,
                             const float* stash, float* dz, float* d_embeds, TcPackView pv, float* small_rec, int grid,
                             cudaStream_t st);
int launch_backward_tc_dw(const Dims& d, int64_t N, int64_t E, int64_t B, const void* csr, const float* stash,
                          const float* dz, float* big_rec, cudaStream_t st);
int launch_reduce_tc(const Dims& d, const float* small_rec, int nsmall, const float* big_rec, int G, float* packed_sum,
                     cudaStream_t st);

// Workspace of the tcgen05 backward (floats): weight images | dZ stash | small records | group records | packed sum
struct TcBwdWs { int64_t pack, dz, small_rec, big_rec, sum, total; };
inline TcBwdWs tc_bwd_ws(const Dims& d, int64_t B) {
  GradLayout gl;
  grad_layout(d, &gl);
  TcBwdWs w;
  int64_t o = 0;
  w.pack = o; o += align_up(tc_pack_floats(d), 256);
  w.dz = o; o += tc_dz_bytes(d, B) / 4;
  w.small_rec = o; o += align_up((int64_t)bwd_grid_ctas() * (gl.total - gl.gb), 4);
  w.big_rec = o; o += (int64_t)tc_dw_groups(d) * gl.g0;
  w.sum = o; o += gl.total;
  w.total = o;
  return w;
}

}  // namespace hexgnn
