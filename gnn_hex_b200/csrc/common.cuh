// Shared host/device definitions for libhexgnn.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/hexgnn.h"

namespace hexgnn {

constexpr int kMaxSage = 48;        // body + head SAGE layers supported by the kernel parameter block
constexpr int kThreads = 512;       // CTA size of the fused SIMT kernels (16 warps)
constexpr int kWarps = kThreads / 32;
constexpr int kMaxTileGraphs = 64;  // graphs packed into one tile at most
constexpr int kEdgeCapPerRow = 8;   // tile edge capacity = kEdgeCapPerRow * tile_rows (uint16 local ids)

// Derived dimensions of one model.
struct Dims {
  int H;    // hidden
  int Hp;   // hidden padded to a multiple of 16
  int NJ;   // Hp / 16 == (2*Hp) / 32 : column groups per lane in the [Hp x 2Hp] contraction
  int L;    // body layers
  int Lh;   // head layers
  int Lt;   // L + Lh
  int H2;   // H / 2 (value MLP hidden)
};

// Float offsets into the flat parameter buffer (layout documented in include/hexgnn.h).
struct ParamMap {
  int64_t sage[2][kMaxSage];  // [head][s]: offset of lin_l.weight of SAGE layer s (s<L: body, shared by both heads)
  int32_t cin[kMaxSage];      // input channels of SAGE layer s
  int64_t lin_w[2], lin_b[2]; // head.linear
  int64_t v0_w[2], v0_b[2], v1_w[2], v1_b[2];
  int64_t total;
};

inline __host__ __device__ int64_t sage_wl(const ParamMap& pm, int head, int s) { return pm.sage[head][s]; }
inline __host__ __device__ int64_t sage_bl(const ParamMap& pm, int head, int s, int H) {
  return pm.sage[head][s] + (int64_t)H * pm.cin[s];
}
inline __host__ __device__ int64_t sage_wr(const ParamMap& pm, int head, int s, int H) {
  return pm.sage[head][s] + (int64_t)H * pm.cin[s] + H;
}

inline bool make_dims(const HexModelDesc* d, Dims* o) {
  if (!d || d->hidden < 2 || d->num_layers < 1 || d->num_head_layers < 1 || d->in_channels != 2) return false;
  if (d->num_layers + d->num_head_layers > kMaxSage) return false;
  o->H = d->hidden;
  o->Hp = (d->hidden + 15) / 16 * 16;
  o->NJ = o->Hp / 16;
  o->L = d->num_layers;
  o->Lh = d->num_head_layers;
  o->Lt = o->L + o->Lh;
  o->H2 = d->hidden / 2;
  return o->H2 >= 1;
}

inline void make_param_map(const Dims& d, ParamMap* pm) {
  int64_t off = 0;
  const int H = d.H;
  for (int s = 0; s < d.L; ++s) {
    int cin = s == 0 ? 2 : H;
    pm->cin[s] = cin;
    pm->sage[0][s] = pm->sage[1][s] = off;
    off += (int64_t)H * cin + H + (int64_t)H * cin;
  }
  for (int head = 0; head < 2; ++head) {
    for (int s = d.L; s < d.Lt; ++s) {
      pm->cin[s] = H;
      pm->sage[head][s] = off;
      off += (int64_t)H * H + H + (int64_t)H * H;
    }
    pm->lin_w[head] = off; off += H;
    pm->lin_b[head] = off; off += 1;
    pm->v0_w[head] = off; off += (int64_t)d.H2 * 4 * H;
    pm->v0_b[head] = off; off += d.H2;
    pm->v1_w[head] = off; off += d.H2;
    pm->v1_b[head] = off; off += 1;
  }
  pm->total = off;
}

// ---- csr blob ---------------------------------------------------------------------------
inline int64_t align_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

constexpr int kScanBlockElems = 1024;   // csr_scan_kernel: 256 threads x int4
__host__ __device__ inline int64_t csr_scan_blocks(int64_t N) { return (N + 1 + kScanBlockElems - 1) / kScanBlockElems; }

inline void csr_layout(int64_t N, int64_t E, int64_t B, HexCsrLayout* o) {
  int64_t w = 0;
  o->status = w; w += 1;
  o->ntiles = w; w += 1;
  w = align_up(w, 4);
  o->gptr = w; w += B + 1; w = align_up(w, 4);
  o->rowptr = w; w += N + 1; w = align_up(w, 4);
  o->col = w; w += E; w = align_up(w, 4);
  o->tile_g0 = w; w += B + 1; w = align_up(w, 4);
  o->scratch = w; w += (N + 1) + E; w = align_up(w, 4);  // cursor[N+1], edge ids [E]
  w += 4 + 2 * csr_scan_blocks(N);                       // scan ticket + one 64-bit look-back state per scan block
  w = align_up(w, 4);
  w += B + 1; w = align_up(w, 4);                        // tile_order[B]: processing order of the tiles (heaviest first)
  o->total_words = w;
}

// Device view of the csr blob.
struct CsrView {
  int* status;
  int* ntiles;
  int* gptr;
  int* rowptr;
  int* col;
  int* tile_g0;
  int* cursor;
  int* eid;
  int* scan_ticket;                 // [4] (only [0] used)
  unsigned long long* scan_state;   // [csr_scan_blocks(N)]: flag << 32 | value
  int* tile_order;                  // [ntiles]: position -> tile, by descending cost (8 * rows + edges), ties by index
};

inline CsrView csr_view(void* blob, int64_t N, int64_t E, int64_t B) {
  HexCsrLayout l;
  csr_layout(N, E, B, &l);
  int* w = (int*)blob;
  CsrView v;
  v.status = w + l.status;
  v.ntiles = w + l.ntiles;
  v.gptr = w + l.gptr;
  v.rowptr = w + l.rowptr;
  v.col = w + l.col;
  v.tile_g0 = w + l.tile_g0;
  v.cursor = w + l.scratch;
  v.eid = w + l.scratch + (N + 1);
  v.scan_ticket = w + align_up(l.scratch + (N + 1) + E, 4);
  v.scan_state = (unsigned long long*)(v.scan_ticket + 4);
  v.tile_order = v.scan_ticket + align_up(4 + 2 * csr_scan_blocks(N), 4);
  return v;
}

// ---- forward workspace: packed weights ----------------------------------------------------
// For SAGE layers s = 1..Lt-1 of the TAKEN head (chosen on device from x[0,2]):
//   wcat [s-1][Hp][2Hp]   wcat[k][c] = lin_l.W[c][k] (c<H), wcat[k][Hp+c] = lin_r.W[c][k]; zero padding
//   wcatT[s-1][2Hp][Hp]   transpose (backward: dH = dZ * wcat^T)
//   bias [Lt][Hp]         lin_l.bias, zero padded (entry 0 is layer 0's)
struct PackView {
  float* wcat;
  float* wcatT;
  float* bias;
};
inline int64_t pack_floats(const Dims& d) {
  return (int64_t)(d.Lt - 1) * d.Hp * 2 * d.Hp * 2 + (int64_t)d.Lt * d.Hp;
}
inline PackView pack_view(void* ws, const Dims& d) {
  PackView p;
  p.wcat = (float*)ws;
  p.wcatT = p.wcat + (int64_t)(d.Lt - 1) * d.Hp * 2 * d.Hp;
  p.bias = p.wcatT + (int64_t)(d.Lt - 1) * d.Hp * 2 * d.Hp;
  return p;
}

// ---- backward workspace: per-CTA gradient partials ------------------------------------------
// Packed gradient record of one CTA (floats):
//   gw  [Lt-1][Hp][2Hp]  d wcat        (layers 1..Lt-1)
//   gb  [Lt][Hp]         d bias
//   g0  [4][Hp]          layer 0: d lin_l.W[:,0], d lin_l.W[:,1], d lin_r.W[:,0], d lin_r.W[:,1]
//   glin[Hp], glinb[1 (padded to 4)]
//   gv0 [H2][4H] (padded to multiple of 4), gv0b[H2p], gv1[H2p], gv1b[4]
struct GradLayout {
  int64_t gw, gb, g0, glin, glinb, gv0, gv0b, gv1, gv1b, total;
};
inline void grad_layout(const Dims& d, GradLayout* g) {
  int64_t o = 0;
  g->gw = o; o += (int64_t)(d.Lt - 1) * d.Hp * 2 * d.Hp;
  g->gb = o; o += (int64_t)d.Lt * d.Hp;
  g->g0 = o; o += 4 * d.Hp;
  g->glin = o; o += d.Hp;
  g->glinb = o; o += 4;
  g->gv0 = o; o += align_up((int64_t)d.H2 * 4 * d.H, 4);
  g->gv0b = o; o += align_up(d.H2, 4);
  g->gv1 = o; o += align_up(d.H2, 4);
  g->gv1b = o; o += 4;
  g->total = align_up(o, 4);
}

int bwd_grid_ctas();  // number of persistent CTAs the backward uses (= partial records)

// ---- error plumbing -------------------------------------------------------------------------
void set_error(const char* fmt, ...);
#define HEX_CUDA_CHECK(expr)                                                            \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess) {                                                            \
      ::hexgnn::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return HEXGNN_E_CUDA;                                                             \
    }                                                                                   \
  } while (0)

// ---- per-kernel device timing for bench.py (api.cu) -------------------------------------------
// ProfScope brackets one launch with CUDA events on `st` when profiling is enabled on this thread.
struct ProfScope {
  ProfScope(int which, cudaStream_t st);
  ~ProfScope();
  int slot_;
  cudaStream_t st_;
};

// ---- launchers implemented in the .cu files ---------------------------------------------------
int launch_build_csr(const int64_t* edge_index, int64_t E, const int64_t* batch, const int64_t* ptr, int64_t N,
                     int64_t B, int tile_rows, void* csr, cudaStream_t st);
int launch_pack(const Dims& d, const ParamMap& pm, const float* params, const float* x, PackView pv, bool with_transpose,
                cudaStream_t st);
int launch_forward_simt(const Dims& d, const ParamMap& pm, const float* params, const float* x, int64_t N, int64_t E,
                        int64_t B, const void* csr, int tile_rows, int flags, float* out_nodes, float* out_value,
                        float* stash, PackView pv, cudaStream_t st);
int launch_backward_simt(const Dims& d, const ParamMap& pm, const float* params, const float* x, int64_t N, int64_t E,
                         int64_t B, const void* csr, int tile_rows, int flags, const float* d_nodes,
                         const float* d_value, const float* stash, float* d_embeds, PackView pv, float* partials,
                         cudaStream_t st);
int launch_reduce_grads(const Dims& d, const ParamMap& pm, const float* x, const float* partials, int nparts,
                        const int* ntiles, float* packed_sum, float* grad_params, cudaStream_t st);

int launch_unpack_grads(const Dims& d, const ParamMap& pm, const float* x, const float* packed_sum, float* grad_params,
                        cudaStream_t st);
struct TdLossArgs {
  const float* q; const int64_t* ptr; const int64_t* action; const float* target; const float* weight;
  int64_t N; int B, kind; float scale; float* loss; float* td_err; float* d_q;
};
int launch_td_loss(const float* q, const int64_t* ptr, const int64_t* action, const float* target, const float* weight,
                   int64_t N, int64_t B, int kind, float scale, float* loss, float* td_err, float* d_q, cudaStream_t st);
int launch_adam_step(const Dims& d, const ParamMap& pm, float* params, float* grads, float* exp_avg, float* exp_avg_sq,
                     int* state, const float* x, float lr, float beta1, float beta2, float eps, float max_norm,
                     int skip_untaken, float* grad_norm, float* ws, cudaStream_t st);
int adam_ws_floats();
int launch_noisy_compose(const Dims& d, const ParamMap& pm, float* params, const float* noisy, cudaStream_t st);
int launch_noisy_grads(const Dims& d, const ParamMap& pm, const float* grads, const float* noisy, float* d_noisy,
                       int accumulate, cudaStream_t st);
int64_t collate_ws_bytes(int64_t B);
int launch_collate(const float* const* x_ptrs, const int64_t* const* ei_ptrs, const int64_t* n_nodes, const int64_t* n_edges,
                   int64_t B, float* x_out, int64_t* ei_out, int64_t* batch_out, int64_t* ptr_out, void* ws, void* host_ws,
                   cudaStream_t st);
int launch_forward_big(const Dims& d, const ParamMap& pm, const float* params, const float* x, int64_t N, int64_t E,
                       int64_t B, const void* csr, int tile_rows, int flags, int max_rows, float* out_nodes,
                       float* out_value, float* ws, cudaStream_t st);
int launch_allreduce_p2p(const float* const* peers_dev, int world, int64_t n, float* out, cudaStream_t st);
int launch_allreduce_mc(float* mc_buf, int* const* peer_flags_dev, int rank, int world, int64_t n, int epoch, int* counter,
                        cudaStream_t st);
int launch_mask_argmax(float* q, const int64_t* ptr, int64_t N, int64_t B, int mask_terminals, float mask_value,
                       float* out_max, int64_t* out_arg, cudaStream_t st);
int launch_td_target(const float* q_target, const int64_t* best, const float* reward, const float* done, float gamma_n,
                     int64_t N, int64_t B, float* target, cudaStream_t st);
int launch_tc_selftest(const float* A, const float* B, float* D, int N, int K, int modes, cudaStream_t st);

}  // namespace hexgnn
