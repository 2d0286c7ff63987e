// C-ABI of libhexgnn.so (declared in include/hexgnn.h).
#include <stdarg.h>
#include <string.h>
#include <vector>
#include "common.cuh"
#include "sage_tc.cuh"
#include "lw.cuh"

namespace hexgnn {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int bwd_grid_ctas() {
  static int sms = 0;  // immutable after first use
  if (sms == 0) {
    int dev = 0, v = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0)
      sms = v;
    else
      sms = 148;
  }
  return sms;
}

// ---- profiling: thread-local event pairs --------------------------------------------------------
struct ProfState {
  bool on = false;
  std::vector<cudaEvent_t> ev[3];  // start, stop, start, stop, ...
  size_t used[3] = {0, 0, 0};
};
static ProfState g_prof;  // process-wide: autograd runs hexgnn_backward on its own thread
static thread_local int g_prof_mute = 0;  // > 0: an enclosing scope brackets several overlapped launches

ProfScope::ProfScope(int which, cudaStream_t st) : slot_(-1), st_(st) {
  if (!g_prof.on || g_prof_mute > 0 || which < 0 || which > 2) return;
  auto& v = g_prof.ev[which];
  size_t& u = g_prof.used[which];
  if (u + 2 > v.size()) {
    cudaEvent_t a, b;
    if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) return;
    v.push_back(a);
    v.push_back(b);
  }
  slot_ = which * (1 << 28) + (int)u;
  cudaEventRecord(v[u], st);
  u += 2;
}
ProfScope::~ProfScope() {
  if (slot_ < 0) return;
  const int which = slot_ >> 28, u = slot_ & ((1 << 28) - 1);
  cudaEventRecord(g_prof.ev[which][u + 1], st_);
}

static bool tile_rows_ok(int tile_rows) { return tile_rows == 128; }

// One decision for forward AND backward (their stash formats differ): the tcgen05 kernels unless the caller
// forces the CUDA-core path or the hidden size leaves no room for the ones-channel (Hp = 128).
static bool use_tc(const Dims& d, int flags) { return !(flags & HEXGNN_PATH_SIMT) && tc_bwd_supported(d); }

}  // namespace hexgnn

using namespace hexgnn;

extern "C" {

int hexgnn_abi_version(void) { return HEXGNN_ABI_VERSION; }
const char* hexgnn_last_error(void) { return g_err; }

int64_t hexgnn_param_count(const HexModelDesc* desc) {
  Dims d;
  if (!make_dims(desc, &d)) { set_error("bad model descriptor"); return HEXGNN_E_BADARG; }
  ParamMap pm;
  make_param_map(d, &pm);
  return pm.total;
}

int32_t hexgnn_param_offsets(const HexModelDesc* desc, int64_t* offsets, int32_t max_entries) {
  Dims d;
  if (!make_dims(desc, &d)) { set_error("bad model descriptor"); return HEXGNN_E_BADARG; }
  ParamMap pm;
  make_param_map(d, &pm);
  int32_t n = 0;
  auto put = [&](int64_t o) { if (offsets && n < max_entries) offsets[n] = o; ++n; };
  for (int s = 0; s < d.L; ++s) { put(sage_wl(pm, 0, s)); put(sage_bl(pm, 0, s, d.H)); put(sage_wr(pm, 0, s, d.H)); }
  for (int h = 0; h < 2; ++h) {
    for (int s = d.L; s < d.Lt; ++s) { put(sage_wl(pm, h, s)); put(sage_bl(pm, h, s, d.H)); put(sage_wr(pm, h, s, d.H)); }
    put(pm.lin_w[h]); put(pm.lin_b[h]); put(pm.v0_w[h]); put(pm.v0_b[h]); put(pm.v1_w[h]); put(pm.v1_b[h]);
  }
  return n;
}

int hexgnn_csr_layout(int64_t N, int64_t E, int64_t B, HexCsrLayout* out) {
  if (!out || N < 0 || E < 0 || B < 0) { set_error("csr_layout: bad argument"); return HEXGNN_E_BADARG; }
  csr_layout(N, E, B, out);
  return HEXGNN_OK;
}

int hexgnn_sizes(const HexModelDesc* desc, int64_t N, int64_t E, int64_t B, int32_t tile_rows, HexSizes* out) {
  Dims d;
  if (!make_dims(desc, &d) || !out || N < 0 || E < 0 || B < 0) { set_error("sizes: bad argument"); return HEXGNN_E_BADARG; }
  if (!tile_rows_ok(tile_rows) || d.NJ > 8) {
    set_error("sizes: unsupported tile_rows=%d or hidden=%d", tile_rows, d.H);
    return HEXGNN_E_UNSUPPORTED;
  }
  HexCsrLayout l;
  csr_layout(N, E, B, &l);
  GradLayout gl;
  grad_layout(d, &gl);
  out->csr_bytes = l.total_words * 4;
  out->fwd_ws_bytes = (pack_floats(d) > tc_pack_floats(d) ? pack_floats(d) : tc_pack_floats(d)) * 4;
  const int64_t simt_stash = (int64_t)d.Lt * N * d.Hp * 4;
  const int64_t simt_bwd = (pack_floats(d) + (int64_t)(bwd_grid_ctas() + 1) * gl.total) * 4;
  const bool tc = tc_bwd_supported(d);
  const int64_t tc_stash = tc ? tc_stash_bytes(d, N, B) : 0, tc_bwd = tc ? tc_bwd_ws(d, B).total * 4 : 0;
  out->stash_bytes = simt_stash > tc_stash ? simt_stash : tc_stash;
  out->bwd_ws_bytes = simt_bwd > tc_bwd ? simt_bwd : tc_bwd;
  out->hidden_padded = d.Hp;
  return HEXGNN_OK;
}

int hexgnn_build_csr(const int64_t* edge_index, int64_t E, const int64_t* batch, const int64_t* ptr, int64_t N,
                     int64_t B, int32_t tile_rows, void* csr, size_t csr_bytes, void* stream) {
  if (N < 0 || E < 0 || B < 1 || !csr || (E > 0 && !edge_index)) { set_error("build_csr: bad argument"); return HEXGNN_E_BADARG; }
  if (N >= (1ll << 31) - 8 || E >= (1ll << 31) - 8) { set_error("build_csr: N/E exceed int32"); return HEXGNN_E_UNSUPPORTED; }
  if (!batch && !ptr && B != 1) { set_error("build_csr: B must be 1 when neither batch nor ptr is given"); return HEXGNN_E_BADARG; }
  if (!tile_rows_ok(tile_rows)) { set_error("build_csr: unsupported tile_rows=%d", tile_rows); return HEXGNN_E_UNSUPPORTED; }
  HexCsrLayout l;
  csr_layout(N, E, B, &l);
  if ((int64_t)csr_bytes < l.total_words * 4) { set_error("build_csr: csr blob too small (%zu < %lld)", csr_bytes, (long long)l.total_words * 4); return HEXGNN_E_WORKSPACE; }
  return launch_build_csr(edge_index, E, batch, ptr, N, B, tile_rows, csr, (cudaStream_t)stream);
}

int hexgnn_forward(const HexModelDesc* desc, const float* params, const float* x, int64_t N, int64_t E, int64_t B,
                   const void* csr, int32_t tile_rows, int32_t flags, float* out_nodes, float* out_value,
                   float* stash, void* ws, size_t ws_bytes, void* stream) {
  Dims d;
  if (!make_dims(desc, &d)) { set_error("forward: bad model descriptor"); return HEXGNN_E_BADARG; }
  if (!params || !x || !csr || !out_nodes || !ws || N < 1 || B < 1 || E < 0) { set_error("forward: bad argument"); return HEXGNN_E_BADARG; }
  if ((flags & HEXGNN_SEPARATE) && !out_value) { set_error("forward: SEPARATE needs out_value"); return HEXGNN_E_BADARG; }
  if ((flags & HEXGNN_TF32X1) && (stash || !use_tc(d, flags))) { set_error("forward: TF32X1 is an inference mode of the tensor-core path (no stash)"); return HEXGNN_E_BADARG; }
  if (!tile_rows_ok(tile_rows) || d.NJ > 8) { set_error("forward: unsupported tile_rows=%d or hidden=%d", tile_rows, d.H); return HEXGNN_E_UNSUPPORTED; }
  const int64_t need = (pack_floats(d) > tc_pack_floats(d) ? pack_floats(d) : tc_pack_floats(d)) * 4;
  if ((int64_t)ws_bytes < need) { set_error("forward: workspace too small (%zu < %lld)", ws_bytes, (long long)need); return HEXGNN_E_WORKSPACE; }
  ParamMap pm;
  make_param_map(d, &pm);
  cudaStream_t st = (cudaStream_t)stream;
  if (!use_tc(d, flags)) {
    PackView pv = pack_view(ws, d);
    int rc = launch_pack(d, pm, params, x, pv, false, st);
    if (rc) return rc;
    return launch_forward_simt(d, pm, params, x, N, E, B, csr, tile_rows, flags, out_nodes, out_value, stash, pv, st);
  }
  // default: tcgen05 path (3xTF32, fp32 parity)
  TcPackView tpv = tc_pack_view(ws, d);
  int rc = launch_pack_tc(d, pm, params, x, tpv, false, st);
  if (rc) return rc;
  return launch_forward_tc(d, pm, params, x, N, E, B, csr, flags, out_nodes, out_value, stash, tpv, st);
}

static int backward_impl(const HexModelDesc* desc, const float* params, const float* x, int64_t N, int64_t E, int64_t B,
                         const void* csr, int32_t tile_rows, int32_t flags, const float* d_nodes, const float* d_value,
                         const float* stash, float* grad_params, float* d_embeds, float* d_x, void* ws, size_t ws_bytes,
                         void* stream);

int hexgnn_backward(const HexModelDesc* desc, const float* params, const float* x, int64_t N, int64_t E, int64_t B,
                    const void* csr, int32_t tile_rows, int32_t flags, const float* d_nodes, const float* d_value,
                    const float* stash, float* grad_params, float* d_embeds, void* ws, size_t ws_bytes,
                    void* stream) {
  return backward_impl(desc, params, x, N, E, B, csr, tile_rows, flags, d_nodes, d_value, stash, grad_params, d_embeds,
                       nullptr, ws, ws_bytes, stream);
}

int hexgnn_backward_dx(const HexModelDesc* desc, const float* params, const float* x, int64_t N, int64_t E, int64_t B,
                       const void* csr, int32_t tile_rows, int32_t flags, const float* d_nodes, const float* d_value,
                       const float* stash, float* grad_params, float* d_embeds, float* d_x, void* ws, size_t ws_bytes,
                       void* stream) {
  if (!d_x) { set_error("backward_dx: d_x is NULL (use hexgnn_backward)"); return HEXGNN_E_BADARG; }
  return backward_impl(desc, params, x, N, E, B, csr, tile_rows, flags, d_nodes, d_value, stash, grad_params, d_embeds,
                       d_x, ws, ws_bytes, stream);
}

static int backward_impl(const HexModelDesc* desc, const float* params, const float* x, int64_t N, int64_t E, int64_t B,
                         const void* csr, int32_t tile_rows, int32_t flags, const float* d_nodes, const float* d_value,
                         const float* stash, float* grad_params, float* d_embeds, float* d_x, void* ws, size_t ws_bytes,
                         void* stream) {
  Dims d;
  if (!make_dims(desc, &d)) { set_error("backward: bad model descriptor"); return HEXGNN_E_BADARG; }
  if (!params || !x || !csr || !d_nodes || !stash || !grad_params || !ws || N < 1 || B < 1 || E < 0) { set_error("backward: bad argument"); return HEXGNN_E_BADARG; }
  if (!tile_rows_ok(tile_rows) || d.NJ > 8) { set_error("backward: unsupported tile_rows=%d or hidden=%d", tile_rows, d.H); return HEXGNN_E_UNSUPPORTED; }
  GradLayout gl;
  grad_layout(d, &gl);
  ParamMap pm;
  make_param_map(d, &pm);
  cudaStream_t st = (cudaStream_t)stream;
  if (use_tc(d, flags)) {
    const TcBwdWs w = tc_bwd_ws(d, B);
    if ((int64_t)ws_bytes < w.total * 4) { set_error("backward: workspace too small (%zu < %lld)", ws_bytes, (long long)w.total * 4); return HEXGNN_E_WORKSPACE; }
    float* wsf = (float*)ws;
    TcPackView tpv = tc_pack_view(wsf + w.pack, d);
    int* sync = (int*)(wsf + w.sync);
    int rc = launch_pack_tc(d, pm, params, x, tpv, true, st, sync, (int)tc_sync_ints(d, B));   // also zeroes the flags
    if (rc) return rc;
    const int grid = bwd_grid_ctas();
    {
      // The dH chain and the dW GEMM overlap (programmatic dependent launch, per-tile flags): ONE event pair around
      // the two launches -- an event between them would serialise them.
      ProfScope prof(1, st);
      rc = launch_backward_tc_chain(d, pm, params, x, N, E, B, csr, flags, d_nodes, d_value, stash, wsf + w.dz, d_embeds,
                                    tpv, wsf + w.small_rec, wsf + w.head_rec, sync, grid, st, nullptr, d_x);
      if (rc) return rc;
      rc = launch_backward_tc_dw(d, N, E, B, csr, stash, wsf + w.dz, wsf + w.big_rec, sync, st);
      if (rc) return rc;
    }
    rc = launch_reduce_tc(d, wsf + w.small_rec, csr_view(const_cast<void*>(csr), N, E, B).ntiles, wsf + w.big_rec, tc_dw_groups(d), wsf + w.head_rec, B, wsf + w.sum, st);
    if (rc) return rc;
    return launch_unpack_grads(d, pm, x, wsf + w.sum, grad_params, st);
  }
  if (d_x) { set_error("backward_dx: the input gradient is produced by the tensor-core path only"); return HEXGNN_E_UNSUPPORTED; }
  const int nparts = bwd_grid_ctas();
  const int64_t need = (pack_floats(d) + (int64_t)(nparts + 1) * gl.total) * 4;
  if ((int64_t)ws_bytes < need) { set_error("backward: workspace too small (%zu < %lld)", ws_bytes, (long long)need); return HEXGNN_E_WORKSPACE; }
  PackView pv = pack_view(ws, d);
  float* partials = (float*)ws + pack_floats(d);
  float* packed_sum = partials + (int64_t)nparts * gl.total;
  int rc = launch_pack(d, pm, params, x, pv, true, st);
  if (rc) return rc;
  rc = launch_backward_simt(d, pm, params, x, N, E, B, csr, tile_rows, flags, d_nodes, d_value, stash, d_embeds, pv,
                            partials, st);
  if (rc) return rc;
  CsrView cv = csr_view(const_cast<void*>(csr), N, E, B);
  return launch_reduce_grads(d, pm, x, partials, nparts, cv.ntiles, packed_sum, grad_params, st);
}

int hexgnn_train_step(const HexModelDesc* desc, const float* params, const float* x, int64_t N, int64_t E, int64_t B,
                      const void* csr, int32_t tile_rows, int32_t flags, const int64_t* ptr, const int64_t* action,
                      const float* target, const float* weight, int32_t kind, float scale, float* q, float* loss,
                      float* td_err, float* stash, float* grad_params, void* ws, size_t ws_bytes, void* stream) {
  Dims d;
  if (!make_dims(desc, &d)) { set_error("train_step: bad model descriptor"); return HEXGNN_E_BADARG; }
  if (!params || !x || !csr || !ptr || !action || !target || !q || !loss || !stash || !grad_params || !ws ||
      N < 1 || B < 1 || E < 0) { set_error("train_step: bad argument"); return HEXGNN_E_BADARG; }
  if (kind != HEXGNN_LOSS_MSE && kind != HEXGNN_LOSS_SMOOTH_L1) { set_error("train_step: unknown loss kind %d", kind); return HEXGNN_E_BADARG; }
  if (flags & (HEXGNN_ADV_ONLY | HEXGNN_SEPARATE | HEXGNN_PATH_SIMT)) { set_error("train_step: only the default Q output on the tensor-core path"); return HEXGNN_E_UNSUPPORTED; }
  if (!tile_rows_ok(tile_rows) || !tc_bwd_supported(d)) { set_error("train_step: unsupported tile_rows=%d or hidden=%d", tile_rows, d.H); return HEXGNN_E_UNSUPPORTED; }
  const TcBwdWs w = tc_bwd_ws(d, B);
  if ((int64_t)ws_bytes < w.total * 4) { set_error("train_step: workspace too small (%zu < %lld)", ws_bytes, (long long)w.total * 4); return HEXGNN_E_WORKSPACE; }
  ParamMap pm;
  make_param_map(d, &pm);
  cudaStream_t st = (cudaStream_t)stream;
  float* wsf = (float*)ws;
  int* sync = (int*)(wsf + w.sync);
  // everything the three overlapped kernels need besides each other's tiles is produced first
  TcPackView tpv = tc_pack_view(wsf + w.pack, d);   // one pack launch writes the forward AND the backward images
  int rc = launch_pack_tc(d, pm, params, x, tpv, true, st, sync, (int)tc_sync_ints(d, B));   // ... and zeroes the flags
  if (rc) return rc;
  const int grid = bwd_grid_ctas();
  int* fwd_flag = sync + tc_sync_fwd_off(d, B);
  const TdFuse fuse{action, target, weight, q, fwd_flag, kind, scale};
  {
    // forward -> dH chain -> dW GEMM, each a programmatic dependent launch of the one before, handing tiles over
    // through flags: ONE event pair around the three (slot 2)
    ProfScope prof(2, st);
    ++g_prof_mute;
    rc = launch_forward_tc(d, pm, params, x, N, E, B, csr, flags, q, nullptr, stash, tpv, st, fwd_flag);
    if (!rc) rc = launch_backward_tc_chain(d, pm, params, x, N, E, B, csr, flags, nullptr, nullptr, stash, wsf + w.dz, nullptr,
                                           tpv, wsf + w.small_rec, wsf + w.head_rec, sync, grid, st, &fuse);
    if (!rc) rc = launch_backward_tc_dw(d, N, E, B, csr, stash, wsf + w.dz, wsf + w.big_rec, sync, st);
    --g_prof_mute;
    if (rc) return rc;
  }
  // the loss value and the TD errors are one more block of the reduction kernel
  const TdLossArgs tdl{q, ptr, action, target, weight, N, (int)B, kind, scale, loss, td_err, nullptr};
  rc = launch_reduce_tc(d, wsf + w.small_rec, csr_view(const_cast<void*>(csr), N, E, B).ntiles, wsf + w.big_rec, tc_dw_groups(d), wsf + w.head_rec, B, wsf + w.sum, st, &tdl);
  if (rc) return rc;
  return launch_unpack_grads(d, pm, x, wsf + w.sum, grad_params, st);
}

int64_t hexgnn_collate_ws_bytes(int64_t B) { return B < 0 ? 0 : collate_ws_bytes(B); }

int hexgnn_collate(const float* const* x_ptrs, const int64_t* const* ei_ptrs, const int64_t* n_nodes, const int64_t* n_edges,
                   int64_t B, float* x_out, int64_t* ei_out, int64_t* batch_out, int64_t* ptr_out, void* ws, size_t ws_bytes,
                   void* host_ws, size_t host_ws_bytes, void* stream) {
  if (!x_ptrs || !ei_ptrs || !n_nodes || !n_edges || B < 1 || !x_out || !batch_out || !ptr_out || !ws || !host_ws) { set_error("collate: bad argument"); return HEXGNN_E_BADARG; }
  if ((int64_t)ws_bytes < collate_ws_bytes(B) || (int64_t)host_ws_bytes < collate_ws_bytes(B)) { set_error("collate: workspace too small"); return HEXGNN_E_WORKSPACE; }
  int64_t E = 0;
  for (int64_t g = 0; g < B; ++g) {
    if (n_nodes[g] < 0 || n_edges[g] < 0 || (n_nodes[g] > 0 && !x_ptrs[g]) || (n_edges[g] > 0 && !ei_ptrs[g])) { set_error("collate: bad graph %lld", (long long)g); return HEXGNN_E_BADARG; }
    E += n_edges[g];
  }
  if (E > 0 && !ei_out) { set_error("collate: ei_out is NULL"); return HEXGNN_E_BADARG; }
  return launch_collate(x_ptrs, ei_ptrs, n_nodes, n_edges, B, x_out, ei_out, batch_out, ptr_out, ws, host_ws, (cudaStream_t)stream);
}

int64_t hexgnn_adam_ws_bytes(void) { return (int64_t)adam_ws_floats() * 4; }

int hexgnn_adam_step(const HexModelDesc* desc, float* params, float* grads, float* exp_avg, float* exp_avg_sq,
                     int32_t* state, const float* x, float lr, float beta1, float beta2, float eps, float max_grad_norm,
                     int32_t flags, float* grad_norm, void* ws, size_t ws_bytes, void* stream) {
  Dims d;
  if (!make_dims(desc, &d)) { set_error("adam_step: bad model descriptor"); return HEXGNN_E_BADARG; }
  if (!params || !grads || !exp_avg || !exp_avg_sq || !state || !x || !ws) { set_error("adam_step: bad argument"); return HEXGNN_E_BADARG; }
  if (!(lr >= 0.f) || !(beta1 >= 0.f && beta1 < 1.f) || !(beta2 >= 0.f && beta2 < 1.f) || !(eps >= 0.f)) { set_error("adam_step: bad hyper-parameter"); return HEXGNN_E_BADARG; }
  if ((int64_t)ws_bytes < hexgnn_adam_ws_bytes()) { set_error("adam_step: workspace too small"); return HEXGNN_E_WORKSPACE; }
  ParamMap pm;
  make_param_map(d, &pm);
  return launch_adam_step(d, pm, params, grads, exp_avg, exp_avg_sq, state, x, lr, beta1, beta2, eps, max_grad_norm,
                          (flags & HEXGNN_ADAM_SKIP_UNTAKEN) ? 1 : 0, grad_norm, (float*)ws, (cudaStream_t)stream);
}

int hexgnn_noisy_compose(const HexModelDesc* desc, float* params, const float* noisy, void* stream) {
  Dims d;
  if (!make_dims(desc, &d)) { set_error("noisy_compose: bad model descriptor"); return HEXGNN_E_BADARG; }
  if (!params || !noisy) { set_error("noisy_compose: bad argument"); return HEXGNN_E_BADARG; }
  ParamMap pm;
  make_param_map(d, &pm);
  return launch_noisy_compose(d, pm, params, noisy, (cudaStream_t)stream);
}

int hexgnn_noisy_grads(const HexModelDesc* desc, const float* grads, const float* noisy, float* d_noisy, int32_t accumulate,
                       void* stream) {
  Dims d;
  if (!make_dims(desc, &d)) { set_error("noisy_grads: bad model descriptor"); return HEXGNN_E_BADARG; }
  if (!grads || !noisy || !d_noisy) { set_error("noisy_grads: bad argument"); return HEXGNN_E_BADARG; }
  ParamMap pm;
  make_param_map(d, &pm);
  return launch_noisy_grads(d, pm, grads, noisy, d_noisy, accumulate ? 1 : 0, (cudaStream_t)stream);
}

int hexgnn_td_loss(const float* q, const int64_t* ptr, const int64_t* action, const float* target, const float* weight,
                   int64_t N, int64_t B, int32_t kind, float scale, float* loss, float* td_err, float* d_q, void* stream) {
  if (!q || !ptr || !action || !target || !loss || N < 1 || B < 1) { set_error("td_loss: bad argument"); return HEXGNN_E_BADARG; }
  if (kind != HEXGNN_LOSS_MSE && kind != HEXGNN_LOSS_SMOOTH_L1) { set_error("td_loss: unknown loss kind %d", kind); return HEXGNN_E_BADARG; }
  if (B >= (1ll << 31)) { set_error("td_loss: B exceeds int32"); return HEXGNN_E_UNSUPPORTED; }
  return launch_td_loss(q, ptr, action, target, weight, N, B, kind, scale, loss, td_err, d_q, (cudaStream_t)stream);
}

int hexgnn_forward_big(const HexModelDesc* desc, const float* params, const float* x, int64_t N, int64_t E, int64_t B,
                       const void* csr, int32_t tile_rows, int32_t flags, int32_t max_rows, float* out_nodes,
                       float* out_value, void* ws, size_t ws_bytes, void* stream) {
  Dims d;
  if (!make_dims(desc, &d)) { set_error("forward_big: bad model descriptor"); return HEXGNN_E_BADARG; }
  if (!params || !x || !csr || !out_nodes || !ws || N < 1 || B < 1 || E < 0 || max_rows < 1) { set_error("forward_big: bad argument"); return HEXGNN_E_BADARG; }
  if ((flags & HEXGNN_SEPARATE) && !out_value) { set_error("forward_big: SEPARATE needs out_value"); return HEXGNN_E_BADARG; }
  if (!tile_rows_ok(tile_rows) || max_rows > 65535 * 32 || B > 0x7fffffff) { set_error("forward_big: unsupported tile_rows / max_rows"); return HEXGNN_E_UNSUPPORTED; }
  const int64_t need = (2 * N * d.Hp + pack_floats(d)) * 4;
  if ((int64_t)ws_bytes < need) { set_error("forward_big: workspace too small (%zu < %lld)", ws_bytes, (long long)need); return HEXGNN_E_WORKSPACE; }
  ParamMap pm;
  make_param_map(d, &pm);
  return launch_forward_big(d, pm, params, x, N, E, B, csr, tile_rows, flags, max_rows, out_nodes, out_value, (float*)ws,
                            (cudaStream_t)stream);
}

int hexgnn_allreduce_mc(float* mc_buf, int32_t* const* peer_flags, int32_t rank, int32_t world, int64_t n, int32_t epoch,
                        int32_t* counter, void* stream) {
  if (!mc_buf || !peer_flags || !counter || world < 1 || world > 64 || rank < 0 || rank >= world || n < 1 || epoch < 1) {
    set_error("allreduce_mc: bad argument");
    return HEXGNN_E_BADARG;
  }
  if (((uintptr_t)mc_buf & 15) != 0 || n % (4 * (int64_t)world) != 0) {
    set_error("allreduce_mc: the buffer must be 16-byte aligned and n a multiple of 4 * world (pad it)");
    return HEXGNN_E_BADARG;
  }
  return launch_allreduce_mc(mc_buf, (int* const*)peer_flags, rank, world, n, epoch, (int*)counter, (cudaStream_t)stream);
}

int hexgnn_allreduce_p2p(const float* const* peer_bufs, int32_t world, int64_t n, float* out, void* stream) {
  if (!peer_bufs || !out || world < 1 || world > 64 || n < 1) { set_error("allreduce_p2p: bad argument"); return HEXGNN_E_BADARG; }
  if (((uintptr_t)out & 15) != 0) { set_error("allreduce_p2p: out must be 16-byte aligned"); return HEXGNN_E_BADARG; }
  return launch_allreduce_p2p(peer_bufs, world, n, out, (cudaStream_t)stream);
}

int hexgnn_mask_argmax(float* q, const int64_t* ptr, int64_t N, int64_t B, int32_t mask_terminals, float mask_value,
                       float* out_max, int64_t* out_arg, void* stream) {
  if (!q || !ptr || N < 1 || B < 1 || B >= (1ll << 26)) { set_error("mask_argmax: bad argument"); return HEXGNN_E_BADARG; }
  return launch_mask_argmax(q, ptr, N, B, mask_terminals, mask_value, out_max, out_arg, (cudaStream_t)stream);
}

int hexgnn_td_target(const float* q_target, const int64_t* best, const float* reward, const float* done, float gamma_n,
                     int64_t N, int64_t B, float* target, void* stream) {
  if (!q_target || !best || !reward || !done || !target || N < 1 || B < 1 || B >= (1ll << 31)) { set_error("td_target: bad argument"); return HEXGNN_E_BADARG; }
  return launch_td_target(q_target, best, reward, done, gamma_n, N, B, target, (cudaStream_t)stream);
}

int hexgnn_stash_embed(const HexModelDesc* desc, int64_t N, int64_t B, int32_t flags, int64_t* offset_floats,
                       int64_t* row_pitch) {
  Dims d;
  if (!make_dims(desc, &d) || !offset_floats || !row_pitch || N < 0 || B < 0) { set_error("stash_embed: bad argument"); return HEXGNN_E_BADARG; }
  *row_pitch = d.Hp;
  *offset_floats = use_tc(d, flags) ? (int64_t)d.Lt * B * tc_act_blocks(d.Hp) * (kImgBlockBytes / 4) : (int64_t)(d.L - 1) * N * d.Hp;
  return HEXGNN_OK;
}

int hexgnn_selftest_mma(const float* A, const float* B, float* D, int32_t N, int32_t K, int32_t modes, void* stream) {
  if (!A || !B || !D) { set_error("selftest: null pointer"); return HEXGNN_E_BADARG; }
  return launch_tc_selftest(A, B, D, N, K, modes, (cudaStream_t)stream);
}

int hexgnn_debug_phases(int32_t which, int64_t* host_out, int32_t n) {
  if (!host_out || n < 1) { set_error("debug_phases: bad argument"); return HEXGNN_E_BADARG; }
  if (which == 0) return debug_read_fwd_phases((long long*)host_out, n);
  if (which == 1) return debug_read_bwd_phases((long long*)host_out, n);
  if (which == 2) return debug_read_lw_phases((long long*)host_out, n);
  set_error("debug_phases: unknown kernel %d", which);
  return HEXGNN_E_BADARG;
}

int hexgnn_profile_enable(int on) {
  g_prof.on = on != 0;
  if (!on) g_prof.used[0] = g_prof.used[1] = g_prof.used[2] = 0;
  return HEXGNN_OK;
}

int hexgnn_profile_read(int which, double* total_ms, int64_t* launches) {
  if (which < 0 || which > 2 || !total_ms || !launches) { set_error("profile_read: bad argument"); return HEXGNN_E_BADARG; }
  double tot = 0.0;
  const size_t u = g_prof.used[which];
  for (size_t i = 0; i + 1 < u; i += 2) {
    HEX_CUDA_CHECK(cudaEventSynchronize(g_prof.ev[which][i + 1]));
    float ms = 0.f;
    HEX_CUDA_CHECK(cudaEventElapsedTime(&ms, g_prof.ev[which][i], g_prof.ev[which][i + 1]));
    tot += ms;
  }
  *total_ms = tot;
  *launches = (int64_t)(u / 2);
  g_prof.used[which] = 0;
  return HEXGNN_OK;
}

}  // extern "C"
