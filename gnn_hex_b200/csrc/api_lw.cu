// C-ABI of the layer-at-a-time tensor-core path (declared in include/hexgnn.h; kernels in lw_*.cu).
#include "lw_model.cuh"

using namespace hexgnn;

extern "C" {

int hexgnn_lw_sizes(const HexModelDesc* desc, int64_t N, int64_t E, int64_t B, int32_t norm, HexLwSizes* out) {
  Dims d;
  if (!make_dims(desc, &d) || !out || N < 0 || E < 0 || B < 0) { set_error("lw_sizes: bad argument"); return HEXGNN_E_BADARG; }
  if (d.Hp > kLwMaxK) { set_error("lw_sizes: hidden=%d not supported (<= %d)", d.H, kLwMaxK); return HEXGNN_E_UNSUPPORTED; }
  out->fwd_ws_bytes = lw_model_ws(d, N, false).total;
  out->bwd_ws_bytes = lw_bwd_ws_bytes(d, N, B, norm != 0);
  out->stash_bytes = lw_stash_floats(d, N, norm != 0) * 4;
  out->norm_floats = lw_norm_floats(d);
  out->embed_offset_floats = norm ? (int64_t)2 * d.Lt * N * d.Hp : (int64_t)(d.L - 1) * N * d.Hp;
  out->hidden_padded = d.Hp;
  return HEXGNN_OK;
}

int hexgnn_forward_lw(const HexModelDesc* desc, const float* params, const float* norm_params, const float* x, int64_t N,
                      int64_t E, int64_t B, const void* csr, int32_t flags, float* out_nodes, float* out_value, float* stash,
                      void* ws, size_t ws_bytes, void* stream) {
  Dims d;
  if (!make_dims(desc, &d)) { set_error("forward_lw: bad model descriptor"); return HEXGNN_E_BADARG; }
  if (!params || !x || !csr || !out_nodes || !ws || N < 1 || B < 1 || E < 0) { set_error("forward_lw: bad argument"); return HEXGNN_E_BADARG; }
  if ((flags & HEXGNN_SEPARATE) && !out_value) { set_error("forward_lw: SEPARATE needs out_value"); return HEXGNN_E_BADARG; }
  if (d.Hp > kLwMaxK || B > 0x7fffffff) { set_error("forward_lw: unsupported hidden=%d", d.H); return HEXGNN_E_UNSUPPORTED; }
  const int64_t need = lw_model_ws(d, N, false).total;
  if ((int64_t)ws_bytes < need) { set_error("forward_lw: workspace too small (%zu < %lld)", ws_bytes, (long long)need); return HEXGNN_E_WORKSPACE; }
  ParamMap pm;
  make_param_map(d, &pm);
  return launch_forward_lw(d, pm, params, norm_params, x, N, E, B, csr, flags, out_nodes, out_value, stash, ws,
                           (cudaStream_t)stream);
}

int hexgnn_backward_lw(const HexModelDesc* desc, const float* params, const float* norm_params, const float* x, int64_t N,
                       int64_t E, int64_t B, const void* csr, int32_t flags, const float* d_nodes, const float* d_value,
                       const float* stash, float* grad_params, float* grad_norm, float* d_embeds, void* ws, size_t ws_bytes,
                       void* stream) {
  Dims d;
  if (!make_dims(desc, &d)) { set_error("backward_lw: bad model descriptor"); return HEXGNN_E_BADARG; }
  if (!params || !x || !csr || !d_nodes || !stash || !grad_params || !ws || N < 1 || B < 1 || E < 0 ||
      (norm_params && !grad_norm)) { set_error("backward_lw: bad argument"); return HEXGNN_E_BADARG; }
  if (d.Hp > kLwMaxK || B > 0x7fffffff) { set_error("backward_lw: unsupported hidden=%d", d.H); return HEXGNN_E_UNSUPPORTED; }
  const int64_t need = lw_bwd_ws_bytes(d, N, B, norm_params != nullptr);
  if ((int64_t)ws_bytes < need) { set_error("backward_lw: workspace too small (%zu < %lld)", ws_bytes, (long long)need); return HEXGNN_E_WORKSPACE; }
  ParamMap pm;
  make_param_map(d, &pm);
  return launch_backward_lw(d, pm, params, norm_params, x, N, E, B, csr, flags, d_nodes, d_value, stash, grad_params,
                            grad_norm, d_embeds, ws, (cudaStream_t)stream);
}

}  // extern "C"
