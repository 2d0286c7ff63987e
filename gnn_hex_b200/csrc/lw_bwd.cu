// Backward of modern_two_headed on the layer-at-a-time path (see lw.cuh).
#include "lw_model.cuh"

namespace hexgnn {

int64_t lw_bwd_ws_bytes(const Dims& d, int64_t N, int64_t B, bool norm) { return lw_model_ws(d, N, true).total; }

int launch_backward_lw(const Dims& d, const ParamMap& pm, const float* params, const float* normp, const float* x, int64_t N,
                       int64_t E, int64_t B, const void* csr, int flags, const float* d_nodes, const float* d_value,
                       const float* stash, float* grad_params, float* grad_norm, float* d_embeds, void* ws, cudaStream_t st) {
  set_error("backward_lw: not built yet");
  return HEXGNN_E_UNSUPPORTED;
}

}  // namespace hexgnn
