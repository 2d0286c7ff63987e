// hexgnn_forward_static (declared in include/hexgnn.h): hexgnn_forward for inference with NO caller workspace and no
// allocation of any kind.  It has no counterpart in the reference; it was added for
// callers that cannot allocate (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// The one thing hexgnn_forward keeps in its workspace is scratch: the fp16-split tensor-core images of the SAGE weights and
// the padded bias rows, rebuilt from `params` by pack_tc_kernel in front of every forward.  Here that scratch is a
// file-scope __device__ array of HEXGNN_STATIC_WS_BYTES (every GPU that loads libhexgnn.so holds it until the process
// exits).  In host code the array's NAME is only a host-side stand-in for the device symbol, not a device address, so its
// address is looked up once per device with cudaGetSymbolAddress (that device current) and the pointer is what the kernels
// get.  All calls on a device share the array: it keeps the previous call's images, every byte the forward reads is
// rewritten by the pack kernel of the same call (nothing needs to start at zero), and two calls that overlap on the GPU
// (different streams, no dependency) race on it.
#include <mutex>

#include "common.cuh"
#include "sage_tc.cuh"

namespace hexgnn {
namespace {

static __device__ __align__(256) unsigned char g_static_ws[HEXGNN_STATIC_WS_BYTES];

constexpr int kMaxDevices = 64;
void* g_ws_dev[kMaxDevices];      // device address of g_static_ws per device ordinal (nullptr: not looked up yet)
std::mutex g_ws_mu;

// Device address of the scratch array on the CURRENT device (set-up: once per device).
int static_ws_current_device(unsigned char** out) {
  int dev = -1;
  HEX_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) { set_error("forward_static: device ordinal %d out of range", dev); return HEXGNN_E_UNSUPPORTED; }
  std::lock_guard<std::mutex> lock(g_ws_mu);
  if (!g_ws_dev[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_static_ws);
    if (e != cudaSuccess || !p) {
      set_error("forward_static: cudaGetSymbolAddress(static scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
      return HEXGNN_E_CUDA;
    }
    g_ws_dev[dev] = p;
  }
  *out = (unsigned char*)g_ws_dev[dev];
  return HEXGNN_OK;
}

// compact layout inside the scratch: forward images | bias rows (the transposed images of training are not needed)
inline int64_t static_img_bytes(const Dims& d) { return (int64_t)(d.Lt - 1) * 2 * (d.Hp / kTcKF) * (int64_t)d.Hp * 64; }
inline int64_t static_need_bytes(const Dims& d) { return align_up(static_img_bytes(d), 256) + (int64_t)d.Lt * d.Hp * 4; }

}  // namespace
}  // namespace hexgnn

using namespace hexgnn;

extern "C" int hexgnn_forward_static(const HexModelDesc* desc, const float* params, const float* x, int64_t N, int64_t E,
                                     int64_t B, const void* csr, int32_t tile_rows, int32_t flags, float* out_nodes,
                                     float* out_value, void* stream) {
  Dims d;
  if (!make_dims(desc, &d)) { set_error("forward_static: bad model descriptor"); return HEXGNN_E_BADARG; }
  if (!params || !x || !csr || !out_nodes || N < 1 || B < 1 || E < 0) { set_error("forward_static: bad argument"); return HEXGNN_E_BADARG; }
  if ((flags & HEXGNN_SEPARATE) && !out_value) { set_error("forward_static: SEPARATE needs out_value"); return HEXGNN_E_BADARG; }
  if (flags & HEXGNN_PATH_SIMT) { set_error("forward_static: tensor-core path only"); return HEXGNN_E_UNSUPPORTED; }
  if (tile_rows != 128 || d.NJ > 8 || !tc_bwd_supported(d)) {
    set_error("forward_static: unsupported tile_rows=%d or hidden=%d", tile_rows, d.H);
    return HEXGNN_E_UNSUPPORTED;
  }
  const int64_t need = static_need_bytes(d);
  if (need > (int64_t)HEXGNN_STATIC_WS_BYTES) {
    set_error("forward_static: model (hidden %d, %d + %d layers) needs %lld bytes of weight-image scratch, the static array "
              "holds HEXGNN_STATIC_WS_BYTES = %lld: use hexgnn_forward with a caller workspace",
              d.H, d.L, d.Lh, (long long)need, (long long)HEXGNN_STATIC_WS_BYTES);
    return HEXGNN_E_UNSUPPORTED;
  }
  unsigned char* ws = nullptr;
  int rc = static_ws_current_device(&ws);
  if (rc) return rc;
  TcPackView pv;
  pv.wimg = (float*)ws;
  pv.wimgT = nullptr;                                          // never touched without with_transpose
  pv.bias = (float*)(ws + align_up(static_img_bytes(d), 256));
  ParamMap pm;
  make_param_map(d, &pm);
  cudaStream_t st = (cudaStream_t)stream;
  rc = launch_pack_tc(d, pm, params, x, pv, false, st);
  if (rc) return rc;
  return launch_forward_tc(d, pm, params, x, N, E, B, csr, flags, out_nodes, out_value, nullptr, pv, st);
}
