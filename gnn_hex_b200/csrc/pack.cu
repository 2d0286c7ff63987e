// Weight packing (torch layout -> kernel layout) and deterministic gradient reduction/unpacking.
#include "common.cuh"

namespace hexgnn {

// wcat[s-1][k][c'] / wcatT[s-1][c'][k] for SAGE layers s>=1 of the head selected by x[0,2]
// (GN0/models.py:556: `if is_maker==1` -> maker_head else breaker_head), bias[s][c] for all s.
__global__ void pack_weights_kernel(Dims d, ParamMap pm, const float* __restrict__ params, const float* __restrict__ x,
                                    PackView pv, int with_transpose) {
  const int head = (x[2] == 1.0f) ? 0 : 1;
  const int Hp = d.Hp, H = d.H, W2 = 2 * d.Hp;
  const int64_t per_layer = (int64_t)Hp * W2;
  const int64_t total = per_layer * (d.Lt - 1);
  const int64_t nth = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += nth) {
    const int s = (int)(i / per_layer) + 1;
    const int r = (int)(i % per_layer);
    const int k = r / W2, c2 = r % W2;
    float val = 0.f;
    if (k < H) {
      if (c2 < Hp) {
        if (c2 < H) val = params[sage_wl(pm, head, s) + (int64_t)c2 * H + k];
      } else {
        const int c = c2 - Hp;
        if (c < H) val = params[sage_wr(pm, head, s, H) + (int64_t)c * H + k];
      }
    }
    pv.wcat[i] = val;
    if (with_transpose) pv.wcatT[(int64_t)(s - 1) * per_layer + (int64_t)c2 * Hp + k] = val;
  }
  const int nb = d.Lt * Hp;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nb; i += nth) {
    const int s = (int)(i / Hp), c = (int)(i % Hp);
    pv.bias[i] = c < H ? params[sage_bl(pm, head, s, H) + c] : 0.f;
  }
}

int launch_pack(const Dims& d, const ParamMap& pm, const float* params, const float* x, PackView pv, bool with_transpose,
                cudaStream_t st) {
  const int64_t total = (int64_t)(d.Lt - 1) * d.Hp * 2 * d.Hp;
  int blocks = (int)((total + 255) / 256);
  if (blocks < 1) blocks = 1;
  if (blocks > 1184) blocks = 1184;
  pack_weights_kernel<<<blocks, 256, 0, st>>>(d, pm, params, x, pv, with_transpose ? 1 : 0);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

// ---- gradient reduction --------------------------------------------------------------------------
// Stage 1: sum the per-CTA packed records in a fixed order (record 0, 1, 2, ...): deterministic.
__global__ void reduce_partials_kernel(const float4* __restrict__ partials, int64_t rec_vec4, int nparts_max,
                                       const int* __restrict__ ntiles, float4* __restrict__ out) {
  int np = *ntiles;
  if (np > nparts_max) np = nparts_max;
  const int64_t nth = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < rec_vec4; i += nth) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int p = 0; p < np; ++p) {
      const float4 v = partials[(int64_t)p * rec_vec4 + i];
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    out[i] = acc;
  }
}

// Stage 2: packed sum -> flat torch layout.  The head not taken gets zeros.
__global__ void unpack_grads_kernel(Dims d, ParamMap pm, GradLayout gl, const float* __restrict__ x,
                                    const float* __restrict__ sum, float* __restrict__ grad) {
  const int head = (x[2] == 1.0f) ? 0 : 1;
  const int H = d.H, Hp = d.Hp, W2 = 2 * d.Hp;
  const int64_t nth = (int64_t)gridDim.x * blockDim.x;
  const int64_t body0 = 5 * (int64_t)H;                 // layer 0: Wl[H,2], b[H], Wr[H,2]
  const int64_t bs = 2 * (int64_t)H * H + H;            // layers >= 1
  const int64_t body_end = body0 + (int64_t)(d.L - 1) * bs;
  const int64_t head_sz = (int64_t)d.Lh * bs + H + 1 + (int64_t)d.H2 * 4 * H + d.H2 + d.H2 + 1;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < pm.total; i += nth) {
    float val = 0.f;
    int64_t r = i;
    int s = -1;  // SAGE layer index if the element belongs to one
    if (r < body0) {
      // layer 0
      if (r < 2 * H) { const int c = (int)(r / 2), q = (int)(r % 2); val = sum[gl.g0 + q * Hp + c]; }
      else if (r < 3 * H) { val = sum[gl.gb + (r - 2 * H)]; }
      else { r -= 3 * H; const int c = (int)(r / 2), q = (int)(r % 2); val = sum[gl.g0 + (2 + q) * Hp + c]; }
    } else if (r < body_end) {
      r -= body0;
      s = 1 + (int)(r / bs);
      r %= bs;
    } else {
      r -= body_end;
      const int h = (int)(r / head_sz);
      r %= head_sz;
      if (h == head) {
        if (r < (int64_t)d.Lh * bs) { s = d.L + (int)(r / bs); r %= bs; }
        else {
          r -= (int64_t)d.Lh * bs;
          if (r < H) val = sum[gl.glin + r];
          else if (r < H + 1) val = sum[gl.glinb];
          else {
            r -= H + 1;
            const int64_t nv0 = (int64_t)d.H2 * 4 * H;
            if (r < nv0) val = sum[gl.gv0 + r];
            else if (r < nv0 + d.H2) val = sum[gl.gv0b + (r - nv0)];
            else if (r < nv0 + 2 * d.H2) val = sum[gl.gv1 + (r - nv0 - d.H2)];
            else val = sum[gl.gv1b];
          }
        }
      }
    }
    if (s >= 1) {
      const int64_t base = gl.gw + (int64_t)(s - 1) * Hp * W2;
      if (r < (int64_t)H * H) { const int c = (int)(r / H), k = (int)(r % H); val = sum[base + (int64_t)k * W2 + c]; }
      else if (r < (int64_t)H * H + H) { val = sum[gl.gb + (int64_t)s * Hp + (r - (int64_t)H * H)]; }
      else { r -= (int64_t)H * H + H; const int c = (int)(r / H), k = (int)(r % H); val = sum[base + (int64_t)k * W2 + Hp + c]; }
    }
    grad[i] = val;
  }
}

int launch_unpack_grads(const Dims& d, const ParamMap& pm, const float* x, const float* packed_sum, float* grad_params,
                        cudaStream_t st) {
  GradLayout gl;
  grad_layout(d, &gl);
  int b2 = (int)((pm.total + 255) / 256);
  if (b2 > 1184) b2 = 1184;
  unpack_grads_kernel<<<b2, 256, 0, st>>>(d, pm, gl, x, packed_sum, grad_params);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

int launch_reduce_grads(const Dims& d, const ParamMap& pm, const float* x, const float* partials, int nparts,
                        const int* ntiles, float* packed_sum, float* grad_params, cudaStream_t st) {
  GradLayout gl;
  grad_layout(d, &gl);
  const int64_t v4 = gl.total / 4;
  int blocks = (int)((v4 + 255) / 256);
  if (blocks > 1184) blocks = 1184;
  reduce_partials_kernel<<<blocks, 256, 0, st>>>((const float4*)partials, v4, nparts, ntiles, (float4*)packed_sum);
  int b2 = (int)((pm.total + 255) / 256);
  if (b2 > 1184) b2 = 1184;
  unpack_grads_kernel<<<b2, 256, 0, st>>>(d, pm, gl, x, packed_sum, grad_params);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

}  // namespace hexgnn
