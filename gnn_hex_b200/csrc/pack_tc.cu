// Weight images for the tcgen05 kernels (layout documented in sage_tc.cuh): torch layout -> per layer,
// per 16-wide k chunk, a tf32-hi half and a tf32-lo half in the K-major SWIZZLE_NONE canonical form the
// tensor core reads, so the kernels fetch each ring stage with ONE contiguous bulk copy.
#include "sage_tc.cuh"

namespace hexgnn {

__global__ void pack_tc_kernel(Dims d, ParamMap pm, const float* __restrict__ params, const float* __restrict__ x,
                               TcPackView pv, int with_transpose) {
  const int head = (x[2] == 1.0f) ? 0 : 1;
  const int Hp = d.Hp, H = d.H, N2 = 2 * d.Hp;
  const int NQ = Hp / kTcKC;
  const int64_t half_f = (int64_t)kTcKC * N2;     // floats per half
  const int64_t layer_f = 2 * half_f * NQ;        // floats per layer image
  const int64_t per_layer = (int64_t)Hp * N2;     // logical elements per layer
  const int64_t total = per_layer * (d.Lt - 1);
  const int64_t nth = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += nth) {
    const int s = (int)(i / per_layer) + 1;
    const int r = (int)(i % per_layer);
    const int n = r / Hp, k = r % Hp;   // k fastest: coalesced reads of W[c][k]
    // forward: B[n][k] = (n < Hp ? lin_l.W : lin_r.W)[c = n % Hp][k]
    float val = 0.f;
    const int c = n < Hp ? n : n - Hp;
    if (c < H && k < H)
      val = params[(n < Hp ? sage_wl(pm, head, s) : sage_wr(pm, head, s, H)) + (int64_t)c * H + k];
    float hi, lo;
    tc::split_tf32(val, hi, lo);
    {
      const int q = k / kTcKC, kq = k % kTcKC;
      const int64_t off = (int64_t)(s - 1) * layer_f + (int64_t)q * 2 * half_f + (int64_t)(kq / 4) * (N2 * 4) +
                          (n / 8) * 32 + (n % 8) * 4 + (kq % 4);
      pv.wimg[off] = hi;
      pv.wimg[off + half_f] = lo;
    }
    if (with_transpose) {
      // backward: dH[i][k] = sum_c dZ_which[i][c] * W_which[c][k]  =>  B[n = k][kk = c], which 0 = lin_r, 1 = lin_l
      const int which = n < Hp ? 1 : 0;
      const int64_t halfB = (int64_t)kTcKC * Hp;                  // floats per half of a (which, q) stage
      const int q = c / kTcKC, cq = c % kTcKC;
      const int64_t off = (((int64_t)(s - 1) * 2 + which) * NQ + q) * 2 * halfB + (int64_t)(cq / 4) * (Hp * 4) +
                          (k / 8) * 32 + (k % 8) * 4 + (cq % 4);
      pv.wimgT[off] = hi;
      pv.wimgT[off + halfB] = lo;
    }
  }
  const int nb = d.Lt * Hp;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nb; i += nth) {
    const int s = (int)(i / Hp), c = (int)(i % Hp);
    pv.bias[i] = c < H ? params[sage_bl(pm, head, s, H) + c] : 0.f;
  }
}

int launch_pack_tc(const Dims& d, const ParamMap& pm, const float* params, const float* x, TcPackView pv,
                   bool with_transpose, cudaStream_t st) {
  const int64_t total = (int64_t)(d.Lt - 1) * d.Hp * 2 * d.Hp;
  int blocks = (int)((total + 255) / 256);
  if (blocks < 1) blocks = 1;
  if (blocks > 1184) blocks = 1184;
  pack_tc_kernel<<<blocks, 256, 0, st>>>(d, pm, params, x, pv, with_transpose ? 1 : 0);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

// Fixed-order reduction of the tcgen05 backward's partial records into one packed record (GradLayout):
//   gw and gb[s>=1]  <- the G group records of sage_dw_tc_kernel   (big_rec [G][gl.g0])
//   everything else  <- the per-CTA records of sage_bwd_tc_kernel  (small_rec [nsmall][gl.total - gl.gb])
__global__ void reduce_tc_kernel(GradLayout gl, int Hp, const float* __restrict__ small_rec, int nsmall,
                                 const float* __restrict__ big_rec, int G, float* __restrict__ out) {
  const int64_t nth = (int64_t)gridDim.x * blockDim.x;
  const int64_t recn = gl.total - gl.gb;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < gl.total; i += nth) {
    float acc = 0.f;
    const bool from_big = i < gl.gb || (i >= gl.gb + Hp && i < gl.g0);
    if (from_big) {
      for (int g = 0; g < G; ++g) acc += big_rec[(int64_t)g * gl.g0 + i];
    } else {
      for (int p = 0; p < nsmall; ++p) acc += small_rec[(int64_t)p * recn + (i - gl.gb)];
    }
    out[i] = acc;
  }
}

int launch_reduce_tc(const Dims& d, const float* small_rec, int nsmall, const float* big_rec, int G, float* packed_sum,
                     cudaStream_t st) {
  GradLayout gl;
  grad_layout(d, &gl);
  int blocks = (int)((gl.total + 255) / 256);
  if (blocks > 1184) blocks = 1184;
  reduce_tc_kernel<<<blocks, 256, 0, st>>>(gl, d.Hp, small_rec, nsmall, big_rec, G, packed_sum);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

}  // namespace hexgnn
