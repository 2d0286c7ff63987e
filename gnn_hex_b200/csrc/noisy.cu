// FactorizedNoisyLinear on the advantage Linear(H, 1) of the two heads (`--noisy_dqn=True`):
// GN0/models.py:84-141 (the layer), :331-334 (HeadNetwork puts it in place of `self.linear`), :928-947 (factory args).
//
//   w = weight_mu + weight_sigma * weight_epsilon;  b = bias_mu + bias_sigma * bias_epsilon      (models.py:134-141)
//
// The fused kernels read the advantage weights from the flat parameter buffer, so the noisy layer is a composition step
// in front of them and a chain-rule step behind the backward:
//   compose:  params[linear.W | linear.b of head h] = mu + sigma * eps
//   grads:    d_mu = g,  d_sigma = g * eps           (g = gradient of the composed weights, from the flat gradient)
// `noisy` is [2 heads][3: mu, sigma, epsilon][H + 1] (weight row then the bias), `d_noisy` is [2][2: mu, sigma][H + 1].
// One launch each, both heads; element-wise, deterministic.
#include "common.cuh"

namespace hexgnn {

__global__ void noisy_compose_kernel(float* __restrict__ params, const float* __restrict__ noisy, int H, int64_t w0,
                                     int64_t b0, int64_t w1, int64_t b1) {
  const int h = blockIdx.x;
  const float* mu = noisy + (int64_t)h * 3 * (H + 1);
  const float* sg = mu + (H + 1);
  const float* ep = sg + (H + 1);
  const int64_t w = h ? w1 : w0, b = h ? b1 : b0;
  for (int i = threadIdx.x; i <= H; i += blockDim.x) {
    const float v = mu[i] + sg[i] * ep[i];      // two roundings, as torch's `mu + sigma * eps` (no fused multiply-add)
    params[i < H ? w + i : b] = v;
  }
}

__global__ void noisy_grads_kernel(const float* __restrict__ grads, const float* __restrict__ noisy,
                                   float* __restrict__ d_noisy, int accumulate, int H, int64_t w0, int64_t b0, int64_t w1,
                                   int64_t b1) {
  const int h = blockIdx.x;
  const float* ep = noisy + ((int64_t)h * 3 + 2) * (H + 1);
  float* dmu = d_noisy + (int64_t)h * 2 * (H + 1);
  float* dsg = dmu + (H + 1);
  const int64_t w = h ? w1 : w0, b = h ? b1 : b0;
  for (int i = threadIdx.x; i <= H; i += blockDim.x) {
    const float g = grads[i < H ? w + i : b];
    const float gs = __fmul_rn(g, ep[i]);
    if (accumulate) { dmu[i] += g; dsg[i] += gs; }
    else { dmu[i] = g; dsg[i] = gs; }
  }
}

int launch_noisy_compose(const Dims& d, const ParamMap& pm, float* params, const float* noisy, cudaStream_t st) {
  noisy_compose_kernel<<<2, 128, 0, st>>>(params, noisy, d.H, pm.lin_w[0], pm.lin_b[0], pm.lin_w[1], pm.lin_b[1]);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

int launch_noisy_grads(const Dims& d, const ParamMap& pm, const float* grads, const float* noisy, float* d_noisy,
                       int accumulate, cudaStream_t st) {
  noisy_grads_kernel<<<2, 128, 0, st>>>(grads, noisy, d_noisy, accumulate, d.H, pm.lin_w[0], pm.lin_b[0], pm.lin_w[1],
                                        pm.lin_b[1]);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

}  // namespace hexgnn
