// Learner optimizer step over the FLAT parameter / gradient buffers: global-norm gradient clipping + Adam in two
// launches (the reference runs nn.utils.clip_grad_norm_ and torch.optim.Adam over ~70 parameter tensors:
// GN0/RainbowDQN/Rainbow/common/rainbow.py:79 (Adam(lr, eps)), :320 (clip_grad_norm_), :325-326 (opt.step(); opt.zero_grad())).
//
//   total_norm = || g ||_2 over all parameters          (torch.nn.utils.clip_grad_norm_, norm_type 2)
//   clip       = min(1, max_norm / (total_norm + 1e-6)); g <- clip * g   (in place, as the reference does)
//   m <- b1 m + (1 - b1) g;  v <- b2 v + (1 - b2) g^2;  p <- p - (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
//
// torch.optim.Adam keeps one step counter per parameter and skips parameters whose .grad is None.  The two-headed
// network only ever gives gradients to the head selected by x[0,2] (models.py:556), so three counters live in `state`:
// body, maker head, breaker head.  A head's counter starts with the first step that takes it; from then on it advances
// every step (the reference's torch 1.13 `opt.zero_grad()` leaves ZERO gradients, not None: Adam keeps applying the
// momentum) unless HEXGNN_ADAM_SKIP_UNTAKEN is set (torch >= 2.0 `set_to_none=True` semantics: untouched while not taken).
// Deterministic: fixed-order two-level sum of squares, every block re-reduces the same partials.
// Guard (no reference counterpart; the reference has no size limit to trip over): a non-finite global norm -- e.g. the NaN
// record a tile poisons when a graph exceeds the tile capacity -- skips the WHOLE step: parameters, moments, the in-place
// clip and the step counters stay untouched, `grad_norm` reports the non-finite value.  The norm kernel therefore only
// PROPOSES the new counters (in the workspace); block 0 of the apply kernel commits them once the norm is known.
#include "common.cuh"

namespace hexgnn {

constexpr int kOptThreads = 256;
constexpr int kOptBlocks = 296;   // 2 x 148

// state (int32): [0] body steps, [1] maker-head steps, [2] breaker-head steps, [3] spare
__global__ void __launch_bounds__(kOptThreads) adam_norm_kernel(const float* __restrict__ g, int64_t n, const float* __restrict__ x,
                                                                const int* __restrict__ state, int skip_untaken,
                                                                float* __restrict__ partial, int* __restrict__ proposed) {
  __shared__ float red[kOptThreads];
  const int tid = threadIdx.x;
  const int64_t per = (n + gridDim.x - 1) / gridDim.x;
  const int64_t i0 = (int64_t)blockIdx.x * per, i1 = (i0 + per < n) ? i0 + per : n;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  int64_t i = i0 + tid;
  for (; i + 3 * kOptThreads < i1; i += 4 * kOptThreads) {
    const float v0 = g[i], v1 = g[i + kOptThreads], v2 = g[i + 2 * kOptThreads], v3 = g[i + 3 * kOptThreads];
    a0 = fmaf(v0, v0, a0); a1 = fmaf(v1, v1, a1); a2 = fmaf(v2, v2, a2); a3 = fmaf(v3, v3, a3);
  }
  for (; i < i1; i += kOptThreads) { const float v = g[i]; a0 = fmaf(v, v, a0); }
  red[tid] = (a0 + a1) + (a2 + a3);
  __syncthreads();
  for (int o = kOptThreads / 2; o > 0; o >>= 1) {
    if (tid < o) red[tid] += red[tid + o];
    __syncthreads();
  }
  if (tid == 0) partial[blockIdx.x] = red[0];
  if (blockIdx.x == 0 && tid == 0) {
    const int taken = (x[2] == 1.0f) ? 1 : 2;           // maker head : breaker head (models.py:556)
    proposed[0] = state[0] + 1;
    for (int h = 1; h <= 2; ++h) {
      int t = state[h];
      if (h == taken) t += 1;
      else if (t > 0 && !skip_untaken) t += 1;
      proposed[h] = t;
    }
    proposed[3] = taken;
  }
}

struct AdamArgs {
  float* p; float* g; float* m; float* v;
  int64_t n, seg1, seg2;          // [0, seg1) body, [seg1, seg2) maker head, [seg2, n) breaker head
  int* state;
  const int* proposed;            // counters of THIS step (adam_norm_kernel), committed to `state` when the norm is finite
  const float* partial; int nparts;
  float lr, beta1, beta2, eps, max_norm;
  int skip_untaken;
  float* grad_norm;
};

__global__ void __launch_bounds__(kOptThreads) adam_apply_kernel(const AdamArgs a) {
  __shared__ float red[kOptThreads];
  __shared__ float clip_s;
  __shared__ int ok_s;
  const int tid = threadIdx.x;
  float s = 0.f;
  for (int i = tid; i < a.nparts; i += kOptThreads) s += a.partial[i];   // same order in every block
  red[tid] = s;
  __syncthreads();
  for (int o = kOptThreads / 2; o > 0; o >>= 1) {
    if (tid < o) red[tid] += red[tid + o];
    __syncthreads();
  }
  if (tid == 0) {
    const float norm = sqrtf(red[0]);
    float c = 1.f;
    if (a.max_norm > 0.f) { c = a.max_norm / (norm + 1e-6f); if (c > 1.f) c = 1.f; }
    clip_s = c;
    ok_s = isfinite(norm) ? 1 : 0;
    if (blockIdx.x == 0 && a.grad_norm) a.grad_norm[0] = norm;
    if (blockIdx.x == 0 && ok_s) {
#pragma unroll
      for (int k = 0; k < 4; ++k) a.state[k] = a.proposed[k];   // nobody reads `state` in this launch
    }
  }
  __syncthreads();
  if (!ok_s) return;                                            // non-finite gradient: the step does not happen
  const float clip = clip_s;
  // per-segment scalars, computed in double like the Python side of torch.optim.Adam
  float step_size[3], bc2s[3];
  bool on[3];
  const int taken = a.proposed[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int t = a.proposed[k];
    on[k] = t > 0 && !(a.skip_untaken && k > 0 && k != taken);
    const double bc1 = 1.0 - pow((double)a.beta1, (double)(t > 0 ? t : 1));
    const double bc2 = 1.0 - pow((double)a.beta2, (double)(t > 0 ? t : 1));
    step_size[k] = (float)((double)a.lr / bc1);
    bc2s[k] = (float)sqrt(bc2);
  }
  const int64_t nth = (int64_t)gridDim.x * kOptThreads;
  for (int64_t i = (int64_t)blockIdx.x * kOptThreads + tid; i < a.n; i += nth) {
    const int k = i < a.seg1 ? 0 : (i < a.seg2 ? 1 : 2);
    float gi = a.g[i];
    if (clip < 1.f) { gi *= clip; a.g[i] = gi; }
    if (!on[k]) continue;
    const float m = a.beta1 * a.m[i] + (1.f - a.beta1) * gi;        // exp_avg.mul_(b1).add_(g, alpha = 1 - b1)
    const float v = a.beta2 * a.v[i] + (1.f - a.beta2) * gi * gi;   // exp_avg_sq.mul_(b2).addcmul_(g, g, value = 1 - b2)
    a.m[i] = m;
    a.v[i] = v;
    const float denom = sqrtf(v) / bc2s[k] + a.eps;
    a.p[i] = a.p[i] - step_size[k] * (m / denom);
  }
}

int launch_adam_step(const Dims& d, const ParamMap& pm, float* params, float* grads, float* exp_avg, float* exp_avg_sq,
                     int* state, const float* x, float lr, float beta1, float beta2, float eps, float max_norm,
                     int skip_untaken, float* grad_norm, float* ws, cudaStream_t st) {
  int* proposed = (int*)(ws + kOptBlocks);
  adam_norm_kernel<<<kOptBlocks, kOptThreads, 0, st>>>(grads, pm.total, x, state, skip_untaken, ws, proposed);
  AdamArgs a;
  a.p = params; a.g = grads; a.m = exp_avg; a.v = exp_avg_sq;
  a.n = pm.total; a.seg1 = sage_wl(pm, 0, d.L); a.seg2 = sage_wl(pm, 1, d.L);
  a.state = state; a.proposed = proposed; a.partial = ws; a.nparts = kOptBlocks;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.max_norm = max_norm;
  a.skip_untaken = skip_untaken; a.grad_norm = grad_norm;
  adam_apply_kernel<<<kOptBlocks, kOptThreads, 0, st>>>(a);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

int adam_ws_floats() { return kOptBlocks + 4; }

}  // namespace hexgnn
