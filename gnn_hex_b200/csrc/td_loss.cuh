// The TD-loss CTA (rainbow.py:301-312), shared by td_loss_kernel (td_loss.cu) and the extra block of reduce_tc_kernel
// (pack_tc.cu) that hexgnn_train_step uses instead of a separate launch.
#pragma once
#include "common.cuh"

namespace hexgnn {

// One CTA of 1024 threads (also run as an extra block of reduce_tc_kernel inside hexgnn_train_step).
__device__ __forceinline__ void td_loss_block(const TdLossArgs a) {
  const float* __restrict__ q = a.q; const int64_t* __restrict__ ptr = a.ptr; const int64_t* __restrict__ action = a.action;
  const float* __restrict__ target = a.target; const float* __restrict__ weight = a.weight;
  const int64_t N = a.N; const int B = a.B, kind = a.kind; const float scale = a.scale;
  float* __restrict__ loss = a.loss; float* __restrict__ td_err = a.td_err; float* __restrict__ d_q = a.d_q;
  __shared__ float red[1024];
  const int tid = threadIdx.x;
  if (d_q) {
    float4* d4 = (float4*)d_q;
    const int64_t n4 = N >> 2;
    for (int64_t i = tid; i < n4; i += 1024) d4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t i = (n4 << 2) + tid; i < N; i += 1024) d_q[i] = 0.f;
  }
  __syncthreads();   // the zero fill is ordered before the B scattered entries (same CTA)
  float part = 0.f;
  const float invB = scale / (float)B;
  for (int g = tid; g < B; g += 1024) {
    const int64_t idx = ptr[g] + action[g];
    const float w = weight ? weight[g] : 1.f;
    float l, dl, e;
    if (idx < 0 || idx >= N) {
      e = l = dl = __int_as_float(0x7fc00000);   // out-of-range action: poison instead of reading out of bounds
    } else {
      e = q[idx] - target[g];
      if (kind == 0) { l = e * e; dl = 2.f * e; }
      else { const float a = fabsf(e); l = a < 1.f ? 0.5f * e * e : a - 0.5f; dl = a < 1.f ? e : (e > 0.f ? 1.f : -1.f); }
      if (d_q) d_q[idx] += w * dl * invB;        // two graphs never share a row; += keeps even that case exact
    }
    if (td_err) td_err[g] = e;
    part += w * l;
  }
  red[tid] = part;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if (tid < o) red[tid] += red[tid + o];
    __syncthreads();
  }
  if (tid == 0) loss[0] = red[0] * invB;
}


}  // namespace hexgnn
