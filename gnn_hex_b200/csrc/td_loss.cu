// Rainbow learner loss on device: the glue between the fused forward and the fused backward
// (GN0/RainbowDQN/Rainbow/common/rainbow.py:301-312):
//     td_est[g] = Q[ptr[g] + action[g]]                     rainbow.py:301,303
//     td_err[g] = td_est[g] - td_tgt[g]                     rainbow.py:307 (PER priorities = |td_err| + 1e-6)
//     loss      = scale * mean_g( w[g] * l(td_tgt[g], td_est[g]) ),  l = MSE or SmoothL1(beta=1)   rainbow.py:85,311-312
// and dloss/dQ, which is zero except at the B selected rows, is written in the same launch, so the learner step
// needs no index / elementwise / reduction kernels between the two fused kernels.  One CTA; the mean is a
// fixed-order tree (deterministic).
#include "common.cuh"
#include "td_loss.cuh"

namespace hexgnn {

__global__ void __launch_bounds__(1024) td_loss_kernel(const float* __restrict__ q, const int64_t* __restrict__ ptr,
                                                       const int64_t* __restrict__ action,
                                                       const float* __restrict__ target, const float* __restrict__ weight,
                                                       int64_t N, int B, int kind, float scale, float* __restrict__ loss,
                                                       float* __restrict__ td_err, float* __restrict__ d_q) {
  td_loss_block(TdLossArgs{q, ptr, action, target, weight, N, B, kind, scale, loss, td_err, d_q});
}

int launch_td_loss(const float* q, const int64_t* ptr, const int64_t* action, const float* target, const float* weight,
                   int64_t N, int64_t B, int kind, float scale, float* loss, float* td_err, float* d_q, cudaStream_t st) {
  td_loss_kernel<<<1, 1024, 0, st>>>(q, ptr, action, target, weight, N, (int)B, kind, scale, loss, td_err, d_q);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

}  // namespace hexgnn

// ---- acting / target glue (rainbow.py:222-225 act, 263-269 td_target) -------------------------------------------
namespace hexgnn {

// One warp per graph: mask the two terminal rows, then (max, first arg max) over the graph's rows.
__global__ void mask_argmax_kernel(float* __restrict__ q, const int64_t* __restrict__ ptr, int64_t N, int B,
                                   int mask_terminals, float mask_value, float* __restrict__ out_max,
                                   int64_t* __restrict__ out_arg) {
  const int lane = threadIdx.x & 31;
  const int g = (int)((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5);
  if (g >= B) return;
  const int64_t r0 = ptr[g], r1 = ptr[g + 1];
  float best = -INFINITY;
  int64_t arg = N;   // empty graph: out-of-range index (torch_scatter's fill for empty groups)
  for (int64_t i = r0 + lane; i < r1; i += 32) {
    float v = q[i];
    if (mask_terminals && i < r0 + 2) { v = mask_value; q[i] = v; }
    if (v > best || (v == best && i < arg)) { best = v; arg = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int64_t oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (ob > best || (ob == best && oa < arg)) { best = ob; arg = oa; }
  }
  if (lane == 0) {
    if (out_max) out_max[g] = r1 > r0 ? best : 0.f;
    if (out_arg) out_arg[g] = arg;
  }
}

__global__ void td_target_kernel(const float* __restrict__ q_target, const int64_t* __restrict__ best,
                                 const float* __restrict__ reward, const float* __restrict__ done, float gamma_n,
                                 int64_t N, int B, float* __restrict__ target) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= B) return;
  const int64_t i = best[g];
  const float nq = (i >= 0 && i < N) ? q_target[i] : __int_as_float(0x7fc00000);
  target[g] = reward[g] + gamma_n * nq * (1.f - done[g]);
}

int launch_mask_argmax(float* q, const int64_t* ptr, int64_t N, int64_t B, int mask_terminals, float mask_value,
                       float* out_max, int64_t* out_arg, cudaStream_t st) {
  const int threads = 256;
  const int64_t blocks = (B * 32 + threads - 1) / threads;
  mask_argmax_kernel<<<(unsigned)blocks, threads, 0, st>>>(q, ptr, N, (int)B, mask_terminals, mask_value, out_max, out_arg);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

int launch_td_target(const float* q_target, const int64_t* best, const float* reward, const float* done, float gamma_n,
                     int64_t N, int64_t B, float* target, cudaStream_t st) {
  td_target_kernel<<<(unsigned)((B + 255) / 256), 256, 0, st>>>(q_target, best, reward, done, gamma_n, N, (int)B, target);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

}  // namespace hexgnn

// ---- data-parallel gradient all-reduce over NVLink peer memory (one-shot, rank-ordered) --------------------------
// Every rank's flat gradient lives in a peer-mapped (symmetric) buffer; each rank reads all W of them with coalesced
// 16-byte loads over NVLink / NVSwitch and adds them in RANK ORDER, so all ranks end with bit-identical sums whatever
// the topology.  1.95 MB per rank: latency-, not bandwidth-bound -- one launch instead of a ring / tree schedule.
namespace hexgnn {

__global__ void __launch_bounds__(256) allreduce_p2p_kernel(const float* const* __restrict__ peers, int world, int64_t n4,
                                                            int64_t n, float* __restrict__ out) {
  const int64_t nth = (int64_t)gridDim.x * blockDim.x;
  const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (int64_t i = t0; i < n4; i += nth) {
    float4 acc = ((const float4*)peers[0])[i];
    for (int r = 1; r < world; ++r) {
      const float4 v = ((const float4*)peers[r])[i];
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    ((float4*)out)[i] = acc;
  }
  for (int64_t i = n4 * 4 + t0; i < n; i += nth) {
    float acc = peers[0][i];
    for (int r = 1; r < world; ++r) acc += peers[r][i];
    out[i] = acc;
  }
}

int launch_allreduce_p2p(const float* const* peers_dev, int world, int64_t n, float* out, cudaStream_t st) {
  const int64_t n4 = n / 4;
  int blocks = (int)((n4 + 255) / 256);
  if (blocks < 1) blocks = 1;
  if (blocks > 592) blocks = 592;
  allreduce_p2p_kernel<<<blocks, 256, 0, st>>>(peers_dev, world, n4, n, out);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

}  // namespace hexgnn
