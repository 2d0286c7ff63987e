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

// ---- the same all-reduce through the NVSwitch (two-shot, in place) --------------------------------------------------------
// The flat gradients sit in a symmetric allocation that is also bound to a MULTICAST address (NVLS).  Rank r owns slice r of
// the buffer: one `multimem.ld_reduce` per 16 bytes makes the switch fetch that granule from all W GPUs and return the fp32
// sum, one `multimem.st` writes the sum back into all W buffers.  Per GPU 2 x n / W floats cross its links (the one-shot
// kernel above pulls W x n).  The block that finishes last then tells every peer "my slice is in your buffer" (release store
// into the peer's flag word) and waits for the same word from every peer, so kernel completion on a rank means its buffer
// holds the complete sum.  Each slice is reduced by one rank and broadcast: all ranks end with identical bits.
__device__ __forceinline__ float4 multimem_ld_reduce_add(const float4* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st(float4* p, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
               :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

__device__ __forceinline__ unsigned long long mc_now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void mc_wait_flags(const int* flags, int world, int want) {
  if ((int)threadIdx.x < world) {
    const int* f = flags + threadIdx.x;
    int v;
    unsigned spins = 0;
    unsigned long long t0 = 0;
    do {
      asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
      if ((++spins & 0xfffu) == 0) {                   // a peer that has not arrived after 60 s never will: fail loudly
        const unsigned long long now = mc_now_ns();    // instead of hanging the box
        if (t0 == 0) t0 = now;
        else if (now - t0 > 60ull * 1000000000ull) __trap();
      }
    } while (v < want);
  }
  __syncthreads();
}

// Flag protocol (one int32 word per (owner rank, writer rank), values only grow): call number e writes 2e - 1 = "my gradient
// of this step is in my buffer" before the reduction and 2e = "my slice of the sums is in your buffer" after it.
__global__ void __launch_bounds__(256) allreduce_mc_kernel(float4* __restrict__ mc, int* const* __restrict__ peer_flags, int rank,
                                                           int world, int64_t n4, int epoch, int* __restrict__ counter) {
  __shared__ int last_s;
  // ---- barrier in: every rank's backward has written its buffer (the writes precede this kernel in stream order) ----
  if (blockIdx.x == 0 && (int)threadIdx.x < world) {
    __threadfence_system();
    asm volatile("st.release.sys.global.s32 [%0], %1;" :: "l"(peer_flags[threadIdx.x] + rank), "r"(2 * epoch - 1) : "memory");
  }
  mc_wait_flags(peer_flags[rank], world, 2 * epoch - 1);
  // ---- this rank's slice: the switch sums the W copies, the sum goes back into all W buffers ----------------------------
  const int64_t per = n4 / world;                      // n4 is a multiple of world (the host pads the buffer)
  const int64_t lo = per * rank, hi = lo + per;
  const int64_t nth = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += 4 * nth) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) if (i + u * nth < hi) v[u] = multimem_ld_reduce_add(mc + i + u * nth);
#pragma unroll
    for (int u = 0; u < 4; ++u) if (i + u * nth < hi) multimem_st(mc + i + u * nth, v[u]);
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) last_s = atomicAdd(counter, 1) == (int)gridDim.x - 1;
  __syncthreads();
  if (!last_s) return;
  // ---- barrier out (last block only): kernel completion == every peer's slice has landed in this rank's buffer ----------
  __threadfence_system();
  if ((int)threadIdx.x < world)
    asm volatile("st.release.sys.global.s32 [%0], %1;" :: "l"(peer_flags[threadIdx.x] + rank), "r"(2 * epoch) : "memory");
  mc_wait_flags(peer_flags[rank], world, 2 * epoch);
  if (threadIdx.x == 0) *counter = 0;                  // ready for the next call on this stream
}

int launch_allreduce_mc(float* mc_buf, int* const* peer_flags_dev, int rank, int world, int64_t n, int epoch, int* counter,
                        cudaStream_t st) {
  const int64_t n4 = n / 4;
  const int64_t per = n4 / world;
  int blocks = (int)((per + 255) / 256);
  if (blocks < 1) blocks = 1;
  if (blocks > bwd_grid_ctas()) blocks = bwd_grid_ctas();   // all blocks spin on the barrier in: they must be co-resident
  allreduce_mc_kernel<<<blocks, 256, 0, st>>>((float4*)mc_buf, peer_flags_dev, rank, world, n4, epoch, counter);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
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
