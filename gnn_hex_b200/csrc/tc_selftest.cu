// Self-test of the tcgen05 building blocks (tc_common.cuh): one CTA computes
// D[128,N] = A[128,K] * B[K,N] in 3xTF32 with A staged in tensor memory and B in shared memory in
// either canonical SWIZZLE_NONE layout.  tests/test_gpu_tc.py checks it against fp64.
#include "common.cuh"
#include "tc_common.cuh"

namespace hexgnn {
using namespace tc;

__global__ void __launch_bounds__(128, 1) tc_selftest_kernel(const float* __restrict__ A, const float* __restrict__ B,
                                                             float* __restrict__ D, int N, int K, int mn_major) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  uint8_t* bhi = smem_raw;
  uint8_t* blo = smem_raw + (size_t)K * N * 4;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (warp == 0) tmem_alloc(&tmem_slot, 512);
  if (tid == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  const uint32_t lbo = mn_major ? 128u : (uint32_t)N * 16u;
  const uint32_t sbo = mn_major ? (uint32_t)K * 16u : 128u;
  for (int idx = tid; idx < K * N; idx += 128) {
    const int k = idx / N, n = idx - k * N;
    float hi, lo;
    split_tf32(B[idx], hi, lo);
    uint32_t off;
    if (!mn_major) off = (n & 7) * 16 + (n >> 3) * sbo + (k >> 2) * lbo + (k & 3) * 4;
    else off = (n & 3) * 4 + (n >> 2) * sbo + (k & 7) * 16 + (k >> 3) * lbo;
    *(float*)(bhi + off) = hi;
    *(float*)(blo + off) = lo;
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_slot;
  const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
  const uint32_t col_ahi = 0, col_alo = (uint32_t)K, col_d = 256;
  for (int k0 = 0; k0 < K; k0 += 8) {
    float hi[8], lo[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) split_tf32(A[(size_t)tid * K + k0 + i], hi[i], lo[i]);
    tmem_st8(tbase + lane_base + col_ahi + k0, hi);
    tmem_st8(tbase + lane_base + col_alo + k0, lo);
  }
  tmem_wait_st();
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tc_fence_after();
    const uint32_t idesc = make_idesc_tf32(128, N, mn_major != 0);
    const uint32_t kstep_bytes = mn_major ? 128u : 2u * lbo;
    for (int ks = 0; ks < K / 8; ++ks) {
      const uint64_t dhi = make_smem_desc(smem_u32(bhi) + ks * kstep_bytes, lbo, sbo);
      const uint64_t dlo = make_smem_desc(smem_u32(blo) + ks * kstep_bytes, lbo, sbo);
      mma_tf32_ts(tbase + col_d, tbase + col_alo + ks * 8, dhi, idesc, ks > 0);
      mma_tf32_ts(tbase + col_d, tbase + col_ahi + ks * 8, dlo, idesc, true);
      mma_tf32_ts(tbase + col_d, tbase + col_ahi + ks * 8, dhi, idesc, true);
    }
    mma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  for (int n0 = 0; n0 < N; n0 += 8) {
    float v[8];
    tmem_ld8(tbase + lane_base + col_d + n0, v);
    tmem_wait_ld();
#pragma unroll
    for (int i = 0; i < 8; ++i) D[(size_t)tid * N + n0 + i] = v[i];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 512);
}

int launch_tc_selftest(const float* A, const float* B, float* D, int N, int K, int mn_major, cudaStream_t st) {
  const size_t smem = (size_t)K * N * 8;
  if (N % 16 || N < 16 || N > 256 || K % 8 || K < 8 || K > 128 || smem > 200 * 1024) {
    set_error("selftest: need N%%16==0<=256, K%%8==0<=128, 8*K*N<=200KB");
    return HEXGNN_E_UNSUPPORTED;
  }
  HEX_CUDA_CHECK(cudaFuncSetAttribute(tc_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  tc_selftest_kernel<<<1, 128, smem, st>>>(A, B, D, N, K, mn_major);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

}  // namespace hexgnn
