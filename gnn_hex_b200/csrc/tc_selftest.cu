// Self-test of the tcgen05 building blocks (tc_common.cuh): one CTA computes
// D[128,N] = A[128,K] * B[K,N] in 3xTF32.  Operand staging is selectable so every layout a kernel wants
// to rely on is first proven against fp64 (tests/test_gpu_tc.py):
//   a_mode 0: A in tensor memory   1: A in smem, K-major SWIZZLE_NONE   3: A in smem, MN-major SWIZZLE_128B_BASE32B
//   b_mode 0: B in smem, K-major SWIZZLE_NONE                           3: B in smem, MN-major SWIZZLE_128B_BASE32B
//   (mode 2, MN-major with plain SWIZZLE_128B, is kept only to document that the hardware returns zeros for it)
#include "common.cuh"
#include "tc_common.cuh"

namespace hexgnn {
using namespace tc;

// byte offset of element (mn, k) inside an operand image holding MN x K elements
__device__ __forceinline__ uint32_t img_off(int mode, int mn, int k, int MN, int K) {
  if (mode == 3) {   // MN-major SWIZZLE_128B_BASE32B: 4 k rows x 128 B atoms, 32-byte chunks XOR (k%4)
    const uint32_t blk = (uint32_t)(K / 4) * 512u;
    return (uint32_t)(mn >> 5) * blk + (uint32_t)(k >> 2) * 512u + (uint32_t)(k & 3) * 128u +
           ((((uint32_t)(mn & 31) >> 3) ^ (uint32_t)(k & 3)) << 5) + (uint32_t)(mn & 7) * 4u;
  }
  if (mode == 2) {
    const uint32_t blk = (uint32_t)(K / 8) * 1024u;
    return (uint32_t)(mn >> 5) * blk + (uint32_t)(k >> 3) * 1024u + (uint32_t)(k & 7) * 128u +
           ((((uint32_t)(mn & 31) >> 2) ^ (uint32_t)(k & 7)) << 4) + (uint32_t)(mn & 3) * 4u;
  }
  return (uint32_t)(mn & 7) * 16u + (uint32_t)(mn >> 3) * 128u + (uint32_t)(k >> 2) * ((uint32_t)MN * 16u) + (uint32_t)(k & 3) * 4u;
}
__device__ __forceinline__ uint64_t img_desc(int mode, uint32_t base, int ks, int MN, int K) {
  if (mode == 3)
    return make_smem_desc(base + ks * 1024u, (uint32_t)(K / 4) * 512u, 512u) | ((uint64_t)1 << 61);
  if (mode == 2) return make_smem_desc_sw128(base + ks * 1024u, (uint32_t)(K / 8) * 1024u, 1024u);
  const uint32_t lbo = (uint32_t)MN * 16u;
  return make_smem_desc(base + ks * 2u * lbo, lbo, 128u);
}

__global__ void __launch_bounds__(128, 1) tc_selftest_kernel(const float* __restrict__ A, const float* __restrict__ B,
                                                             float* __restrict__ D, int N, int K, int a_mode, int b_mode) {
  extern __shared__ uint8_t smem_dyn[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  uint8_t* smem_raw = (uint8_t*)(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
  const int Nst = (N + 31) / 32 * 32;                  // storage width of B (whole 32-blocks for the swizzled form)
  const uint32_t b_bytes = (uint32_t)K * Nst * 4, a_bytes = 128u * K * 4;
  uint8_t* bhi = smem_raw;
  uint8_t* blo = bhi + b_bytes;
  uint8_t* ahi = blo + b_bytes;
  uint8_t* alo = ahi + a_bytes;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (warp == 0) tmem_alloc(&tmem_slot, 512);
  if (tid == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  for (int idx = tid; idx < K * N; idx += 128) {
    const int k = idx / N, n = idx - k * N;
    float hi, lo;
    split_tf32(B[idx], hi, lo);
    const uint32_t off = img_off(b_mode, n, k, Nst, K);
    *(float*)(bhi + off) = hi;
    *(float*)(blo + off) = lo;
  }
  if (a_mode != 0) {
    for (int idx = tid; idx < 128 * K; idx += 128) {
      const int m = idx / K, k = idx - m * K;
      float hi, lo;
      split_tf32(A[idx], hi, lo);
      const uint32_t off = img_off(a_mode, m, k, 128, K);
      *(float*)(ahi + off) = hi;
      *(float*)(alo + off) = lo;
    }
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_slot;
  const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
  const uint32_t col_ahi = 0, col_alo = (uint32_t)K, col_d = 256;
  if (a_mode == 0) {
    for (int k0 = 0; k0 < K; k0 += 8) {
      float hi[8], lo[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) split_tf32(A[(size_t)tid * K + k0 + i], hi[i], lo[i]);
      tmem_st8(tbase + lane_base + col_ahi + k0, hi);
      tmem_st8(tbase + lane_base + col_alo + k0, lo);
    }
    tmem_wait_st();
  }
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tc_fence_after();
    const uint32_t idesc = make_idesc_tf32(128, N, b_mode >= 2, a_mode >= 2);
    for (int ks = 0; ks < K / 8; ++ks) {
      const uint64_t dbh = img_desc(b_mode, smem_u32(bhi), ks, Nst, K);
      const uint64_t dbl = img_desc(b_mode, smem_u32(blo), ks, Nst, K);
      if (a_mode == 0) {
        mma_tf32_ts(tbase + col_d, tbase + col_alo + ks * 8, dbh, idesc, ks > 0);
        mma_tf32_ts(tbase + col_d, tbase + col_ahi + ks * 8, dbl, idesc, true);
        mma_tf32_ts(tbase + col_d, tbase + col_ahi + ks * 8, dbh, idesc, true);
      } else {
        const uint64_t dah = img_desc(a_mode, smem_u32(ahi), ks, 128, K);
        const uint64_t dal = img_desc(a_mode, smem_u32(alo), ks, 128, K);
        mma_tf32_ss(tbase + col_d, dal, dbh, idesc, ks > 0);
        mma_tf32_ss(tbase + col_d, dah, dbl, idesc, true);
        mma_tf32_ss(tbase + col_d, dah, dbh, idesc, true);
      }
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

int launch_tc_selftest(const float* A, const float* B, float* D, int N, int K, int modes, cudaStream_t st) {
  const int a_mode = modes >> 4, b_mode = modes & 15;
  const int Nst = (N + 31) / 32 * 32;
  const size_t smem = (size_t)K * Nst * 8 + (a_mode ? (size_t)128 * K * 8 : 0) + 1024;
  if (N % 16 || N < 16 || N > 256 || K % 8 || K < 8 || K > 128 || smem > 220 * 1024 || a_mode < 0 || a_mode > 3 || (b_mode != 0 && b_mode != 2 && b_mode != 3)) {
    set_error("selftest: need N%%16==0<=256, K%%8==0<=128, images <= 220KB, a_mode in {0,1,2}, b_mode in {0,2}");
    return HEXGNN_E_UNSUPPORTED;
  }
  HEX_CUDA_CHECK(cudaFuncSetAttribute(tc_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  tc_selftest_kernel<<<1, 128, smem, st>>>(A, B, D, N, K, a_mode, b_mode);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

}  // namespace hexgnn
