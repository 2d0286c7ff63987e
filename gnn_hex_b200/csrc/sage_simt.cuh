// Pieces shared by the fused SIMT forward and backward kernels.
#pragma once
#include "common.cuh"

namespace hexgnn {

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Shared-memory carve-up of one tile (R rows of Hp channels).
template <int NJ, int RPW>
struct TileSmem {
  static constexpr int Hp = 16 * NJ;
  static constexpr int W2 = 32 * NJ;
  static constexpr int R = kWarps * RPW;
  static constexpr int KC = 16;                       // k rows per staged weight chunk
  static constexpr int ECAP = kEdgeCapPerRow * R;
  // float regions
  static constexpr int off_a = 0;                     // [R][Hp]   h (fwd) / dZr (bwd)
  static constexpr int off_b = off_a + R * Hp;        // [R][Hp]   Zl (fwd) / dZl (bwd)
  static constexpr int off_w = off_b + R * Hp;        // [2][KC][W2] weight staging
  static constexpr int off_inv = off_w + 2 * KC * W2; // [R] 1/max(deg,1)
  static constexpr int off_x0 = off_inv + R;          // [R] x[:,0]
  static constexpr int off_x1 = off_x0 + R;           // [R] x[:,1]
  static constexpr int off_m0 = off_x1 + R;           // [R] mean-aggregated x0
  static constexpr int off_m1 = off_m0 + R;           // [R]
  static constexpr int off_adv = off_m1 + R;          // [R] advantages / du
  static constexpr int off_pool = off_adv + R;        // [4*Hp] pooled P (sum|max|min|mean) or dP
  static constexpr int off_hid = off_pool + 4 * Hp;   // [Hp] value-MLP hidden / dt
  static constexpr int off_misc = off_hid + Hp;       // [16] scalars
  static constexpr int off_rowptr = off_misc + 16;    // int [R+1] (+pad)
  static constexpr int off_gp = off_rowptr + R + 4;   // int [kMaxTileGraphs+1] (+pad) local graph row starts
  static constexpr int off_col = off_gp + kMaxTileGraphs + 4;  // uint16 [ECAP]
  static constexpr int floats_total = off_col + ECAP / 2;
  static constexpr size_t bytes = (size_t)floats_total * 4;
};

// Load one tile's CSR slice and input features into shared memory.  Returns false (uniformly) when
// the tile does not fit; the caller poisons the outputs.
template <class S>
__device__ __forceinline__ bool load_tile(float* sm, const CsrView& v, const float* __restrict__ x, float flag,
                                          int t, int& g0, int& g1, int& r0, int& n) {
  int* rowptr_s = (int*)(sm + S::off_rowptr);
  int* gp_s = (int*)(sm + S::off_gp);
  unsigned short* col_s = (unsigned short*)(sm + S::off_col);
  g0 = v.tile_g0[t];
  g1 = v.tile_g0[t + 1];
  r0 = v.gptr[g0];
  const int r1 = v.gptr[g1];
  n = r1 - r0;
  const int e0 = v.rowptr[r0];
  const int ne = v.rowptr[r1] - e0;
  if (n > S::R || ne > S::ECAP || n < 0 || ne < 0) return false;
  const int tid = threadIdx.x;
  int bad = 0;
  for (int i = tid; i <= n; i += kThreads) rowptr_s[i] = v.rowptr[r0 + i] - e0;
  for (int g = tid; g <= g1 - g0; g += kThreads) gp_s[g] = v.gptr[g0 + g] - r0;
  for (int e = tid; e < ne; e += kThreads) {
    int c = v.col[e0 + e] - r0;
    if (c < 0 || c >= n) { bad |= HEXGNN_ST_EDGE_CROSSES; c = 0; }
    col_s[e] = (unsigned short)c;
  }
  for (int i = tid; i < n; i += kThreads) {
    const float* xr = x + (int64_t)(r0 + i) * 3;
    sm[S::off_x0 + i] = xr[0];
    sm[S::off_x1 + i] = xr[1];
    if (xr[2] != flag) bad |= HEXGNN_ST_FLAG_MIXED;
  }
  if (bad) atomicOr(v.status, bad);
  __syncthreads();
  for (int i = tid; i < n; i += kThreads) {
    const int b = rowptr_s[i], e = rowptr_s[i + 1];
    const int deg = e - b;
    sm[S::off_inv + i] = 1.0f / (float)(deg > 0 ? deg : 1);
    float s0 = 0.f, s1 = 0.f;
    for (int p = b; p < e; ++p) { const int j = col_s[p]; s0 += sm[S::off_x0 + j]; s1 += sm[S::off_x1 + j]; }
    const float inv = 1.0f / (float)(deg > 0 ? deg : 1);
    sm[S::off_m0 + i] = s0 * inv;
    sm[S::off_m1 + i] = s1 * inv;
  }
  __syncthreads();
  return true;
}

// acc[r][j] += A[rb+r][0:K] * Wg[0:K][lane+32j]  with A in shared memory (row pitch lda floats) and
// Wg [K][WCOLS] in global memory streamed through a double-buffered cp.async stage of KC rows.
// All kThreads threads must call (barriers inside); `active` warps do the math.
template <int RPW, int NCJ, int WCOLS, int KC>
__device__ __forceinline__ void tile_gemm(float (&acc)[RPW][NCJ], const float* __restrict__ A, int lda, int rb,
                                          const float* __restrict__ Wg, int K, float* wstage, bool active) {
  const int tid = threadIdx.x, lane = tid & 31;
  constexpr int chunk_f4 = KC * WCOLS / 4;
  const int nchunks = K / KC;
  auto issue = [&](int kc, int buf) {
    const float4* src = (const float4*)(Wg + (int64_t)kc * KC * WCOLS);
    float4* dst = (float4*)(wstage + buf * KC * WCOLS);
    for (int i = tid; i < chunk_f4; i += kThreads) cp_async16(dst + i, src + i);
    cp_async_commit();
  };
  issue(0, 0);
  for (int kc = 0; kc < nchunks; ++kc) {
    if (kc + 1 < nchunks) { issue(kc + 1, (kc + 1) & 1); cp_async_wait<1>(); }
    else cp_async_wait<0>();
    __syncthreads();
    if (active) {
      const float* wb = wstage + (kc & 1) * KC * WCOLS;
#pragma unroll
      for (int k4 = 0; k4 < KC / 4; ++k4) {
        float4 hv[RPW];
#pragma unroll
        for (int r = 0; r < RPW; ++r) hv[r] = *(const float4*)(A + (rb + r) * lda + kc * KC + k4 * 4);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          float wv[NCJ];
#pragma unroll
          for (int j = 0; j < NCJ; ++j) {
            const int col = lane + 32 * j;
            wv[j] = (col < WCOLS) ? wb[(k4 * 4 + kk) * WCOLS + col] : 0.f;
          }
#pragma unroll
          for (int r = 0; r < RPW; ++r) {
            const float a = kk == 0 ? hv[r].x : kk == 1 ? hv[r].y : kk == 2 ? hv[r].z : hv[r].w;
#pragma unroll
            for (int j = 0; j < NCJ; ++j) acc[r][j] = fmaf(a, wv[j], acc[r][j]);
          }
        }
      }
    }
    __syncthreads();
  }
}

}  // namespace hexgnn
