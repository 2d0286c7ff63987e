// Weight images for the tcgen05 kernels (layout documented in sage_tc.cuh): torch layout -> per layer,
// per 16-wide k chunk, a tf32-hi half and a tf32-lo half in the K-major SWIZZLE_NONE canonical form the
// tensor core reads, so the kernels fetch each ring stage with ONE contiguous bulk copy.
#include <cuda_fp16.h>
#include "sage_tc.cuh"
#include "td_loss.cuh"

namespace hexgnn {

__global__ void pack_tc_kernel(Dims d, ParamMap pm, const float* __restrict__ params, const float* __restrict__ x,
                               TcPackView pv, int with_transpose, int* __restrict__ zero_ints, int n_zero) {
  // hexgnn_train_step / hexgnn_backward: the hand-off flags and tickets of this call start at zero (saves a memset node)
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_zero; i += (int64_t)gridDim.x * blockDim.x) zero_ints[i] = 0;
  const int head = (x[2] == 1.0f) ? 0 : 1;
  const int Hp = d.Hp, H = d.H, N2 = 2 * d.Hp;
  const int NQ = Hp / kTcKC;
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
      // forward: stage (which, p), c = out channel of lin_l (which 0) / lin_r (which 1); fp16 hi | 2^11 lo (sage_tc.cuh)
      const int which = n < Hp ? 0 : 1;
      const int pq = k / kTcKF, kk = k % kTcKF;
      char* stage = (char*)pv.wimg + (((int64_t)(s - 1) * 2 + which) * (Hp / kTcKF) + pq) * (int64_t)(Hp * 64);
      const int64_t off = (int64_t)(kk / 8) * (Hp * 16) + (int64_t)(c / 8) * 128 + (c % 8) * 16 + (kk % 8) * 2;
      const __half h16 = __float2half_rn(val);
      *(__half*)(stage + off) = h16;
      *(__half*)(stage + Hp * 32 + off) = __float2half_rn((val - __half2float(h16)) * 2048.0f);
    }
    if (with_transpose) {
      // backward: dH[i][k] = sum_c dZ_which[i][c] * W_which[c][k]  =>  B[n = k][kk = c], which 0 = lin_r, 1 = lin_l
      const int which = n < Hp ? 1 : 0;
      // stage (which, q) = [fp16 hi | fp16 2^11 lo] of B[n = k][kk = c in 16-chunk q], K-major SWIZZLE_NONE:
      //   byte = (kk/8) * (Hp*16) + (k/8)*128 + (k%8)*16 + (kk%8)*2          (all-fp16 3-product scheme, sage_tc.cuh)
      const int q = c / kTcKC, cq = c % kTcKC;
      char* stage = (char*)pv.wimgT + ((((int64_t)(s - 1) * 2 + which) * NQ + q) * (int64_t)(Hp * 64));
      const int64_t off = (int64_t)(cq / 8) * (Hp * 16) + (int64_t)(k / 8) * 128 + (k % 8) * 16 + (cq % 8) * 2;
      const __half h16 = __float2half_rn(val);
      *(__half*)(stage + off) = h16;
      *(__half*)(stage + Hp * 32 + off) = __float2half_rn((val - __half2float(h16)) * 2048.0f);
    }
  }
  const int nb = d.Lt * Hp;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nb; i += nth) {
    const int s = (int)(i / Hp), c = (int)(i % Hp);
    pv.bias[i] = c < H ? params[sage_bl(pm, head, s, H) + c] : 0.f;
  }
}

int launch_pack_tc(const Dims& d, const ParamMap& pm, const float* params, const float* x, TcPackView pv,
                   bool with_transpose, cudaStream_t st, int* zero_ints, int n_zero) {
  const int64_t total = (int64_t)(d.Lt - 1) * d.Hp * 2 * d.Hp;
  int blocks = (int)((total + 255) / 256);
  if (blocks < 1) blocks = 1;
  if (blocks > 1184) blocks = 1184;
  pack_tc_kernel<<<blocks, 256, 0, st>>>(d, pm, params, x, pv, with_transpose ? 1 : 0, zero_ints, n_zero);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

// Fixed-order reduction of the tcgen05 backward's partial records into one packed record (GradLayout):
//   gw and gb[s>=1]  <- the G group records of sage_dw_tc_kernel   (big_rec [G][gl.g0])
//   gv0              <- sum over graphs (in graph order) of dt_g (x) pooled_g from the head records
//   everything else  <- the per-TILE records of sage_bwd_tc_kernel  (small_rec [ntiles][gl.total - gl.gb])
constexpr int kRedThreads = 1024;   // gv0 blocks: 32 q x 32 graph ranges
__global__ void __launch_bounds__(kRedThreads, 2) reduce_tc_kernel(GradLayout gl, int Hp, int H, int H2,
                                                                const float* __restrict__ small_rec, const int* __restrict__ ntiles,
                                                                const float* __restrict__ big_rec, int G,
                                                                const float* __restrict__ head_rec, int hrec_n, int B,
                                                                int nmain, int ngv0, float* __restrict__ out,
                                                                const TdLossArgs td) {
  // Block order = schedule order: the latency-bound sections go FIRST (TD-loss block, then the per-graph gv0 chains, then
  // the per-tile record chains) so that they run under the waves of the short one-element `main` blocks instead of after
  // them.  `bx` is the block's index in the section layout [main | gv0 | small | td] the code below is written in.
  const int nfront = (int)gridDim.x - nmain;                 // td + gv0 + small blocks
  const int has_td = td.q ? 1 : 0;
  int bx;
  if ((int)blockIdx.x >= nfront) bx = (int)blockIdx.x - nfront;                              // main
  else if ((int)blockIdx.x < has_td) bx = (int)gridDim.x - 1;                                // td
  else bx = nmain + ((int)blockIdx.x - has_td);                                              // gv0, then small
#ifndef HEXGNN_RED_SKIP
#define HEXGNN_RED_SKIP 0      // profiling builds only: bit 0 td block, 1 gv0, 2 per-tile records, 3 group records skipped
#endif
  if (bx == (int)gridDim.x - 1 && td.q) {   // hexgnn_train_step: loss value + TD errors ride along (no extra launch)
    if (!(HEXGNN_RED_SKIP & 1)) td_loss_block(td);
    return;
  }
  if ((HEXGNN_RED_SKIP & 2) && bx >= nmain && bx < nmain + ngv0) return;
  if ((HEXGNN_RED_SKIP & 4) && bx >= nmain + ngv0) return;
  if ((HEXGNN_RED_SKIP & 8) && bx < nmain) return;
  if (bx >= nmain + ngv0) {
    // per-tile records of the chain kernel.  Only three short ranges of the record are live here -- layer-0 bias
    // [gb, gb+Hp), layer-0 weights + head linear [g0, gv0), value-MLP biases and output layer [gv0b, total) -- so the blocks
    // index those ~1 k elements compactly: block = 32 elements x 32 contiguous ranges of tiles (4 interleaved chains inside
    // a range, ranges combined in range order): a fixed summation order that does not depend on the kernels' tile
    // schedule, and two dependent round trips per thread for 209 tiles instead of seven.
    __shared__ float sp_part[32][32];
    const int sb = bx - nmain - ngv0;
    const int ql = threadIdx.x & 31, sl = threadIdx.x >> 5;
    const int64_t n1 = Hp, n2 = gl.gv0 - gl.g0, n3 = gl.total - gl.gv0b;
    const int64_t e = (int64_t)sb * 32 + ql;
    const bool mine = e < n1 + n2 + n3;
    const int64_t i = e < n1 ? gl.gb + e : (e < n1 + n2 ? gl.g0 + (e - n1) : gl.gv0b + (e - n1 - n2));
    const int64_t recn = gl.total - gl.gb;
    const int nsmall = *ntiles;
    const int per = (nsmall + 31) / 32;
    const int pa = sl * per < nsmall ? sl * per : nsmall, pb = (pa + per < nsmall) ? pa + per : nsmall;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    if (mine) {
      const float* sp = small_rec + (i - gl.gb);
      int p = pa;
      for (; p + 4 <= pb; p += 4) {
        a0 += sp[(int64_t)p * recn]; a1 += sp[(int64_t)(p + 1) * recn];
        a2 += sp[(int64_t)(p + 2) * recn]; a3 += sp[(int64_t)(p + 3) * recn];
      }
      for (; p < pb; ++p) a0 += sp[(int64_t)p * recn];
    }
    sp_part[sl][ql] = (a0 + a1) + (a2 + a3);
    __syncthreads();
    if (sl == 0 && mine) {
      float u = sp_part[0][ql];
#pragma unroll
      for (int t = 1; t < 32; ++t) u += sp_part[t][ql];
      out[i] = u;
    }
    return;
  }
  if (bx >= nmain) {
    // gv0[k][q] = sum_g dt_g[k] * pooled_g[q]: block = 32 q x 8 k; its 32 warps take 32 contiguous ranges of the graphs
    // (sequential inside a range, ranges combined in range order: deterministic).  The per-thread chain of dependent-looking
    // L2 round trips is what bounds this kernel (measured by compiling sections out: 17.1 us with, 10.7 us without this
    // section when a thread walked 32 graphs), so the ranges are short -- B/32 graphs -- and there are 4x the blocks.
    __shared__ float part[32][8][32];
    const int H2p = (H2 + 3) & ~3, Q = 4 * H;
    const int nqb = (Q + 31) / 32;
    const int bid = bx - nmain, qb = bid % nqb, kb = bid / nqb;
    const int ql = threadIdx.x & 31, q = qb * 32 + ql, sl = threadIdx.x >> 5, k0 = kb * 8;
    const int Bs = (B + 31) / 32;
    const int ga = sl * Bs < B ? sl * Bs : B, gb = (ga + Bs < B) ? ga + Bs : B;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    if (q < Q) {
      const float* hr = head_rec + (int64_t)ga * hrec_n;
#pragma unroll 4
      for (int g = ga; g < gb; ++g, hr += hrec_n) {
        const float pv = hr[H2p + q];
        const float4 d0 = *(const float4*)(hr + k0), d1 = *(const float4*)(hr + k0 + 4);   // padded with zeros
        acc[0] = fmaf(d0.x, pv, acc[0]); acc[1] = fmaf(d0.y, pv, acc[1]);
        acc[2] = fmaf(d0.z, pv, acc[2]); acc[3] = fmaf(d0.w, pv, acc[3]);
        acc[4] = fmaf(d1.x, pv, acc[4]); acc[5] = fmaf(d1.y, pv, acc[5]);
        acc[6] = fmaf(d1.z, pv, acc[6]); acc[7] = fmaf(d1.w, pv, acc[7]);
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) part[sl][j][ql] = acc[j];
    __syncthreads();
    if (sl < 8 && q < Q) {                   // warp j sums output row k0 + j over the 32 ranges, in range order
      const int j = sl;
      float u = part[0][j][ql];
#pragma unroll
      for (int t = 1; t < 32; ++t) u += part[t][j][ql];
      if (k0 + j < H2) out[gl.gv0 + (int64_t)(k0 + j) * Q + q] = u;
    }
    return;
  }
  const int64_t nth = (int64_t)nmain * blockDim.x;
  for (int64_t i = (int64_t)bx * blockDim.x + threadIdx.x; i < gl.total; i += nth) {
    const bool from_big = i < gl.gb || (i >= gl.gb + Hp && i < gl.g0);
    if (!from_big) continue;                       // gv0 blocks / small-record blocks
    // all G loads requested before the first add (summed in group order: the bits are unchanged)
    constexpr int kBatch = 12;
    float vbuf[kBatch];
#pragma unroll
    for (int g = 0; g < kBatch; ++g) vbuf[g] = g < G ? __ldg(big_rec + (int64_t)g * gl.g0 + i) : 0.f;
    float acc = 0.f;
#pragma unroll
    for (int g = 0; g < kBatch; ++g) if (g < G) acc += vbuf[g];
    for (int g = kBatch; g < G; ++g) acc += big_rec[(int64_t)g * gl.g0 + i];
    out[i] = acc;
  }
}

int launch_reduce_tc(const Dims& d, const float* small_rec, const int* ntiles, const float* big_rec, int G,
                     const float* head_rec, int64_t B, float* packed_sum, cudaStream_t st, const TdLossArgs* td) {
  GradLayout gl;
  grad_layout(d, &gl);
  int blocks = (int)((gl.total + kRedThreads - 1) / kRedThreads);
  if (blocks > 592) blocks = 592;
  const int gv0_blocks = ((4 * d.H + 31) / 32) * ((d.H2 + 7) / 8);
  const int small_blocks = (int)((d.Hp + (gl.gv0 - gl.g0) + (gl.total - gl.gv0b) + 31) / 32);   // the live ranges, compact
  TdLossArgs tdl{};
  if (td) tdl = *td;
  reduce_tc_kernel<<<blocks + gv0_blocks + small_blocks + (td ? 1 : 0), kRedThreads, 0, st>>>(
      gl, d.Hp, d.H, d.H2, small_rec, ntiles, big_rec, G, head_rec, tc_head_rec_floats(d), (int)B, blocks, gv0_blocks,
      packed_sum, tdl);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

}  // namespace hexgnn
