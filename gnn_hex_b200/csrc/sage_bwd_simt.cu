// Fused backward of modern_two_headed, fp32 CUDA-core path.
//
// Reverse mode of sage_fwd_simt.cu, one persistent CTA per tile, activations read from the stash the
// forward wrote.  Per SAGE layer s (output h_{s+1}, input h_s):
//     dpre = dH (.) (h_{s+1} > 0)                     ReLU
//     db  += sum_i dpre_i
//     dZr  = dpre ;  dZl_j = sum_{i in N(j)} dpre_i / deg(i)      (the graph is symmetric, so the
//            transposed CSR is the CSR: the adjoint of the mean-gather is again a gather, no atomics)
//     dW  += h_s^T [dZl | dZr]                        accumulated into this CTA's partial record
//     dH'  = [dZl | dZr] * wcat^T
// Weight gradients: every CTA owns one packed record in the workspace (first tile assigns, later
// tiles add, fixed tile order), pack.cu reduces the records in a fixed order => deterministic.
// max/min pooling routes its gradient to the FIRST row attaining the extremum (torch_scatter CPU
// semantics; GN0/models.py:380-382).
#include "sage_simt.cuh"

namespace hexgnn {

template <int NJ, int RPW>
struct BwdSmem {
  static constexpr int Hp = 16 * NJ;
  static constexpr int W2 = 32 * NJ;
  static constexpr int R = kWarps * RPW;
  static constexpr int KC = 16;
  static constexpr int ECAP = kEdgeCapPerRow * R;
  static constexpr int off_dz = 0;                     // [R][W2]  row i: dZl (0..Hp) | dZr = dpre = dH (Hp..2Hp)
  static constexpr int off_h = off_dz + R * W2;        // [R][Hp]  h_{s+1}, then h_s
  static constexpr int off_w = off_h + R * Hp;         // [2][KC][W2] staging (uses KC*Hp per buffer)
  static constexpr int off_inv = off_w + 2 * KC * W2;
  static constexpr int off_x0 = off_inv + R;
  static constexpr int off_x1 = off_x0 + R;
  static constexpr int off_m0 = off_x1 + R;
  static constexpr int off_m1 = off_m0 + R;
  static constexpr int off_adv = off_m1 + R;           // [R] du_i
  static constexpr int off_th = off_adv + R;           // [R] tanh(u_i)
  static constexpr int off_pool = off_th + R;          // [4*Hp] P, later dP
  static constexpr int off_hid = off_pool + 4 * Hp;    // [Hp] hid / dt
  static constexpr int off_arg = off_hid + Hp;         // int [2*Hp] argmax | argmin rows
  static constexpr int off_misc = off_arg + 2 * Hp;    // [16]
  static constexpr int off_rowptr = off_misc + 16;
  static constexpr int off_gp = off_rowptr + R + 4;
  static constexpr int off_col = off_gp + kMaxTileGraphs + 4;
  static constexpr int floats_total = off_col + ECAP / 2;
  static constexpr size_t bytes = (size_t)floats_total * 4;
};

struct BwdArgs {
  Dims d;
  ParamMap pm;
  GradLayout gl;
  const float* params;
  const float* x;
  int64_t N, E, B;
  CsrView csr;
  int flags;
  const float* d_nodes;
  const float* d_value;
  const float* stash;
  float* d_embeds;
  PackView pv;
  float* partials;  // [gridDim.x][gl.total]
};

template <int NJ, int RPW>
__global__ void __launch_bounds__(kThreads, 1) sage_bwd_simt_kernel(const __grid_constant__ BwdArgs a) {
  using S = BwdSmem<NJ, RPW>;
  constexpr int Hp = S::Hp, W2 = S::W2;
  constexpr int NCJ = (Hp + 31) / 32;
  extern __shared__ __align__(16) float sm[];
  float* dz = sm + S::off_dz;
  float* dzr = dz + Hp;  // pitch W2
  float* hb = sm + S::off_h;
  float* wstage = sm + S::off_w;
  const int* rowptr_s = (const int*)(sm + S::off_rowptr);
  const int* gp_s = (const int*)(sm + S::off_gp);
  const unsigned short* col_s = (const unsigned short*)(sm + S::off_col);
  int* arg_s = (int*)(sm + S::off_arg);
  float* pool = sm + S::off_pool;
  float* hid = sm + S::off_hid;
  float* misc = sm + S::off_misc;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int H = a.d.H, H2 = a.d.H2, Lt = a.d.Lt;
  const float flag = a.x[2];
  const int head = (flag == 1.0f) ? 0 : 1;
  const float* P = a.params;
  const int ntiles = *a.csr.ntiles;
  float* rec = a.partials + (int64_t)blockIdx.x * a.gl.total;
  if ((int)blockIdx.x >= ntiles) return;

  // zero the small accumulators of this CTA's record (gw is assigned by the first tile instead)
  for (int64_t i = a.gl.gb + tid; i < a.gl.total; i += kThreads) rec[i] = 0.f;

  const float* wa = P + a.pm.lin_w[head];
  const float ba = P[a.pm.lin_b[head]];
  const float* w0 = P + a.pm.v0_w[head];
  const float* b0 = P + a.pm.v0_b[head];
  const float* w1 = P + a.pm.v1_w[head];
  const float b1 = P[a.pm.v1_b[head]];

  for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const bool first = (t == (int)blockIdx.x);
    int g0, g1, r0, n;
    __syncthreads();
    if (!load_tile<S>(sm, a.csr, a.x, flag, t, g0, g1, r0, n)) {
      // poison: make the failure visible in the gradients
      if (tid == 0) { rec[a.gl.gb] = __int_as_float(0x7fc00000); atomicOr(a.csr.status, HEXGNN_ST_GRAPH_TOO_LARGE); }
      if (first) for (int64_t i = tid; i < a.gl.gb; i += kThreads) rec[i] = 0.f;
      continue;
    }
    // ---- load g = h_Lt -----------------------------------------------------------------------------
    {
      const float4* src = (const float4*)(a.stash + ((int64_t)(Lt - 1) * a.N + r0) * Hp);
      float4* dst = (float4*)hb;
      for (int i = tid; i < n * Hp / 4; i += kThreads) dst[i] = src[i];
    }
    __syncthreads();
    // ---- advantage: th_i = tanh(u_i) -------------------------------------------------------------
    for (int i = wid; i < n; i += kWarps) {
      float u = 0.f;
      for (int c = lane; c < H; c += 32) u = fmaf(wa[c], hb[i * Hp + c], u);
      u = warp_sum(u);
      if (lane == 0) sm[S::off_th + i] = tanhf(u + ba);
    }
    __syncthreads();
    // ---- per graph: du_i and the pooled-value gradient -> dH rows -------------------------------------
    for (int g = 0; g < g1 - g0; ++g) {
      const int rs = gp_s[g], re = gp_s[g + 1];
      const int ng = re - rs;
      if (ng <= 0) continue;
      if (a.flags & HEXGNN_ADV_ONLY) {
        for (int i = rs + tid; i < re; i += kThreads) {
          const float th = sm[S::off_th + i];
          sm[S::off_adv + i] = a.d_nodes[r0 + i] * 2.0f * (1.0f - th * th);
        }
        __syncthreads();
        for (int idx = tid; idx < ng * Hp; idx += kThreads) {
          const int i = rs + idx / Hp, c = idx % Hp;
          dzr[i * W2 + c] = c < H ? sm[S::off_adv + i] * wa[c] : 0.f;
        }
        __syncthreads();
        continue;
      }
      // sum of upstream node gradients (fixed-order warp tree)
      if (wid == 0) {
        float u = 0.f;
        for (int i = rs + lane; i < re; i += 32) u += a.d_nodes[r0 + i];
        u = warp_sum(u);
        if (lane == 0) misc[0] = u;
      }
      // pools with first-occurrence arg
      if (tid < H) {
        float sum = 0.f, mx = -INFINITY, mn = INFINITY;
        int amx = rs, amn = rs;
        for (int i = rs; i < re; ++i) {
          const float v = hb[i * Hp + tid];
          sum += v;
          if (v > mx) { mx = v; amx = i; }
          if (v < mn) { mn = v; amn = i; }
        }
        pool[tid] = sum; pool[H + tid] = mx; pool[2 * H + tid] = mn; pool[3 * H + tid] = sum / (float)ng;
        arg_s[tid] = amx; arg_s[Hp + tid] = amn;
      }
      __syncthreads();
      const float sumdq = misc[0];
      for (int i = rs + tid; i < re; i += kThreads) {
        const float th = sm[S::off_th + i];
        sm[S::off_adv + i] = (a.d_nodes[r0 + i] - sumdq / (float)ng) * 2.0f * (1.0f - th * th);
      }
      for (int k = wid; k < H2; k += kWarps) {
        float u = 0.f;
        const float* wrow = w0 + (int64_t)k * 4 * H;
        for (int q = lane; q < 4 * H; q += 32) u = fmaf(wrow[q], pool[q], u);
        u = warp_sum(u);
        if (lane == 0) hid[k] = fmaxf(u + b0[k], 0.f);
      }
      __syncthreads();
      if (wid == 0) {
        float u = 0.f;
        for (int k = lane; k < H2; k += 32) u = fmaf(w1[k], hid[k], u);
        u = warp_sum(u);
        if (lane == 0) {
          const float v = tanhf(u + b1);
          const float dv = (a.flags & HEXGNN_SEPARATE) ? (a.d_value ? a.d_value[g0 + g] : 0.f) : sumdq;
          misc[1] = dv * (1.0f - v * v);  // ds
        }
      }
      __syncthreads();
      const float ds = misc[1];
      // value-MLP parameter grads; dt_k overwrites hid_k
      if (tid < H2) {
        const float hk = hid[tid];
        rec[a.gl.gv1 + tid] += ds * hk;
        const float dt = hk > 0.f ? ds * w1[tid] : 0.f;
        rec[a.gl.gv0b + tid] += dt;
        hid[tid] = dt;
      }
      if (tid == 0) rec[a.gl.gv1b] += ds;
      __syncthreads();
      for (int idx = tid; idx < H2 * 4 * H; idx += kThreads) {
        const int k = idx / (4 * H), q = idx - k * 4 * H;
        rec[a.gl.gv0 + idx] += hid[k] * pool[q];
      }
      __syncthreads();  // pool (P) fully consumed before it becomes dP
      float dp = 0.f;
      if (tid < 4 * H) {
        for (int k = 0; k < H2; ++k) dp = fmaf(w0[(int64_t)k * 4 * H + tid], hid[k], dp);
      }
      __syncthreads();
      if (tid < 4 * H) pool[tid] = dp;
      __syncthreads();
      for (int idx = tid; idx < ng * Hp; idx += kThreads) {
        const int i = rs + idx / Hp, c = idx % Hp;
        float v = 0.f;
        if (c < H) {
          v = sm[S::off_adv + i] * wa[c] + pool[c] + pool[3 * H + c] / (float)ng;
          if (arg_s[c] == i) v += pool[H + c];
          if (arg_s[Hp + c] == i) v += pool[2 * H + c];
        }
        dzr[i * W2 + c] = v;
      }
      __syncthreads();
    }
    // ---- head.linear grads --------------------------------------------------------------------------------
    if (tid < H) {
      float u = 0.f;
      for (int i = 0; i < n; ++i) u = fmaf(sm[S::off_adv + i], hb[i * Hp + tid], u);
      rec[a.gl.glin + tid] += u;
    } else if (tid == H) {
      float u = 0.f;
      for (int i = 0; i < n; ++i) u += sm[S::off_adv + i];
      rec[a.gl.glinb] += u;
    }
    __syncthreads();

    // ---- SAGE layers, reverse ------------------------------------------------------------------------------
    const int rb = wid * RPW;
    const bool active = rb < n;
    for (int s = Lt - 1; s >= 0; --s) {
      // hb = h_{s+1};  dzr = dL/dh_{s+1}
      if (s == a.d.L - 1 && a.d_embeds) {
        for (int idx = tid; idx < n * Hp; idx += kThreads) {
          const int i = idx / Hp, c = idx % Hp;
          a.d_embeds[(int64_t)(r0 + i) * Hp + c] = dzr[i * W2 + c];
        }
      }
      for (int idx = tid; idx < n * Hp; idx += kThreads) {
        const int i = idx / Hp, c = idx % Hp;
        if (!(hb[idx] > 0.f)) dzr[i * W2 + c] = 0.f;
      }
      __syncthreads();
      // bias grad
      if (tid < Hp) {
        float u = 0.f;
        for (int i = 0; i < n; ++i) u += dzr[i * W2 + tid];
        rec[a.gl.gb + s * Hp + tid] += u;
      }
      if (s == 0) {
        // layer 0: out = relu(Wl m + b + Wr x)
        if (tid < Hp) {
          float u0 = 0.f, u1 = 0.f, u2 = 0.f, u3 = 0.f;
          for (int i = 0; i < n; ++i) {
            const float dpv = dzr[i * W2 + tid];
            u0 = fmaf(dpv, sm[S::off_m0 + i], u0);
            u1 = fmaf(dpv, sm[S::off_m1 + i], u1);
            u2 = fmaf(dpv, sm[S::off_x0 + i], u2);
            u3 = fmaf(dpv, sm[S::off_x1 + i], u3);
          }
          rec[a.gl.g0 + 0 * Hp + tid] += u0;
          rec[a.gl.g0 + 1 * Hp + tid] += u1;
          rec[a.gl.g0 + 2 * Hp + tid] += u2;
          rec[a.gl.g0 + 3 * Hp + tid] += u3;
        }
        break;
      }
      // dZl (gather of dpre / deg) and load h_s
      for (int idx = tid; idx < n * (Hp / 4); idx += kThreads) {
        const int j = idx / (Hp / 4), c4 = idx % (Hp / 4);
        float4 acc4 = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int p = rowptr_s[j]; p < rowptr_s[j + 1]; ++p) {
          const int i = col_s[p];
          const float inv = sm[S::off_inv + i];
          const float4 v = *(const float4*)(dzr + i * W2 + c4 * 4);
          acc4.x = fmaf(v.x, inv, acc4.x); acc4.y = fmaf(v.y, inv, acc4.y);
          acc4.z = fmaf(v.z, inv, acc4.z); acc4.w = fmaf(v.w, inv, acc4.w);
        }
        *(float4*)(dz + j * W2 + c4 * 4) = acc4;
      }
      {
        const float4* src = (const float4*)(a.stash + ((int64_t)(s - 1) * a.N + r0) * Hp);
        float4* dst = (float4*)hb;
        for (int i = tid; i < n * Hp / 4; i += kThreads) dst[i] = src[i];
      }
      __syncthreads();
      // dW[k][c'] = sum_i h_s[i][k] * dz[i][c'] ; warp -> k rows wid*NJ.., lanes -> c' = lane + 32 j
      {
        float acc[NJ][NJ];
#pragma unroll
        for (int r = 0; r < NJ; ++r)
#pragma unroll
          for (int j = 0; j < NJ; ++j) acc[r][j] = 0.f;
        const int kb = wid * NJ;
        for (int i = 0; i < n; ++i) {
          float av[NJ], bv[NJ];
#pragma unroll
          for (int r = 0; r < NJ; ++r) av[r] = hb[i * Hp + kb + r];
#pragma unroll
          for (int j = 0; j < NJ; ++j) bv[j] = dz[i * W2 + lane + 32 * j];
#pragma unroll
          for (int r = 0; r < NJ; ++r)
#pragma unroll
            for (int j = 0; j < NJ; ++j) acc[r][j] = fmaf(av[r], bv[j], acc[r][j]);
        }
        float* gw = rec + a.gl.gw + (int64_t)(s - 1) * Hp * W2;
#pragma unroll
        for (int r = 0; r < NJ; ++r)
#pragma unroll
          for (int j = 0; j < NJ; ++j) {
            float* p = gw + (int64_t)(kb + r) * W2 + lane + 32 * j;
            *p = first ? acc[r][j] : (*p + acc[r][j]);
          }
      }
      // dH_s = dz * wcatT[s-1]   ([n x 2Hp] x [2Hp x Hp])
      {
        float acc[RPW][NCJ];
#pragma unroll
        for (int r = 0; r < RPW; ++r)
#pragma unroll
          for (int j = 0; j < NCJ; ++j) acc[r][j] = 0.f;
        tile_gemm<RPW, NCJ, Hp, S::KC>(acc, dz, W2, rb, a.pv.wcatT + (int64_t)(s - 1) * W2 * Hp, W2, wstage, active);
        if (active) {
#pragma unroll
          for (int r = 0; r < RPW; ++r)
#pragma unroll
            for (int j = 0; j < NCJ; ++j) {
              const int col = lane + 32 * j;
              if (col < Hp) dzr[(rb + r) * W2 + col] = acc[r][j];
            }
        }
      }
      __syncthreads();
    }
  }
}

template <int NJ, int RPW>
static int launch_bwd_t(const BwdArgs& a, int grid, cudaStream_t st) {
  using S = BwdSmem<NJ, RPW>;
  auto kern = sage_bwd_simt_kernel<NJ, RPW>;
  HEX_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::bytes));
  {
    ProfScope prof(1, st);
    kern<<<grid, kThreads, S::bytes, st>>>(a);
  }
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

int launch_backward_simt(const Dims& d, const ParamMap& pm, const float* params, const float* x, int64_t N, int64_t E,
                         int64_t B, const void* csr, int tile_rows, int flags, const float* d_nodes,
                         const float* d_value, const float* stash, float* d_embeds, PackView pv, float* partials,
                         cudaStream_t st) {
  BwdArgs a;
  a.d = d; a.pm = pm; grad_layout(d, &a.gl);
  a.params = params; a.x = x; a.N = N; a.E = E; a.B = B;
  a.csr = csr_view(const_cast<void*>(csr), N, E, B);
  a.flags = flags; a.d_nodes = d_nodes; a.d_value = d_value; a.stash = stash; a.d_embeds = d_embeds;
  a.pv = pv; a.partials = partials;
  const int grid = bwd_grid_ctas();
#define HEX_BWD_CASE(nj)                                                \
  case nj:                                                              \
    if (tile_rows == 128) return launch_bwd_t<nj, 8>(a, grid, st);      \
    break;
  switch (d.NJ) {
    HEX_BWD_CASE(1) HEX_BWD_CASE(2) HEX_BWD_CASE(3) HEX_BWD_CASE(4)
    HEX_BWD_CASE(5) HEX_BWD_CASE(6) HEX_BWD_CASE(7) HEX_BWD_CASE(8)
    default: break;
  }
#undef HEX_BWD_CASE
  set_error("backward(simt): unsupported hidden=%d (padded %d) / tile_rows=%d", d.H, d.Hp, tile_rows);
  return HEXGNN_E_UNSUPPORTED;
}

}  // namespace hexgnn
