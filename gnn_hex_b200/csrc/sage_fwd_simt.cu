// Fused forward of modern_two_headed, fp32 CUDA-core path (exact-fp32 reference path of the library;
// the tcgen05 path in sage_tc.cu is the fast one).
//
// One persistent CTA per tile (a tile = whole graphs packed up to R rows).  Node features stay in
// shared memory across all L+Lh SAGE layers; per layer
//     Z = h * [W_l^T | W_r^T]                 (node-linear contraction, register tiled)
//     h' = relu( mean_{j in N(i)} Zl_j + Zr_i + b )   (CSR gather from shared memory, no atomics)
// which equals PyG's SAGEConv `lin_l(mean_j h_j) + lin_r(h_i)` by linearity
// (GN0/torch_script_models.py:52-73, GN0/models.py:265-288), then the dueling head
// (GN0/models.py:368-384, 567-584) in the same CTA.
#include "sage_simt.cuh"

namespace hexgnn {

struct FwdArgs {
  Dims d;
  ParamMap pm;
  const float* params;
  const float* x;
  int64_t N, E, B;
  CsrView csr;
  int flags;
  float* out_nodes;
  float* out_value;
  float* stash;
  PackView pv;
};

template <int NJ, int RPW>
__global__ void __launch_bounds__(kThreads, 1) sage_fwd_simt_kernel(const __grid_constant__ FwdArgs a) {
  using S = TileSmem<NJ, RPW>;
  constexpr int Hp = S::Hp, W2 = S::W2;
  extern __shared__ __align__(16) float sm[];
  float* hbuf = sm + S::off_a;
  float* zbuf = sm + S::off_b;
  float* wstage = sm + S::off_w;
  const int* rowptr_s = (const int*)(sm + S::off_rowptr);
  const int* gp_s = (const int*)(sm + S::off_gp);
  const unsigned short* col_s = (const unsigned short*)(sm + S::off_col);
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int H = a.d.H;
  const float flag = a.x[2];
  const int head = (flag == 1.0f) ? 0 : 1;
  const float* P = a.params;
  const int ntiles = *a.csr.ntiles;

  for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
    int g0, g1, r0, n;
    __syncthreads();  // previous tile fully consumed
    if (!load_tile<S>(sm, a.csr, a.x, flag, t, g0, g1, r0, n)) {
      const int r1 = a.csr.gptr[g1];
      const float qnan = __int_as_float(0x7fc00000);
      for (int i = r0 + tid; i < r1; i += kThreads) a.out_nodes[i] = qnan;
      if (a.out_value) for (int g = g0 + tid; g < g1; g += kThreads) a.out_value[g] = qnan;
      if (tid == 0) atomicOr(a.csr.status, HEXGNN_ST_GRAPH_TOO_LARGE);
      continue;
    }

    // ---- layer 0 (2 -> H): aggregate-first, rank-2 update -------------------------------------
    {
      const float* wl = P + sage_wl(a.pm, head, 0);      // [H][2]
      const float* bl = P + sage_bl(a.pm, head, 0, H);   // [H]
      const float* wr = P + sage_wr(a.pm, head, 0, H);   // [H][2]
      for (int idx = tid; idx < n * Hp; idx += kThreads) {
        const int i = idx / Hp, c = idx - i * Hp;
        float v = 0.f;
        if (c < H) {
          v = wl[2 * c] * sm[S::off_m0 + i] + wl[2 * c + 1] * sm[S::off_m1 + i] + bl[c] +
              wr[2 * c] * sm[S::off_x0 + i] + wr[2 * c + 1] * sm[S::off_x1 + i];
          v = fmaxf(v, 0.f);
        }
        hbuf[idx] = v;
        if (a.stash) a.stash[(int64_t)(r0 + i) * Hp + c] = v;
      }
      __syncthreads();
    }

    // ---- layers 1 .. Lt-1 -------------------------------------------------------------------------
    const int rb = wid * RPW;
    const bool active = rb < n;
    for (int s = 1; s < a.d.Lt; ++s) {
      float acc[RPW][NJ];
#pragma unroll
      for (int r = 0; r < RPW; ++r)
#pragma unroll
        for (int j = 0; j < NJ; ++j) acc[r][j] = 0.f;
      tile_gemm<RPW, NJ, W2, S::KC>(acc, hbuf, Hp, rb, a.pv.wcat + (int64_t)(s - 1) * Hp * W2, Hp, wstage, active);
      // Zl -> shared
      if (active) {
#pragma unroll
        for (int r = 0; r < RPW; ++r)
#pragma unroll
          for (int j = 0; j < NJ; ++j) {
            const int col = lane + 32 * j;
            if (col < Hp) zbuf[(rb + r) * Hp + col] = acc[r][j];
          }
      }
      __syncthreads();
      // gather + combine: this thread owns Zr(i, c) for its rows and its columns >= Hp
      if (active) {
        const float* bias = a.pv.bias + s * Hp;
        float* st = a.stash ? a.stash + ((int64_t)s * a.N + r0) * Hp : nullptr;
#pragma unroll
        for (int r = 0; r < RPW; ++r) {
          const int i = rb + r;
          if (i < n) {
            const int eb = rowptr_s[i], ee = rowptr_s[i + 1];
            float g[NJ];
#pragma unroll
            for (int j = 0; j < NJ; ++j) g[j] = 0.f;
            for (int p = eb; p < ee; ++p) {
              const float* zr = zbuf + (int)col_s[p] * Hp;
#pragma unroll
              for (int j = 0; j < NJ; ++j) {
                const int col = lane + 32 * j;
                if (col >= Hp) g[j] += zr[col - Hp];
              }
            }
            const float inv = sm[S::off_inv + i];
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
              const int col = lane + 32 * j;
              if (col >= Hp) {
                const int c = col - Hp;
                const float v = fmaxf(fmaf(g[j], inv, acc[r][j] + bias[c]), 0.f);
                hbuf[i * Hp + c] = v;
                if (st) st[(int64_t)i * Hp + c] = v;
              }
            }
          }
        }
      }
      __syncthreads();
    }

    // ---- advantage head: a_i = 2 tanh(w_a . g_i + b_a)  (models.py:374, 567-571) -----------------
    {
      const float* wa = P + a.pm.lin_w[head];
      const float ba = P[a.pm.lin_b[head]];
      for (int i = wid; i < n; i += kWarps) {
        float u = 0.f;
        for (int c = lane; c < H; c += 32) u = fmaf(wa[c], hbuf[i * Hp + c], u);
        u = warp_sum(u);
        if (lane == 0) sm[S::off_adv + i] = 2.0f * tanhf(u + ba);
      }
      __syncthreads();
    }
    if (a.flags & HEXGNN_ADV_ONLY) {
      for (int i = tid; i < n; i += kThreads) a.out_nodes[r0 + i] = sm[S::off_adv + i];
      continue;
    }

    // ---- value head per graph: pools (sum|max|min|mean) -> MLP -> tanh; dueling combine ------------
    const float* w0 = P + a.pm.v0_w[head];  // [H2][4H]
    const float* b0 = P + a.pm.v0_b[head];
    const float* w1 = P + a.pm.v1_w[head];  // [H2]
    const float b1 = P[a.pm.v1_b[head]];
    const int H2 = a.d.H2;
    float* pool = sm + S::off_pool;
    float* hid = sm + S::off_hid;
    float* misc = sm + S::off_misc;
    for (int g = 0; g < g1 - g0; ++g) {
      const int rs = gp_s[g], re = gp_s[g + 1];
      const int ng = re - rs;
      if (tid < H) {
        float sum = 0.f, mx = 0.f, mn = 0.f;  // empty group -> 0 (torch_scatter semantics)
        if (ng > 0) { mx = -INFINITY; mn = INFINITY; }
        for (int i = rs; i < re; ++i) {
          const float v = hbuf[i * Hp + tid];
          sum += v; mx = fmaxf(mx, v); mn = fminf(mn, v);
        }
        pool[tid] = sum;
        pool[H + tid] = mx;
        pool[2 * H + tid] = mn;
        pool[3 * H + tid] = sum / (float)(ng > 0 ? ng : 1);
      }
      __syncthreads();
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
        if (lane == 0) misc[0] = tanhf(u + b1);
      } else if (wid == 1) {
        float u = 0.f;
        for (int i = rs + lane; i < re; i += 32) u += sm[S::off_adv + i];
        u = warp_sum(u);
        if (lane == 0) misc[1] = u / (float)(ng > 0 ? ng : 1);
      }
      __syncthreads();
      const float val = misc[0], amean = misc[1];
      if (a.flags & HEXGNN_SEPARATE) {
        for (int i = rs + tid; i < re; i += kThreads) a.out_nodes[r0 + i] = sm[S::off_adv + i] - amean;
      } else {
        for (int i = rs + tid; i < re; i += kThreads) a.out_nodes[r0 + i] = val + (sm[S::off_adv + i] - amean);
      }
      if (tid == 0 && a.out_value) a.out_value[g0 + g] = val;
      __syncthreads();
    }
  }
}

template <int NJ, int RPW>
static int launch_fwd_t(const FwdArgs& a, cudaStream_t st) {
  using S = TileSmem<NJ, RPW>;
  auto kern = sage_fwd_simt_kernel<NJ, RPW>;
  HEX_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::bytes));
  int grid = (int)(a.B < 1 ? 1 : (a.B > bwd_grid_ctas() ? bwd_grid_ctas() : a.B));
  {
    ProfScope prof(0, st);
    kern<<<grid, kThreads, S::bytes, st>>>(a);
  }
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

int launch_forward_simt(const Dims& d, const ParamMap& pm, const float* params, const float* x, int64_t N, int64_t E,
                        int64_t B, const void* csr, int tile_rows, int flags, float* out_nodes, float* out_value,
                        float* stash, PackView pv, cudaStream_t st) {
  FwdArgs a;
  a.d = d; a.pm = pm; a.params = params; a.x = x; a.N = N; a.E = E; a.B = B;
  a.csr = csr_view(const_cast<void*>(csr), N, E, B);
  a.flags = flags; a.out_nodes = out_nodes; a.out_value = out_value; a.stash = stash; a.pv = pv;
#define HEX_FWD_CASE(nj)                                              \
  case nj:                                                            \
    if (tile_rows == 128) return launch_fwd_t<nj, 8>(a, st);          \
    break;
  switch (d.NJ) {
    HEX_FWD_CASE(1) HEX_FWD_CASE(2) HEX_FWD_CASE(3) HEX_FWD_CASE(4)
    HEX_FWD_CASE(5) HEX_FWD_CASE(6) HEX_FWD_CASE(7) HEX_FWD_CASE(8)
    default: break;
  }
#undef HEX_FWD_CASE
  set_error("forward(simt): unsupported hidden=%d (padded %d) / tile_rows=%d", d.H, d.Hp, tile_rows);
  return HEXGNN_E_UNSUPPORTED;
}

}  // namespace hexgnn
