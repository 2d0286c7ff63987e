// Backward of modern_two_headed on the layer-at-a-time path (lw.cuh): replaces loss.backward() through the model
// (Rainbow/common/rainbow.py:316) for batches the fused kernels cannot take (graphs beyond one tile, norm=True).
//   head        lw_head_bwd_kernel   per graph: dQ -> (dueling combine, tanh, advantage Linear, 4-way pooling, value MLP)
//                                    -> dG [N][H] + one small parameter record per graph          (GN0/models.py:368-384,567-584)
//   layer s     lw_sage_kernel       adjoint direction: [dpre_s | sum_i dpre_i / deg_i] * [W_r ; W_l] -> dH_{s-1}, masked by
//                                    ReLU' of the layer below (norm=False) or followed by the two LayerNorm passes (norm=True)
//   weights     lw_dw_kernel         dWcat_s = [h_{s-1} | 1]^T [dpre_s | dZl_s] for ALL layers in one launch, split-K tcgen05
//                                    GEMM (3xTF32, two accumulators), one record per (layer, group)
//   reduce      lw_reduce_*          records -> the flat torch-layout gradient, fixed order (deterministic, no atomics)
#include "lw_model.cuh"

namespace hexgnn {
using namespace tc;

namespace {

__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ float wsum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double wsum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- head -----------------------------------------------------------------------------------------------------
// per-graph record (floats): glin [Hp] | glinb [4] | dt [H2p] | pooled [4H pad 4] | gv1 [H2p] | gv1b [4]
struct HeadRec { int glin, glinb, dt, pooled, gv1, gv1b, total; };
__host__ __device__ inline HeadRec head_rec(const Dims& d) {
  HeadRec r;
  const int H2p = (d.H2 + 3) & ~3;
  int o = 0;
  r.glin = o; o += d.Hp;
  r.glinb = o; o += 4;
  r.dt = o; o += H2p;
  r.pooled = o; o += (4 * d.H + 3) & ~3;
  r.gv1 = o; o += H2p;
  r.gv1b = o; o += 4;
  r.total = o;
  return r;
}

constexpr int kHbThreads = 256;
struct HeadBwdArgs {
  Dims d; ParamMap pm;
  const float* params; const float* x; const int* gptr;
  const float* hin;          // [N][Hp] final activations g
  int flags;
  const float* d_nodes; const float* d_value;
  float* du;                 // [N] scratch
  float* out; int ldo;       // dG (norm) or dG * (g > 0) (no norm): [N][ldo]
  int apply_mask;
  unsigned* out_absmax;
  float* rec;                // [B][head_rec.total]
};

__global__ void __launch_bounds__(kHbThreads) lw_head_bwd_kernel(const __grid_constant__ HeadBwdArgs a) {
  extern __shared__ float sm[];   // pool [4H] | hid [H2] | dt [H2] | dpool [4H] | amax [H] (int) | amin [H] (int) | misc [8]
  pdl_wait();
  pdl_launch_dependents();
  const int H = a.d.H, Hp = a.d.Hp, H2 = a.d.H2;
  const HeadRec R = head_rec(a.d);
  float* rec = a.rec + (int64_t)blockIdx.x * R.total;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nw = kHbThreads / 32;
  const int r0 = a.gptr[blockIdx.x], n = a.gptr[blockIdx.x + 1] - r0;
  if (n <= 0) {
    for (int i = tid; i < R.total; i += kHbThreads) rec[i] = 0.f;
    return;
  }
  float* pool = sm;
  float* hid = pool + 4 * H;
  float* dts = hid + H2;
  float* dpool = dts + H2;
  int* amax = (int*)(dpool + 4 * H);
  int* amin = amax + H;
  float* misc = (float*)(amin + H);
  const int head = (a.x[2] == 1.0f) ? 0 : 1;
  const float* wa = a.params + a.pm.lin_w[head];
  const float ba = a.params[a.pm.lin_b[head]];
  const bool adv_only = (a.flags & HEXGNN_ADV_ONLY) != 0, sep = (a.flags & HEXGNN_SEPARATE) != 0;
  // (b) sum of the incoming node gradients
  if (wid == 0) {
    float s = 0.f;
    for (int i = lane; i < n; i += 32) s += a.d_nodes[r0 + i];
    s = wsum(s);
    if (lane == 0) misc[0] = s;
  }
  __syncthreads();
  const float dsum = misc[0];
  const float dmean = adv_only ? 0.f : dsum / (float)n;
  const float dv = adv_only ? 0.f : (sep ? (a.d_value ? a.d_value[blockIdx.x] : 0.f) : dsum);
  // (a) du_i = (d_i - mean d) * 2 (1 - tanh^2(u_i))
  for (int i = wid; i < n; i += nw) {
    const float* gr = a.hin + (int64_t)(r0 + i) * Hp;
    float u = 0.f;
    for (int c = lane; c < H; c += 32) u = fmaf(wa[c], gr[c], u);
    u = wsum(u);
    if (lane == 0) {
      const float th = tanhf(u + ba);
      a.du[r0 + i] = (a.d_nodes[r0 + i] - dmean) * 2.0f * (1.0f - th * th);
    }
  }
  // (c) pooling with the first row attaining each extremum (torch_scatter CPU semantics)
  if (!adv_only) {
    for (int c = tid; c < H; c += kHbThreads) {
      float sum = 0.f, mx = -INFINITY, mn = INFINITY;
      int imx = 0, imn = 0;
#pragma unroll 4
      for (int i = 0; i < n; ++i) {
        const float v = a.hin[(int64_t)(r0 + i) * Hp + c];
        sum += v;
        if (v > mx) { mx = v; imx = i; }
        if (v < mn) { mn = v; imn = i; }
      }
      pool[c] = sum; pool[H + c] = mx; pool[2 * H + c] = mn; pool[3 * H + c] = sum / (float)n;
      amax[c] = imx; amin[c] = imn;
    }
  }
  __syncthreads();   // du of this graph is visible to the whole CTA from here on
  float dvp = 0.f;
  if (!adv_only) {
    const float* w0 = a.params + a.pm.v0_w[head];
    const float* b0 = a.params + a.pm.v0_b[head];
    const float* w1 = a.params + a.pm.v1_w[head];
    const float b1 = a.params[a.pm.v1_b[head]];
    for (int k = wid; k < H2; k += nw) {
      float u = 0.f;
      const float* wrow = w0 + (int64_t)k * 4 * H;
      for (int q = lane; q < 4 * H; q += 32) u = fmaf(wrow[q], pool[q], u);
      u = wsum(u);
      if (lane == 0) hid[k] = fmaxf(u + b0[k], 0.f);
    }
    __syncthreads();
    if (wid == 0) {
      float u = 0.f;
      for (int k = lane; k < H2; k += 32) u = fmaf(w1[k], hid[k], u);
      u = wsum(u);
      if (lane == 0) { const float v = tanhf(u + b1); misc[1] = dv * (1.0f - v * v); }
    }
    __syncthreads();
    dvp = misc[1];
    for (int k = tid; k < H2; k += kHbThreads) dts[k] = hid[k] > 0.f ? dvp * w1[k] : 0.f;
    __syncthreads();
    for (int q = tid; q < 4 * H; q += kHbThreads) {
      float u = 0.f;
      for (int k = 0; k < H2; ++k) u = fmaf(w0[(int64_t)k * 4 * H + q], dts[k], u);
      dpool[q] = u;
    }
  } else {
    for (int q = tid; q < 4 * H; q += kHbThreads) dpool[q] = 0.f;
    for (int k = tid; k < H2; k += kHbThreads) { dts[k] = 0.f; hid[k] = 0.f; }
    for (int q = tid; q < 4 * H; q += kHbThreads) pool[q] = 0.f;
    for (int c = tid; c < H; c += kHbThreads) { amax[c] = -1; amin[c] = -1; }
  }
  __syncthreads();
  // (e) the record
  for (int c = tid; c < Hp; c += kHbThreads) {
    float s = 0.f;
    if (c < H) {
#pragma unroll 4
      for (int i = 0; i < n; ++i) s = fmaf(a.du[r0 + i], a.hin[(int64_t)(r0 + i) * Hp + c], s);
    }
    rec[R.glin + c] = s;
  }
  if (wid == 0) {
    float s = 0.f;
    for (int i = lane; i < n; i += 32) s += a.du[r0 + i];
    s = wsum(s);
    if (lane == 0) { rec[R.glinb] = s; rec[R.gv1b] = dvp; }
  }
  for (int k = tid; k < ((H2 + 3) & ~3); k += kHbThreads) {
    rec[R.dt + k] = k < H2 ? dts[k] : 0.f;
    rec[R.gv1 + k] = k < H2 ? dvp * hid[k] : 0.f;
  }
  for (int q = tid; q < ((4 * H + 3) & ~3); q += kHbThreads) rec[R.pooled + q] = q < 4 * H ? pool[q] : 0.f;
  // (h) dG
  float mxabs = 0.f;
  const float invn = 1.0f / (float)n;
  for (int idx = tid; idx < n * Hp; idx += kHbThreads) {
    const int i = idx / Hp, c = idx - i * Hp;
    float v = 0.f;
    if (c < H) {
      v = a.du[r0 + i] * wa[c] + dpool[c] + dpool[3 * H + c] * invn;
      if (i == amax[c]) v += dpool[H + c];
      if (i == amin[c]) v += dpool[2 * H + c];
      if (a.apply_mask && !(a.hin[(int64_t)(r0 + i) * Hp + c] > 0.f)) v = 0.f;
    }
    a.out[(int64_t)(r0 + i) * a.ldo + c] = v;
    mxabs = fmaxf(mxabs, fabsf(v));
  }
  if (a.out_absmax) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mxabs = fmaxf(mxabs, __shfl_xor_sync(0xffffffffu, mxabs, o));
    if (lane == 0 && mxabs > 0.f) atomicMax(a.out_absmax, __float_as_uint(mxabs));
  }
}

// ---- LayerNorm backward (graph mode, statistics over the whole batch) ------------------------------------------------
//   y = gamma * xh + beta, xh = (x - mu) / s, s = sigma + eps;  with dz = dy [* (y_relu > 0)], dxh = dz * gamma:
//   dx = (dxh - S1 / M) / s - xh * S2 / (M sigma),  S1 = sum dxh, S2 = sum dxh * xh;  dgamma_c = sum_i dz xh, dbeta_c = sum_i dz
constexpr int kNbThreads = 256, kNbRows = 64;
struct NormBwdArgs {
  float* d; int ldd;                 // in: dy rows; pass 1 overwrites them with dxh
  const float* xin; int ldx;         // the norm's input x
  const float* mask_src; int ldm;    // ReLU output after the norm (dz = dy where it is > 0), or nullptr
  int64_t N; int H, Hp;
  const double* stats; int nparts;   // partial sums of x (forward)
  const float* gamma[2]; const float* x;
  double* sums;                      // [blocks][2]  partial S1, S2
  float* gb_part;                    // [blocks][2][Hp] partial dgamma, dbeta
  int nblocks;
  float* out; int ldo;               // pass 2: dx
  unsigned* out_absmax;
  float eps;
};
__device__ __forceinline__ void norm_stats(const double* stats, int nparts, int64_t N, int H, float eps, int lane, float* mu,
                                           float* sg, float* s) {
  double t1 = 0.0, t2 = 0.0;
  for (int i = lane; i < nparts; i += 32) { t1 += stats[2 * i]; t2 += stats[2 * i + 1]; }
  t1 = wsum_d(t1); t2 = wsum_d(t2);
  const double M = (double)N * (double)H;
  const double m = t1 / M;
  double var = t2 / M - m * m;
  var = var > 0.0 ? var : 0.0;
  *mu = (float)m;
  *sg = (float)sqrt(var);
  *s = *sg + eps;
}
__global__ void __launch_bounds__(kNbThreads) lw_norm_bwd1_kernel(const __grid_constant__ NormBwdArgs a) {
  __shared__ float st[3];
  __shared__ double red[kNbThreads / 32][2];
  __shared__ float gbs[kNbThreads / 32][2][kLwMaxK];
  pdl_wait();
  pdl_launch_dependents();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (wid == 0) {
    float mu, sg, s;
    norm_stats(a.stats, a.nparts, a.N, a.H, a.eps, lane, &mu, &sg, &s);
    if (lane == 0) { st[0] = mu; st[1] = sg; st[2] = s; }
  }
  __syncthreads();
  const float mu = st[0], s = st[2];
  const int v = (a.x && a.x[2] != 1.0f) ? 1 : 0;
  const float* gamma = a.gamma[v];
  const int64_t row0 = (int64_t)blockIdx.x * kNbRows;
  double S1 = 0.0, S2 = 0.0;
  float dg[4] = {0.f, 0.f, 0.f, 0.f}, db[4] = {0.f, 0.f, 0.f, 0.f};   // channels lane, lane+32, lane+64, lane+96
  // the loads of all eight rows of the warp are issued first (one memory latency per block instead of one per row)
  constexpr int kRW = kNbRows / (kNbThreads / 32);
  float vd[kRW][4], vx[kRW][4], vm[kRW][4];
#pragma unroll
  for (int k = 0; k < kRW; ++k) {
    const int64_t gr = row0 + wid + k * (kNbThreads / 32);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = lane + 32 * j;
      const bool ok = gr < a.N && c < a.H;
      vd[k][j] = ok ? a.d[gr * a.ldd + c] : 0.f;
      vx[k][j] = ok ? a.xin[gr * a.ldx + c] : 0.f;
      vm[k][j] = (ok && a.mask_src) ? a.mask_src[gr * a.ldm + c] : 1.f;
    }
  }
  float gmv[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) gmv[j] = (lane + 32 * j) < a.H ? gamma[lane + 32 * j] : 0.f;
#pragma unroll
  for (int k = 0; k < kRW; ++k) {
    const int64_t gr = row0 + wid + k * (kNbThreads / 32);
    if (gr >= a.N) break;
    float p1 = 0.f, p2 = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = lane + 32 * j;
      if (c >= a.Hp) break;
      float dxh = 0.f;
      if (c < a.H) {
        float dz = vd[k][j];
        if (!(vm[k][j] > 0.f)) dz = 0.f;
        const float xh = (vx[k][j] - mu) / s;
        dg[j] += dz * xh; db[j] += dz;
        dxh = dz * gmv[j];
        p1 += dxh; p2 += dxh * xh;
      }
      a.d[gr * a.ldd + c] = dxh;
    }
    S1 += (double)p1; S2 += (double)p2;
  }
  S1 = wsum_d(S1); S2 = wsum_d(S2);
  if (lane == 0) { red[wid][0] = S1; red[wid][1] = S2; }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c = lane + 32 * j;
    if (c < a.Hp && c < kLwMaxK) { gbs[wid][0][c] = dg[j]; gbs[wid][1][c] = db[j]; }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double t1 = 0.0, t2 = 0.0;
    for (int w = 0; w < kNbThreads / 32; ++w) { t1 += red[w][0]; t2 += red[w][1]; }
    a.sums[2 * (int64_t)blockIdx.x] = t1;
    a.sums[2 * (int64_t)blockIdx.x + 1] = t2;
  }
  for (int i = threadIdx.x; i < 2 * a.Hp; i += kNbThreads) {
    const int which = i / a.Hp, c = i - which * a.Hp;
    float t = 0.f;
    for (int w = 0; w < kNbThreads / 32; ++w) t += gbs[w][which][c];
    a.gb_part[((int64_t)blockIdx.x * 2 + which) * a.Hp + c] = t;
  }
}
__global__ void __launch_bounds__(kNbThreads) lw_norm_bwd2_kernel(const __grid_constant__ NormBwdArgs a) {
  __shared__ float st[5];
  pdl_wait();
  pdl_launch_dependents();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (wid == 0) {
    float mu, sg, s;
    norm_stats(a.stats, a.nparts, a.N, a.H, a.eps, lane, &mu, &sg, &s);
    double t1 = 0.0, t2 = 0.0;
    for (int i = lane; i < a.nblocks; i += 32) { t1 += a.sums[2 * i]; t2 += a.sums[2 * i + 1]; }
    t1 = wsum_d(t1); t2 = wsum_d(t2);
    if (lane == 0) {
      const double M = (double)a.N * (double)a.H;
      st[0] = mu; st[1] = s;
      st[2] = (float)(t1 / M);
      st[3] = (float)(t2 / (M * (double)(sg > 1e-30f ? sg : 1e-30f)));
    }
  }
  __syncthreads();
  const float mu = st[0], s = st[1], m1 = st[2], m2 = st[3];
  const int64_t row0 = (int64_t)blockIdx.x * kNbRows;
  float mx = 0.f;
  constexpr int kRW = kNbRows / (kNbThreads / 32);
  float vd[kRW][4], vx[kRW][4];
#pragma unroll
  for (int k = 0; k < kRW; ++k) {
    const int64_t gr = row0 + wid + k * (kNbThreads / 32);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = lane + 32 * j;
      const bool ok = gr < a.N && c < a.H;
      vd[k][j] = ok ? a.d[gr * a.ldd + c] : 0.f;
      vx[k][j] = ok ? a.xin[gr * a.ldx + c] : 0.f;
    }
  }
#pragma unroll
  for (int k = 0; k < kRW; ++k) {
    const int64_t gr = row0 + wid + k * (kNbThreads / 32);
    if (gr >= a.N) break;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = lane + 32 * j;
      if (c >= a.Hp) break;
      float dx = 0.f;
      if (c < a.H) {
        const float xh = (vx[k][j] - mu) / s;
        dx = (vd[k][j] - m1) / s - xh * m2;
      }
      a.out[gr * a.ldo + c] = dx;
      mx = fmaxf(mx, fabsf(dx));
    }
  }
  if (a.out_absmax) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0 && mx > 0.f) atomicMax(a.out_absmax, __float_as_uint(mx));
  }
}

// ---- weight gradients of layers 1..Lt-1: split-K tcgen05 GEMM -------------------------------------------------------
// Work item = (layer s, group g): the CTA walks the 16-row chunks c = g, g + G, ... of all node rows and accumulates
//   D[k][c'] += sum_rows X[row][k] * Y[row][c'],   X = [h_{s-1} | 1] (M = 128 >= Hp + 1), Y = [dpre_s | dZl_s] (N = 2Hp)
// in TMEM; both operands are row-major fp32 rows in global memory, written by the loader warps into shared memory as MN-major
// SWIZZLE_128B_BASE32B images (the same form sage_dw_tc.cu reads from its stashes): hi = the fp32 value (the tensor core
// ignores the low 13 mantissa bits), lo = rn_tf32(v - trunc_tf32(v)); 3xTF32 with the small products in a second accumulator.
constexpr int kDwLoadWarps = 8;
constexpr int kDwLoadThreads = kDwLoadWarps * 32;
constexpr int kDwThreads2 = kDwLoadThreads + 32;
constexpr int kDwStages2 = 3;
constexpr int kDwChunk = 16;
constexpr int kDwBlkB = kDwChunk * 128;     // bytes of one 32-channel block of a chunk

struct LwDwArgs {
  const float* xtab[kMaxSage];   // use_tab: operand pointers per work-item layer (X = the conv's input rows [N][Hp],
  const float* ytab[kMaxSage];   //          Y = its [dpre | dZl] rows [N][2Hp]); item index = blockIdx.x / G
  int use_tab;
  const float* hs;          // stash base: h[s] = hs + s * N * Hp   (input of layer s = h[s - 1] ...)
  const float* emb2;        // with norm: the input of layer L (after_embed_norm output); else nullptr
  const float* y;           // [Lt][N][2Hp]: Y_s = y + s * N * 2Hp
  int64_t N;
  int Hp, L, Lt, G;
  float* rec;               // [(Lt-1)][G][(Hp + 1)][2Hp]
};

__global__ void __launch_bounds__(kDwThreads2, 1) lw_dw_kernel(const __grid_constant__ LwDwArgs a) {
  extern __shared__ uint8_t smem_dyn[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar_full[kDwStages2], bar_empty[kDwStages2], bar_done;
  __shared__ uint32_t tmem_slot;
  const int Hp = a.Hp, N2 = 2 * Hp, NB = N2 / 32;            // Y blocks of 32 channels
  const int CB = (Hp + 1 + 31) / 32;                          // X blocks (incl. the ones channel)
  const int a_bytes = 4 * kDwBlkB, b_bytes = NB * kDwBlkB, half_bytes = a_bytes + b_bytes, stage_bytes = 2 * half_bytes;
  const uint32_t TM_DS = N2;
  const uint32_t TM_COLS = (2 * N2 <= 64) ? 64 : (2 * N2 <= 128) ? 128 : (2 * N2 <= 256) ? 256 : 512;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int item = blockIdx.x, s = 1 + item / a.G, grp = item % a.G;
  const int64_t nchunks = (a.N + kDwChunk - 1) / kDwChunk;

  if (tid == 0) {
    for (int i = 0; i < kDwStages2; ++i) { mbar_init(&bar_full[i], kDwLoadThreads); mbar_init(&bar_empty[i], 1); }
    mbar_init(&bar_done, 1);
    fence_mbar_init();
  }
  if (wid == kDwLoadWarps) tmem_alloc(&tmem_slot, TM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_slot;
  pdl_wait();
  pdl_launch_dependents();

  const float* X = a.use_tab ? a.xtab[item / a.G] : (a.emb2 && s == a.L) ? a.emb2 : a.hs + (int64_t)(s - 1) * a.N * Hp;
  const float* Y = a.use_tab ? a.ytab[item / a.G] : a.y + (int64_t)s * a.N * N2;

  if (wid == kDwLoadWarps) {
    if (lane == 0) {
      const uint32_t idesc = make_idesc_tf32(128, N2, true, true);
      uint32_t it = 0;
      for (int64_t c = grp; c < nchunks; c += a.G, ++it) {
        const int stg = it % kDwStages2;
        mbar_wait(&bar_full[stg], (it / kDwStages2) & 1);
        tc_fence_after();
        const uint32_t ahi = smem_u32(smem + stg * stage_bytes), bhi = ahi + a_bytes;
        const uint32_t alo = ahi + half_bytes, blo = bhi + half_bytes;
#pragma unroll
        for (int ks = 0; ks < kDwChunk / 8; ++ks) {
          const uint64_t one = (uint64_t)1 << 61;   // SWIZZLE_128B_BASE32B
          const uint64_t dah = make_smem_desc(ahi + ks * 1024, kDwBlkB, 512) | one;
          const uint64_t dal = make_smem_desc(alo + ks * 1024, kDwBlkB, 512) | one;
          const uint64_t dbh = make_smem_desc(bhi + ks * 1024, kDwBlkB, 512) | one;
          const uint64_t dbl = make_smem_desc(blo + ks * 1024, kDwBlkB, 512) | one;
          mma_tf32_ss(tbase + TM_DS, dal, dbh, idesc, (it | ks) != 0);
          mma_tf32_ss(tbase + TM_DS, dah, dbl, idesc, true);
          mma_tf32_ss(tbase, dah, dbh, idesc, (it | ks) != 0);
        }
        mma_commit(&bar_empty[stg]);
      }
      mma_commit(&bar_done);
    }
  } else {
    // ================= loaders: global fp32 rows -> swizzled hi / lo images =================
    // image byte of (channel ch, row r) inside a chunk: (ch/32)*2048 + (r/4)*512 + (r%4)*128 + ((((ch%32)/8) ^ (r%4))*32) + (ch%8)*4
    const int xv = Hp / 4, yv = N2 / 4;                     // float4 per row of X / Y
    const int per_row = xv + yv;
    // A thread moves the same (row, float4) elements of every chunk: the index arithmetic (divisions by a run-time row
    // length, the swizzled image offset) is done once, outside the chunk loop.
    constexpr int kIter = 6;                                // 16 rows x (Hp + 2Hp) / 4 float4 <= 1344 <= 6 x 256 threads
    int64_t goff[kIter];                                    // float offset from the chunk's first row of X / Y; < 0: no element
    uint32_t soff[kIter];                                   // byte offset inside the stage (hi image; lo = + half_bytes)
    int rrow[kIter];
    uint32_t isx_mask = 0;
#pragma unroll
    for (int u = 0; u < kIter; ++u) {
      const int i = tid + u * kDwLoadThreads;
      goff[u] = -1; soff[u] = 0; rrow[u] = 0;
      if (i < kDwChunk * per_row) {
        const int r = i / per_row, q = i - r * per_row;
        const bool isx = q < xv;
        const int ch = 4 * (isx ? q : q - xv);
        goff[u] = isx ? (int64_t)r * Hp + ch : (int64_t)r * N2 + ch;
        soff[u] = (isx ? 0u : (uint32_t)a_bytes) + (uint32_t)(ch >> 5) * kDwBlkB + (uint32_t)(r >> 2) * 512u +
                  (uint32_t)(r & 3) * 128u + ((uint32_t)(((ch & 31) >> 3) ^ (r & 3)) << 5) + (uint32_t)(ch & 7) * 4u;
        rrow[u] = r;
        if (isx) isx_mask |= 1u << u;
      }
    }
    // the rest of the X block that holds channel Hp (ones for live rows, zeros elsewhere; lo image: zeros): 16 x rem <= 512
    const int rem = CB * 32 - Hp;
    uint32_t ooff[2]; int orow[2]; bool oone[2], ohave[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int i = tid + u * kDwLoadThreads;
      ohave[u] = i < kDwChunk * rem;
      const int r = ohave[u] ? i / rem : 0, ch = Hp + (ohave[u] ? i - r * rem : 0);
      ooff[u] = (uint32_t)(ch >> 5) * kDwBlkB + (uint32_t)(r >> 2) * 512u + (uint32_t)(r & 3) * 128u +
                ((uint32_t)(((ch & 31) >> 3) ^ (r & 3)) << 5) + (uint32_t)(ch & 7) * 4u;
      orow[u] = r; oone[u] = ch == Hp;
    }
    uint32_t it = 0;
    for (int64_t c = grp; c < nchunks; c += a.G, ++it) {
      const int stg = it % kDwStages2;
      const int64_t r0 = c * kDwChunk;
      const float* Xc = X + r0 * Hp;
      const float* Yc = Y + r0 * N2;
      // all global loads of the chunk are issued before the ring slot is even free (one L2 / HBM latency per chunk)
      float4 v[kIter];
#pragma unroll
      for (int u = 0; u < kIter; ++u) {
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (goff[u] >= 0 && r0 + rrow[u] < a.N) v[u] = *(const float4*)(((isx_mask >> u) & 1u ? Xc : Yc) + goff[u]);
      }
      mbar_wait(&bar_empty[stg], ((it / kDwStages2) & 1) ^ 1);
      uint8_t* base = smem + stg * stage_bytes;
#pragma unroll
      for (int u = 0; u < kIter; ++u) {
        if (goff[u] < 0) continue;
        uint8_t* dst = base + soff[u];
        float4 l;
        l.x = tf32_rn(v[u].x - tf32_trunc(v[u].x)); l.y = tf32_rn(v[u].y - tf32_trunc(v[u].y));
        l.z = tf32_rn(v[u].z - tf32_trunc(v[u].z)); l.w = tf32_rn(v[u].w - tf32_trunc(v[u].w));
        *(float4*)dst = v[u];
        *(float4*)(dst + half_bytes) = l;
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        if (!ohave[u]) continue;
        *(float*)(base + ooff[u]) = (oone[u] && r0 + orow[u] < a.N) ? 1.0f : 0.f;
        *(float*)(base + half_bytes + ooff[u]) = 0.f;
      }
      fence_proxy_async();
      mbar_arrive(&bar_full[stg]);
    }
    // ================= epilogue =================
    mbar_wait(&bar_done, 0);
    tc_fence_after();
    if (wid < 4) {
      const int m = wid * 32 + lane;
      const uint32_t trow = tbase + ((uint32_t)(wid * 32) << 16);
      float* rec = a.rec + (int64_t)item * (Hp + 1) * N2;
      const bool have = it > 0;
      for (int c0 = 0; c0 < N2; c0 += 8) {
        float v[8], vs[8];
        tmem_ld8(trow + c0, v);
        tmem_ld8(trow + TM_DS + c0, vs);
        tmem_wait_ld();
        if (m <= Hp) {
          float4* dst = (float4*)(rec + (int64_t)m * N2 + c0);
          if (have) {
            dst[0] = make_float4(v[0] + vs[0], v[1] + vs[1], v[2] + vs[2], v[3] + vs[3]);
            dst[1] = make_float4(v[4] + vs[4], v[5] + vs[5], v[6] + vs[6], v[7] + vs[7]);
          } else {
            dst[0] = make_float4(0.f, 0.f, 0.f, 0.f);
            dst[1] = dst[0];
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (wid == kDwLoadWarps) tmem_dealloc(tbase, TM_COLS);
}

// ---- first layer's weight gradients (C_in <= 4) on the CUDA cores -----------------------------------------------------------
//   d lin_l.W[c][k] = sum_i dpre_i[c] mean_i[k],  d lin_l.b[c] = sum_i dpre_i[c],  d lin_r.W[c][k] = sum_i dpre_i[c] x_i[k]
constexpr int kD0Threads = 128, kD0Rows = 64;
struct LwDw0Args {
  const float* x; int ldx, xoff, Cin;
  int64_t N;
  const int* rowptr; const int* col; const float* degf;
  const float* dpre; int ldd;
  int Hp;
  float* part;       // [blocks][2 Cin + 1][Hp]
};
__global__ void __launch_bounds__(kD0Threads) lw_dw0_kernel(const __grid_constant__ LwDw0Args a) {
  __shared__ float feat[kD0Rows][8];   // mean[0..Cin) | self[0..Cin)
  pdl_wait();
  pdl_launch_dependents();
  const int64_t row0 = (int64_t)blockIdx.x * kD0Rows;
  for (int r = threadIdx.x; r < kD0Rows; r += kD0Threads) {
    const int64_t gr = row0 + r;
    float mean[4] = {0.f, 0.f, 0.f, 0.f}, self[4] = {0.f, 0.f, 0.f, 0.f};
    if (gr < a.N) {
      for (int k = 0; k < a.Cin; ++k) self[k] = a.x[gr * a.ldx + a.xoff + k];
      for (int p = a.rowptr[gr]; p < a.rowptr[gr + 1]; ++p) {
        const int64_t j = a.col[p];
        for (int k = 0; k < a.Cin; ++k) mean[k] += a.x[j * a.ldx + a.xoff + k];
      }
      const float dg = a.degf[gr];
      for (int k = 0; k < a.Cin; ++k) mean[k] /= dg;
    }
    for (int k = 0; k < 4; ++k) { feat[r][k] = mean[k]; feat[r][4 + k] = self[k]; }
  }
  __syncthreads();
  const int nrows = (int)((a.N - row0) < kD0Rows ? (a.N - row0) : kD0Rows);
  for (int c = threadIdx.x; c < a.Hp; c += kD0Threads) {
    float gl[4] = {0.f, 0.f, 0.f, 0.f}, gr_[4] = {0.f, 0.f, 0.f, 0.f}, gb = 0.f;
    for (int r = 0; r < nrows; ++r) {
      const float d = a.dpre[(row0 + r) * a.ldd + c];
      gb += d;
#pragma unroll
      for (int k = 0; k < 4; ++k) { gl[k] = fmaf(d, feat[r][k], gl[k]); gr_[k] = fmaf(d, feat[r][4 + k], gr_[k]); }
    }
    float* o = a.part + (int64_t)blockIdx.x * (2 * a.Cin + 1) * a.Hp;
    for (int k = 0; k < a.Cin; ++k) { o[k * a.Hp + c] = gl[k]; o[(a.Cin + 1 + k) * a.Hp + c] = gr_[k]; }
    o[a.Cin * a.Hp + c] = gb;
  }
}

// ---- reductions into the flat torch-layout gradient ---------------------------------------------------------------------
struct LwReduceArgs {
  Dims d; ParamMap pm;
  const float* x;
  float* grad;
  const float* rec; int G;                 // dW records [(Lt-1)][G][(Hp+1)][2Hp]
  const float* part0; int nblk0;           // layer-0 partials [nblk0][5][Hp]
  const float* hrec; int B;                // head records
  float* gnorm; const float* gb_part;      // norm: gradient block, partials [(n_norm)][nbn][2][Hp]
  int nbn, n_norm;
};
__global__ void lw_reduce_sage_kernel(const __grid_constant__ LwReduceArgs a) {
  pdl_wait();
  const int H = a.d.H, Hp = a.d.Hp, N2 = 2 * Hp;
  const int head = (a.x[2] == 1.0f) ? 0 : 1;
  const int s = 1 + blockIdx.y;
  const float* rec = a.rec + (int64_t)(s - 1) * a.G * (Hp + 1) * N2;
  float* wl = a.grad + sage_wl(a.pm, head, s);
  float* bl = a.grad + sage_bl(a.pm, head, s, H);
  float* wr = a.grad + sage_wr(a.pm, head, s, H);
  const int total = (H + 1) * 2 * H;   // (k <= H rows) x (c' in [0, 2H))
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int k = i / (2 * H), cc = i - k * 2 * H;
    const int which = cc / H, c = cc - which * H;           // 0: dpre columns (lin_r, bias), 1: dZl columns (lin_l)
    const int m = k < H ? k : Hp;                            // the ones channel sits at row Hp
    float t = 0.f;
    for (int g = 0; g < a.G; ++g) t += rec[((int64_t)g * (Hp + 1) + m) * N2 + which * Hp + c];
    if (k < H) (which ? wl : wr)[(int64_t)c * H + k] = t;
    else if (!which) bl[c] = t;
  }
}
// value MLP first weight: d W0[k][q] = sum_g dt_g[k] * pooled_g[q], graphs in order (a [H2 x B] x [B x 4H] product, staged
// through shared memory in 128-graph chunks; one block per 32 columns q)
constexpr int kV0Threads = 256, kV0Q = 32, kV0G = 128, kV0K = 64;
__global__ void __launch_bounds__(kV0Threads) lw_reduce_v0w_kernel(const __grid_constant__ LwReduceArgs a) {
  __shared__ float sd[kV0G][kV0K];
  __shared__ float sp[kV0G][kV0Q];
  pdl_wait();
  pdl_launch_dependents();
  const int H = a.d.H, H2 = a.d.H2;
  const int head = (a.x[2] == 1.0f) ? 0 : 1;
  const HeadRec R = head_rec(a.d);
  const int tid = threadIdx.x, ql = tid & 31, kb = tid >> 5;      // outputs of this thread: k = kb + 8u, column q0 + ql
  const int q0 = blockIdx.x * kV0Q;
  float acc[kV0K / 8];
#pragma unroll
  for (int u = 0; u < kV0K / 8; ++u) acc[u] = 0.f;
  for (int g0 = 0; g0 < a.B; g0 += kV0G) {
    __syncthreads();
    for (int i = tid; i < kV0G * kV0K; i += kV0Threads) {
      const int g = i / kV0K, k = i - g * kV0K;
      sd[g][k] = (g0 + g < a.B && k < H2) ? a.hrec[(int64_t)(g0 + g) * R.total + R.dt + k] : 0.f;
    }
    for (int i = tid; i < kV0G * kV0Q; i += kV0Threads) {
      const int g = i / kV0Q, q = q0 + (i - g * kV0Q);
      sp[g][i - g * kV0Q] = (g0 + g < a.B && q < 4 * H) ? a.hrec[(int64_t)(g0 + g) * R.total + R.pooled + q] : 0.f;
    }
    __syncthreads();
    for (int g = 0; g < kV0G; ++g) {
      const float pq = sp[g][ql];
#pragma unroll
      for (int u = 0; u < kV0K / 8; ++u) acc[u] = fmaf(sd[g][kb + 8 * u], pq, acc[u]);
    }
  }
  float* v0w = a.grad + a.pm.v0_w[head];
  if (q0 + ql < 4 * H) {
#pragma unroll
    for (int u = 0; u < kV0K / 8; ++u)
      if (kb + 8 * u < H2) v0w[(int64_t)(kb + 8 * u) * 4 * H + q0 + ql] = acc[u];
  }
}

// sum of n elements p[0], p[stride], ... by one warp: lane l adds elements l, l + 32, ... in order, then a butterfly over
// the lanes -- a fixed order (deterministic), and the loads of a lane are independent (no serial latency chain)
__device__ __forceinline__ float warp_strided_sum(const float* p, int64_t stride, int n, int lane) {
  float t = 0.f;
#pragma unroll 4
  for (int i = lane; i < n; i += 32) t += p[(int64_t)i * stride];
  return wsum(t);
}
__global__ void lw_reduce_small_kernel(const __grid_constant__ LwReduceArgs a) {
  pdl_wait();
  const int H = a.d.H, Hp = a.d.Hp, H2 = a.d.H2;
  const int head = (a.x[2] == 1.0f) ? 0 : 1;
  const HeadRec R = head_rec(a.d);
  const int lane = threadIdx.x & 31;
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;   // one warp per output
  for (int i = gw; i < H + 1 + H2 + H2 + 1; i += nw) {
    int off; float* dst;
    if (i < H) { off = R.glin + i; dst = a.grad + a.pm.lin_w[head] + i; }
    else if (i == H) { off = R.glinb; dst = a.grad + a.pm.lin_b[head]; }
    else if (i < H + 1 + H2) { off = R.dt + (i - H - 1); dst = a.grad + a.pm.v0_b[head] + (i - H - 1); }
    else if (i < H + 1 + 2 * H2) { off = R.gv1 + (i - H - 1 - H2); dst = a.grad + a.pm.v1_w[head] + (i - H - 1 - H2); }
    else { off = R.gv1b; dst = a.grad + a.pm.v1_b[head]; }
    const float t = warp_strided_sum(a.hrec + off, R.total, a.B, lane);
    if (lane == 0) *dst = t;
  }
  // layer 0: partials [nblk0][2 Cin + 1][Hp], Cin = 2: rows 0,1 = lin_l.W[:, k], 2 = bias, 3,4 = lin_r.W[:, k]
  for (int i = gw; i < 5 * H; i += nw) {
    const int row = i / H, c = i - row * H;
    const float t = warp_strided_sum(a.part0 + (int64_t)row * Hp + c, (int64_t)5 * Hp, a.nblk0, lane);
    if (lane != 0) continue;
    if (row < 2) a.grad[sage_wl(a.pm, 0, 0) + (int64_t)c * 2 + row] = t;
    else if (row == 2) a.grad[sage_bl(a.pm, 0, 0, H) + c] = t;
    else a.grad[sage_wr(a.pm, 0, 0, H) + (int64_t)c * 2 + (row - 3)] = t;
  }
  // LayerNorm affine gradients: slot j of gb_part -> norm index (the taken head's variant for head norms)
  if (a.gnorm) {
    const int L = a.d.L, Lh = a.d.Lh;
    for (int i = gw; i < a.n_norm * 2 * H; i += nw) {
      const int j = i / (2 * H), rem = i - j * 2 * H, which = rem / H, c = rem - which * H;
      const float t = warp_strided_sum(a.gb_part + ((int64_t)j * a.nbn * 2 + which) * Hp + c, (int64_t)2 * Hp, a.nbn, lane);
      const int idx = j <= L ? j : j + head * Lh;            // slots: body 0..L-1, after_embed L, head layers L+1..L+Lh
      if (lane == 0) a.gnorm[((int64_t)idx * 2 + which) * H + c] = t;
    }
  }
}

template <typename K, typename A>
int launch_pdl2(K kern, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const A& args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  HEX_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, args));
  return HEXGNN_OK;
}

struct LwBwdWs {
  int64_t degf, zero, imgT, y, tmp, du, rec, part0, hrec, sums, gbpart, total;
  int G, nbn;
};
LwBwdWs lw_bwd_ws(const Dims& d, int64_t N, int64_t B, bool norm) {
  LwBwdWs w;
  int64_t o = 0;
  w.G = bwd_grid_ctas() / (d.Lt - 1 > 0 ? d.Lt - 1 : 1);
  if (w.G < 1) w.G = 1;
  w.nbn = (int)((N + kNbRows - 1) / kNbRows);
  w.degf = o; o += align_up(2 * N * 4, 256);   // max(deg, 1) | its reciprocal
  w.zero = o; o += 256;
  w.imgT = o; o += align_up((int64_t)(d.Lt - 1) * lw_img_bytes(d.Hp, d.Hp), 256);
  w.y = o; o += align_up((int64_t)d.Lt * N * 2 * d.Hp * 4, 256);
  w.tmp = o; o += align_up(N * d.Hp * 4, 256);
  w.du = o; o += align_up(N * 4, 256);
  w.rec = o; o += align_up((int64_t)(d.Lt - 1) * w.G * (d.Hp + 1) * 2 * d.Hp * 4, 256);
  w.part0 = o; o += align_up((int64_t)((N + kD0Rows - 1) / kD0Rows) * 5 * d.Hp * 4, 256);
  w.hrec = o; o += align_up(B * (int64_t)head_rec(d).total * 4, 256);
  w.sums = o; o += align_up((int64_t)w.nbn * 16, 256);
  w.gbpart = o; if (norm) o += align_up((int64_t)(d.Lt + 1) * w.nbn * 2 * d.Hp * 4, 256);
  w.total = o;
  return w;
}

}  // namespace

int64_t lw_bwd_ws_bytes(const Dims& d, int64_t N, int64_t B, bool norm) { return lw_bwd_ws(d, N, B, norm).total; }

// ---- the weight-gradient kernels for other callers (hexara_bwd.cu) ------------------------------------------------------
int lw_dw_groups(int nitems) {
  int g = bwd_grid_ctas() / (nitems > 0 ? nitems : 1);
  return g < 1 ? 1 : g;
}
int64_t lw_dw_rec_floats(int Hp, int nitems) { return (int64_t)nitems * lw_dw_groups(nitems) * (Hp + 1) * 2 * Hp; }
int launch_lw_dw_tab(const float* const* xs, const float* const* ys, int nitems, int64_t N, int Hp, float* rec, cudaStream_t st) {
  if (nitems < 1) return HEXGNN_OK;
  if (nitems > kMaxSage) { set_error("lw_dw: %d layers (max %d)", nitems, kMaxSage); return HEXGNN_E_UNSUPPORTED; }
  LwDwArgs da = {};
  for (int i = 0; i < nitems; ++i) { da.xtab[i] = xs[i]; da.ytab[i] = ys[i]; }
  da.use_tab = 1; da.N = N; da.Hp = Hp; da.L = 0; da.Lt = nitems + 1; da.G = lw_dw_groups(nitems); da.rec = rec;
  const size_t smem = (size_t)kDwStages2 * 2 * (4 * kDwBlkB + (2 * Hp / 32) * kDwBlkB) + 1024;
  HEX_CUDA_CHECK(cudaFuncSetAttribute(lw_dw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  return launch_pdl2(lw_dw_kernel, dim3((unsigned)(nitems * da.G)), dim3(kDwThreads2), smem, st, da);
}
int lw_dw0_blocks(int64_t N) { return (int)((N + kD0Rows - 1) / kD0Rows); }
int launch_lw_dw0(const float* x, int ldx, int Cin, int64_t N, const int* rowptr, const int* col, const float* degf,
                  const float* dpre, int ldd, int Hp, float* part, cudaStream_t st) {
  LwDw0Args a0;
  a0.x = x; a0.ldx = ldx; a0.xoff = 0; a0.Cin = Cin; a0.N = N; a0.rowptr = rowptr; a0.col = col; a0.degf = degf;
  a0.dpre = dpre; a0.ldd = ldd; a0.Hp = Hp; a0.part = part;
  return launch_pdl2(lw_dw0_kernel, dim3((unsigned)lw_dw0_blocks(N)), dim3(kD0Threads), 0, st, a0);
}

int launch_backward_lw(const Dims& d, const ParamMap& pm, const float* params, const float* normp, const float* x, int64_t N,
                       int64_t E, int64_t B, const void* csr, int flags, const float* d_nodes, const float* d_value,
                       const float* stash, float* grad_params, float* grad_norm, float* d_embeds, void* ws, cudaStream_t st) {
  if (d.Hp > kLwMaxK) { set_error("backward_lw: hidden=%d not supported", d.H); return HEXGNN_E_UNSUPPORTED; }
  const CsrView cv = csr_view(const_cast<void*>(csr), N, E, B);
  const bool norm = normp != nullptr;
  const LwBwdWs W = lw_bwd_ws(d, N, B, norm);
  uint8_t* wb = (uint8_t*)ws;
  float* degf = (float*)(wb + W.degf);
  unsigned* absmax = (unsigned*)(wb + W.zero);     // slot s: max |dpre_s|
  uint8_t* imgT = wb + W.imgT;
  float* Y = (float*)(wb + W.y);
  float* tmp = (float*)(wb + W.tmp);
  float* du = (float*)(wb + W.du);
  float* rec = (float*)(wb + W.rec);
  float* part0 = (float*)(wb + W.part0);
  float* hrec = (float*)(wb + W.hrec);
  double* sums = (double*)(wb + W.sums);
  float* gbpart = (float*)(wb + W.gbpart);
  const int H = d.H, Hp = d.Hp, L = d.L, Lt = d.Lt, N2 = 2 * d.Hp;
  const int64_t NH = N * Hp;
  const int64_t parts = lw_stat_parts(N);
  auto hbuf = [&](int s) { return stash + (int64_t)s * NH; };
  auto prebuf = [&](int s) { return stash + (int64_t)(Lt + s) * NH; };
  const float* emb2 = norm ? stash + (int64_t)2 * Lt * NH : nullptr;
  const double* stats = norm ? (const double*)(stash + (int64_t)(2 * Lt + 1) * NH) : nullptr;
  auto stat = [&](int slot) { return stats + (int64_t)slot * parts * 2; };
  auto Ys = [&](int s) { return Y + (int64_t)s * N * N2; };
  auto gamma = [&](int idx, int v) { return normp + ((int64_t)idx + (idx > L ? (int64_t)v * d.Lh : 0)) * 2 * H; };

  HEX_CUDA_CHECK(cudaMemsetAsync(grad_params, 0, (size_t)pm.total * 4, st));
  if (norm) HEX_CUDA_CHECK(cudaMemsetAsync(grad_norm, 0, (size_t)lw_norm_floats(d) * 4, st));
  // ---- transposed weight images (dH = dZ * W) ------------------------------------------------------------
  {
    LwPackArgs pa;
    pa.n = 0; pa.params = params; pa.x = x; pa.img = imgT; pa.bias_out = nullptr;
    for (int s = 1; s < Lt; ++s) {
      if (pa.n == kLwMaxPack) { int rc = launch_lw_pack(pa, st); if (rc) return rc; pa.n = 0; }
      LwPackEntry& e = pa.e[pa.n++];
      for (int v = 0; v < 2; ++v) { e.w_self[v] = sage_wr(pm, v, s, H); e.w_agg[v] = sage_wl(pm, v, s); e.bias[v] = -1; }
      e.ld = H; e.transposed = 1; e.K = H; e.Kp = Hp; e.Nn = H; e.Np = Hp;
      e.img_off = (int64_t)(s - 1) * lw_img_bytes(Hp, Hp);
      e.bias_off = 0;
    }
    int rc = launch_lw_pack(pa, st);
    if (rc) return rc;
  }
  int rc = launch_lw_prepare(cv.rowptr, N, degf, absmax, 64, st);
  if (rc) return rc;

  // LayerNorm backward of one norm: in/out through `tmp` (dy -> dxh) and `out`
  auto norm_bwd = [&](const float* xin, const float* mask, int idx, int slot, int nparts, int gslot, float* out, int ldo,
                      unsigned* amax) {
    NormBwdArgs na = {};
    na.d = tmp; na.ldd = Hp; na.xin = xin; na.ldx = Hp; na.mask_src = mask; na.ldm = Hp;
    na.N = N; na.H = H; na.Hp = Hp; na.stats = stat(slot); na.nparts = nparts;
    for (int v = 0; v < 2; ++v) na.gamma[v] = gamma(idx, v);
    na.x = x; na.sums = sums; na.gb_part = gbpart + (int64_t)gslot * W.nbn * 2 * Hp; na.nblocks = W.nbn;
    na.out = out; na.ldo = ldo; na.out_absmax = amax; na.eps = 1e-5f;
    int r = launch_pdl2(lw_norm_bwd1_kernel, dim3((unsigned)W.nbn), dim3(kNbThreads), 0, st, na);
    if (r) return r;
    return launch_pdl2(lw_norm_bwd2_kernel, dim3((unsigned)W.nbn), dim3(kNbThreads), 0, st, na);
  };
  // gb_part slots: body layer s -> s, after_embed -> L, head layer s >= L -> s + 1   (Lt + 1 slots)

  // ---- head ----------------------------------------------------------------------------------------------------
  {
    HeadBwdArgs ha = {};
    ha.d = d; ha.pm = pm; ha.params = params; ha.x = x; ha.gptr = cv.gptr; ha.hin = hbuf(Lt - 1); ha.flags = flags;
    ha.d_nodes = d_nodes; ha.d_value = d_value; ha.du = du;
    ha.out = norm ? tmp : Ys(Lt - 1); ha.ldo = norm ? Hp : N2;
    ha.apply_mask = norm ? 0 : 1;
    ha.out_absmax = norm ? nullptr : absmax + (Lt - 1);
    ha.rec = hrec;
    const size_t smem = (size_t)(8 * H + 2 * d.H2 + 2 * H + 8) * 4;
    rc = launch_pdl2(lw_head_bwd_kernel, dim3((unsigned)B), dim3(kHbThreads), smem, st, ha);
    if (rc) return rc;
    if (norm) {
      const int s = Lt - 1;
      rc = norm_bwd(prebuf(s), hbuf(s), s < L ? s : s + 1, s, s == 0 ? lw_layer0_blocks(N) : lw_sage_blocks(N),
                    s < L ? s : s + 1, Ys(s), N2, absmax + s);
      if (rc) return rc;
    }
  }
  // ---- layers Lt-1 .. 1: dpre_s -> dpre_{s-1} (and dZl_s for the weight-gradient GEMM) -------------------------------------
  for (int s = Lt - 1; s >= 1; --s) {
    LwSageArgs a = {};
    a.x = Ys(s); a.ldx = N2; a.N = N; a.Kp = Hp; a.Np = Hp; a.Nout = H;
    a.rowptr = cv.rowptr; a.col = cv.col; a.degf = degf; a.rdeg = degf + N; a.adjoint = 1;
    a.wimg = imgT + (int64_t)(s - 1) * lw_img_bytes(Hp, Hp);
    a.in_absmax = absmax + s;
    a.agg_out = Ys(s) + Hp; a.lda = N2;
    a.status = cv.status;
    const bool tap = (s == L);        // the input of layer L is the body embedding (final_conv_acts)
    if (!norm) {
      a.out = Ys(s - 1); a.ldo = N2;
      a.mask_src = hbuf(s - 1); a.ldm = Hp;
      a.out_absmax = absmax + (s - 1);
      if (tap && d_embeds) { a.raw_out = d_embeds; a.ldr = Hp; }
      rc = launch_lw_sage(a, st);
      if (rc) return rc;
    } else {
      a.out = tmp; a.ldo = Hp;
      if (tap && d_embeds) { a.raw_out = d_embeds; a.ldr = Hp; }
      rc = launch_lw_sage(a, st);
      if (rc) return rc;
      if (tap) {   // through after_embed_norm (no ReLU; input h_{L-1}, statistics slot Lt): tmp -> tmp
        rc = norm_bwd(hbuf(L - 1), nullptr, L, Lt, lw_norm_blocks(N), L, tmp, Hp, nullptr);
        if (rc) return rc;
      }
      const int t = s - 1;   // through relu(LN_t(pre_t))
      rc = norm_bwd(prebuf(t), hbuf(t), t < L ? t : t + 1, t, t == 0 ? lw_layer0_blocks(N) : lw_sage_blocks(N),
                    t < L ? t : t + 1, Ys(t), N2, absmax + t);
      if (rc) return rc;
    }
  }
  // ---- weight gradients ---------------------------------------------------------------------------------------------------
  if (Lt > 1) {
    LwDwArgs da = {};
    da.hs = stash; da.emb2 = emb2; da.y = Y; da.N = N; da.Hp = Hp; da.L = L; da.Lt = Lt; da.G = W.G; da.rec = rec;
    const size_t smem = (size_t)kDwStages2 * 2 * (4 * kDwBlkB + (N2 / 32) * kDwBlkB) + 1024;
    static size_t max_set = 0;
    if (smem > max_set) {
      HEX_CUDA_CHECK(cudaFuncSetAttribute(lw_dw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      max_set = smem;
    }
    rc = launch_pdl2(lw_dw_kernel, dim3((unsigned)((Lt - 1) * W.G)), dim3(kDwThreads2), smem, st, da);
    if (rc) return rc;
  }
  {
    LwDw0Args a0;
    a0.x = x; a0.ldx = 3; a0.xoff = 0; a0.Cin = 2; a0.N = N; a0.rowptr = cv.rowptr; a0.col = cv.col; a0.degf = degf;
    a0.dpre = Ys(0); a0.ldd = N2; a0.Hp = Hp; a0.part = part0;
    rc = launch_pdl2(lw_dw0_kernel, dim3((unsigned)((N + kD0Rows - 1) / kD0Rows)), dim3(kD0Threads), 0, st, a0);
    if (rc) return rc;
  }
  LwReduceArgs ra;
  ra.d = d; ra.pm = pm; ra.x = x; ra.grad = grad_params; ra.rec = rec; ra.G = W.G;
  ra.part0 = part0; ra.nblk0 = (int)((N + kD0Rows - 1) / kD0Rows); ra.hrec = hrec; ra.B = (int)B;
  ra.gnorm = norm ? grad_norm : nullptr; ra.gb_part = gbpart; ra.nbn = W.nbn; ra.n_norm = Lt + 1;
  if (Lt > 1) {
    rc = launch_pdl2(lw_reduce_sage_kernel, dim3(32, (unsigned)(Lt - 1)), dim3(256), 0, st, ra);
    if (rc) return rc;
  }
  if (d.H2 > kV0K) { set_error("backward_lw: hidden=%d not supported by the value-MLP reduction", H); return HEXGNN_E_UNSUPPORTED; }
  rc = launch_pdl2(lw_reduce_v0w_kernel, dim3((unsigned)((4 * H + kV0Q - 1) / kV0Q)), dim3(kV0Threads), 0, st, ra);
  if (rc) return rc;
  rc = launch_pdl2(lw_reduce_small_kernel, dim3((unsigned)bwd_grid_ctas()), dim3(256), 0, st, ra);
  return rc;
}

}  // namespace hexgnn
