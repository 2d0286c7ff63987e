// Layer-at-a-time tensor-core SAGE layer (see lw.cuh): out = epi([x_i | aggr_i(x)] * [W_self ; W_agg]) for all node rows,
// 128 rows per CTA.  Math: GN0/torch_script_models.py:52-73 (SAGEConv, mean aggregation, root weight without bias),
// layer loop GN0/models.py:261-294.
//
// CTA = 16 worker warps + 1 control warp.
//   control lane 0   bulk-copies the layer's whole weight image into shared memory (2Kp/16 stages, one mbarrier each; issued
//                    BEFORE griddepcontrol.wait: the images do not depend on the previous layer), later issues the MMAs
//   workers, phase G warp per row: self row + CSR gather of the neighbour rows from global memory (L2), fp32, fixed order
//                    -> shared staging tile [128][2Kp + 4]
//   workers, phase A thread = (row, column slice): staging -> fp16 hi / 2^11 lo split -> tcgen05.st into the A operand in TMEM
//   control          D += hi*hi, Ds += lo*hi + hi*lo per 16-k stage (kind::f16, M = 128, N = Np), commit
//   workers, phase E thread = (row, slice): tcgen05.ld D, Ds -> D + 2^-11 Ds -> unscale, bias, mask / ReLU -> global rows,
//                    optional statistics (LayerNorm) and |max| (operand scale of the next backward layer)
#include "lw.cuh"

namespace hexgnn {
using namespace tc;

namespace {

__device__ __forceinline__ void work_bar() { asm volatile("bar.sync 1, %0;" ::"n"(kLwWorkThreads) : "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

struct LwSmem {
  int off_w, off_stage, off_bar, off_red, off_bias, off_rp, off_col, ncol, total;
  int SP;      // staging row pitch (floats)
  int nst;     // weight stages
};
__host__ __device__ inline LwSmem lw_smem(int Kp, int Np) {
  LwSmem s;
  s.nst = 2 * Kp / 16;
  s.SP = 2 * Kp + 4;
  s.off_w = 0;
  s.off_stage = s.nst * Np * 64;
  s.off_bar = s.off_stage + kLwRows * s.SP * 4;
  s.off_red = s.off_bar + 16 * 8 + 16;                 // <= 14 stage barriers + bar_a + bar_mma, tmem slot
  s.off_bias = s.off_red + kLwWorkWarps * 2 * 8;        // bias row (Np floats)
  s.off_rp = s.off_bias + kLwMaxK * 4;                  // CSR row pointers of the block's rows (<= 129)
  s.off_col = s.off_rp + 132 * 4;                       // ... and their column indices (as many as fit, <= 3072)
  const int room = (232448 - s.off_col) / 4;
  s.ncol = room < 3072 ? (room < 0 ? 0 : room) : 3072;
  s.total = s.off_col + s.ncol * 4;
  return s;
}

#ifdef HEXGNN_PHASES
__device__ unsigned long long g_lw_phase[8];   // summed over CTAs: [0] setup, [1] G1, [2] G2, [3] A, [4] wait MMA, [5] E, [6] CTAs
#define LW_STAMP(k) do { if (tid == 0) { const long long t_ = clock64(); atomicAdd(&g_lw_phase[k], (unsigned long long)(t_ - t_prev)); t_prev = t_; } } while (0)
#else
#define LW_STAMP(k) do { } while (0)
#endif

__global__ void __launch_bounds__(kLwThreads, 1) lw_sage_kernel(const __grid_constant__ LwSageArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int Kp = a.Kp, Np = a.Np;
  const LwSmem L = lw_smem(Kp, Np);
  float* stage = (float*)(smem + L.off_stage);
  uint64_t* bar_w = (uint64_t*)(smem + L.off_bar);
  uint64_t* bar_a = bar_w + 14;        // self half of the A operand is in tensor memory
  uint64_t* bar_mma = bar_w + 15;
  uint64_t* bar_a1 = bar_w + 16;       // aggregated half
  uint32_t* tmem_slot = (uint32_t*)(bar_w + 17);
  double* red = (double*)(smem + L.off_red);
  float* s_bias = (float*)(smem + L.off_bias);
  int* s_rp = (int*)(smem + L.off_rp);
  int* s_col = (int*)(smem + L.off_col);
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int SP = L.SP, nst = L.nst;
  const uint32_t TM_A = 0, TM_D = 2 * Kp, TM_DS = 2 * Kp + Np;
  const uint32_t need = 2 * Kp + 2 * Np;
  const uint32_t TM_COLS = need <= 32 ? 32 : need <= 64 ? 64 : need <= 128 ? 128 : need <= 256 ? 256 : 512;
  const int64_t row0 = (int64_t)blockIdx.x * a.rows;
#ifdef HEXGNN_PHASES
  long long t_prev = clock64();
#endif

  if (tid == 0) {
    for (int i = 0; i < nst; ++i) mbar_init(&bar_w[i], 1);
    mbar_init(bar_a, kLwWorkWarps);
    mbar_init(bar_a1, kLwWorkWarps);
    mbar_init(bar_mma, 1);
    fence_mbar_init();
  }
  if (wid == kLwWorkWarps) tmem_alloc(tmem_slot, TM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = *tmem_slot;
  const int stage_bytes = Np * 64, half_bytes = Np * 32;

  if (wid == kLwWorkWarps) {
    // ================= control warp =================
    if (lane == 0) {
      for (int q = 0; q < nst; ++q) {
        mbar_expect_tx(&bar_w[q], (uint32_t)stage_bytes);
        bulk_g2s(smem + L.off_w + q * stage_bytes, a.wimg + (size_t)q * stage_bytes, (uint32_t)stage_bytes, &bar_w[q]);
      }
      pdl_wait();                 // (the weight images were packed at least two launches ago: see launch order in lw_model.cu)
      pdl_launch_dependents();
      const uint32_t idesc = make_idesc_f16(kLwRows, Np);
      const uint32_t lbo = (uint32_t)Np * 16u, sbo = 128u;
      mbar_wait(bar_a, 0);        // the k-steps of the self half run under the workers' gather
      tc_fence_after();
      for (int q = 0; q < nst; ++q) {
        if (q == nst / 2) { mbar_wait(bar_a1, 0); tc_fence_after(); }
        mbar_wait(&bar_w[q], 0);
        tc_fence_after();
        const uint32_t b0 = smem_u32(smem + L.off_w + q * stage_bytes);
        const uint64_t bhi = make_smem_desc(b0, lbo, sbo), blo = make_smem_desc(b0 + half_bytes, lbo, sbo);
        const uint32_t alo = tbase + TM_A + q * 16, ahi = alo + 8;
        mma_f16_ts(tbase + TM_DS, alo, bhi, idesc, q != 0);
        mma_f16_ts(tbase + TM_DS, ahi, blo, idesc, true);
        mma_f16_ts(tbase + TM_D, ahi, bhi, idesc, q != 0);
      }
      mma_commit(bar_mma);
    }
  } else {
    // ================= workers =================
    pdl_wait();
    LW_STAMP(0);
    // operand scale 2^(6 - floor(log2 max|x|)) for gradient inputs (exact; keeps both fp16 halves inside fp16's range)
    float ascale = 1.0f, unscale = 1.0f;
    if (a.in_absmax) {
      int e = (int)((*a.in_absmax >> 23) & 0xffu);
      if (e != 0) {
        e = e < 10 ? 10 : (e > 250 ? 250 : e);
        ascale = __uint_as_float((uint32_t)(260 - e) << 23);
        unscale = __uint_as_float((uint32_t)(e - 6) << 23);
      }
    }
    // ---- phase G1: this block's own rows -> staging (coalesced; they are the self half of the operand AND the source of
    //      every in-block neighbour below: graphs are contiguous row ranges, so almost all neighbours live in the block) ----
    const int NV = Kp / 4;
    const int R = a.rows;                                       // live rows of this block (<= 128, balanced over the SMs)
    const int64_t rlo = row0, rhi = (row0 + R < a.N) ? row0 + R : a.N;
    const int nrows = (int)(rhi - rlo);
    // the block's slice of the CSR goes to shared memory too (coalesced, one L2 latency for the whole block instead of two
    // dependent ones per row)
    if (tid <= nrows) s_rp[tid] = a.rowptr[rlo + tid];
    if (tid < Np) s_bias[tid] = a.bias ? a.bias[tid] : 0.f;
    {
      // all eight rows of the warp are requested before the first is stored (one L2 latency, not eight in turn)
      constexpr int kRW = kLwRows / kLwWorkWarps;
      float4 selfv[kRW];
      float rdv[kRW];
#pragma unroll
      for (int k = 0; k < kRW; ++k) {
        const int64_t gr = row0 + wid + k * kLwWorkWarps;
        selfv[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        rdv[k] = 0.f;
        if (gr < rhi && lane < NV) selfv[k] = *(const float4*)(a.x + gr * a.ldx + 4 * lane);
        if (gr < rhi && lane == 31) rdv[k] = a.rdeg[gr];
      }
#pragma unroll
      for (int k = 0; k < kRW; ++k) {
        const int r = wid + k * kLwWorkWarps;
        if (lane < NV) *(float4*)(stage + r * SP + 4 * lane) = selfv[k];
        if (lane == 31) stage[r * SP + 2 * Kp] = rdv[k];                      // pad slot: 1 / deg of the row
      }
    }
    work_bar();
    LW_STAMP(1);
    // the column indices are requested now and parked in registers: their latency hides behind the split of the self half
    const int e0 = s_rp[0];
    const int ne = s_rp[nrows] - e0, nc = ne < L.ncol ? ne : L.ncol;
    int colreg[6];
#pragma unroll
    for (int u = 0; u < 6; ++u) {
      const int i = tid + u * kLwWorkThreads;
      colreg[u] = i < nc ? a.col[e0 + i] : 0;
    }
    // ---- phase A (self half): staging -> fp16 split -> tensor memory; its MMAs start while the gather below runs ----------
    const int q4 = wid & 3, slice = wid >> 2;
    const int row = q4 * 32 + lane;
    const uint32_t trow = tbase + ((uint32_t)(q4 * 32) << 16);
    int range_bad = 0;
    // a TMEM lane quarter without live rows (balanced blocks hold fewer than 128 rows) is left as it is: accumulator rows
    // are independent, and nothing of a dead row is read back
    const bool quarter_live = q4 * 32 < nrows;
    auto split_groups = [&](int g_begin, int g_end) {
      for (int g = g_begin + slice; quarter_live && g < g_end; g += 4) {
        const float4 u0 = *(const float4*)(stage + row * SP + 8 * g), u1 = *(const float4*)(stage + row * SP + 8 * g + 4);
        const float v[8] = {u0.x * ascale, u0.y * ascale, u0.z * ascale, u0.w * ascale,
                            u1.x * ascale, u1.y * ascale, u1.z * ascale, u1.w * ascale};
        uint32_t hi4[4], lo4[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          split_f16x2(v[2 * j], v[2 * j + 1], hi4[j], lo4[j]);
          if (fabsf(v[2 * j]) >= 60000.f || fabsf(v[2 * j + 1]) >= 60000.f) range_bad = 1;
        }
        const uint32_t c = trow + TM_A + (g >> 1) * 16 + (g & 1) * 4;
        tmem_st4u(c, lo4);
        tmem_st4u(c + 8, hi4);
      }
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
    };
    split_groups(0, Kp / 8);
    if (lane == 0) mbar_arrive(bar_a);
#pragma unroll
    for (int u = 0; u < 6; ++u) {
      const int i = tid + u * kLwWorkThreads;
      if (i < nc) s_col[i] = colreg[u];
    }
    work_bar();
    // ---- phase G2: aggregated half.  Neighbour rows come from the staged tile (shared memory) when they belong to this
    //      block, from global memory (L2) otherwise; fixed summation order (CSR order) either way ---------------------
    for (int r = wid; r < kLwRows; r += kLwWorkWarps) {
      const int64_t gr = row0 + r;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      if (gr < rhi && lane < NV) {
        const float* xb = a.x + 4 * lane;
        const float* sb = stage + 4 * lane;
        const int b = s_rp[r] - e0, e = s_rp[r + 1] - e0;
        // neighbours in CSR order; a neighbour inside the block comes from shared memory, one outside (a graph that
        // straddles two blocks) from global memory -- both as predicated loads, so the four visits of an unrolled step are
        // in flight together whatever their mix
        if (!a.adjoint) {
#pragma unroll 4
          for (int p = b; p < e; ++p) {
            const int64_t j = p < L.ncol ? s_col[p] : a.col[e0 + p];
            const bool in = j >= rlo && j < rhi;
            float4 v = *(const float4*)(sb + (in ? (int)(j - rlo) : 0) * SP);
            if (!in) v = *(const float4*)(xb + j * a.ldx);
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
          }
          // x * fl(1 / deg): within one ulp of the reference's division (a division per element costs ~45 instructions here)
          const float rd = stage[r * SP + 2 * Kp];
          acc.x *= rd; acc.y *= rd; acc.z *= rd; acc.w *= rd;
        } else {
#pragma unroll 4
          for (int p = b; p < e; ++p) {
            const int64_t j = p < L.ncol ? s_col[p] : a.col[e0 + p];
            const bool in = j >= rlo && j < rhi;
            const int jr = in ? (int)(j - rlo) : 0;
            float4 v = *(const float4*)(sb + jr * SP);
            float rj = stage[jr * SP + 2 * Kp];
            if (!in) { v = *(const float4*)(xb + j * a.ldx); rj = a.rdeg[j]; }
            acc.x = fmaf(v.x, rj, acc.x); acc.y = fmaf(v.y, rj, acc.y); acc.z = fmaf(v.z, rj, acc.z); acc.w = fmaf(v.w, rj, acc.w);
          }
        }
        if (a.agg_out) *(float4*)(a.agg_out + gr * a.lda + 4 * lane) = acc;
      }
      if (lane < NV) *(float4*)(stage + r * SP + Kp + 4 * lane) = acc;
    }
    work_bar();
    LW_STAMP(2);
    // ---- phase A (aggregated half) ----------------------------------------------------------------------------
    split_groups(Kp / 8, 2 * Kp / 8);
    if (lane == 0) mbar_arrive(bar_a1);
    if (range_bad && a.status) atomicOr(a.status, HEXGNN_ST_FP16_RANGE);
    // ---- phase E -------------------------------------------------------------------------------------
    LW_STAMP(3);
    mbar_wait(bar_mma, 0);
    tc_fence_after();
    LW_STAMP(4);
    const int64_t gr = row0 + row;
    const bool live = row < a.rows && gr < a.N;
    float s1 = 0.f, s2 = 0.f, mx = 0.f;
    const float poison = (range_bad && a.status) ? __int_as_float(0x7fc00000) : 0.f;
    for (int g = slice; quarter_live && g < Np / 8; g += 4) {
      float v[8], vs[8];
      tmem_ld8(trow + TM_D + 8 * g, v);
      tmem_ld8(trow + TM_DS + 8 * g, vs);
      tmem_wait_ld();
      if (!live) continue;
      const int c0 = 8 * g;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        float z = (v[u] + vs[u] * (1.0f / 2048.0f)) * unscale;
        z += s_bias[c0 + u];
        if (a.relu) z = fmaxf(z, 0.f);
        v[u] = z + poison;
      }
      if (a.raw_out) {
        float* o = a.raw_out + gr * a.ldr + c0;
        *(float4*)o = make_float4(v[0], v[1], v[2], v[3]);
        *(float4*)(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
      }
      if (a.mask_src) {
        const float4 m0 = *(const float4*)(a.mask_src + gr * a.ldm + c0), m1 = *(const float4*)(a.mask_src + gr * a.ldm + c0 + 4);
        const float m[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = m[u] > 0.f ? v[u] : 0.f;
      }
      if (a.accum) {
        const float4 o0 = *(const float4*)(a.out + gr * a.ldo + c0), o1 = *(const float4*)(a.out + gr * a.ldo + c0 + 4);
        v[0] += o0.x; v[1] += o0.y; v[2] += o0.z; v[3] += o0.w; v[4] += o1.x; v[5] += o1.y; v[6] += o1.z; v[7] += o1.w;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        if (c0 + u >= a.Nout) v[u] = 0.f;
        s1 += v[u]; s2 += v[u] * v[u];
        mx = fmaxf(mx, fabsf(v[u]));
      }
      float* o = a.out + gr * a.ldo + c0;
      *(float4*)o = make_float4(v[0], v[1], v[2], v[3]);
      *(float4*)(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
    if (a.out_absmax) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      if (lane == 0 && mx > 0.f) atomicMax(a.out_absmax, __float_as_uint(mx));
    }
    if (a.stats) {
      const double d1 = warp_sum_d((double)s1), d2 = warp_sum_d((double)s2);
      if (lane == 0) { red[2 * wid] = d1; red[2 * wid + 1] = d2; }
      work_bar();
      if (tid == 0) {
        double t1 = 0.0, t2 = 0.0;
        for (int w = 0; w < kLwWorkWarps; ++w) { t1 += red[2 * w]; t2 += red[2 * w + 1]; }
        a.stats[2 * (int64_t)blockIdx.x] = t1;
        a.stats[2 * (int64_t)blockIdx.x + 1] = t2;
      }
    }
  }
#ifdef HEXGNN_PHASES
  if (wid < kLwWorkWarps) { LW_STAMP(5); if (tid == 0) atomicAdd(&g_lw_phase[6], 1ull); }
#endif
  tc_fence_before();
  __syncthreads();
  if (wid == kLwWorkWarps) tmem_dealloc(tbase, TM_COLS);
}

// ---- weight images --------------------------------------------------------------------------------------------
__global__ void lw_pack_kernel(const __grid_constant__ LwPackArgs a) {
  const LwPackEntry& e = a.e[blockIdx.y];
  const int v = (a.x && a.x[2] != 1.0f) ? 1 : 0;
  const int Kp = e.Kp, Np = e.Np;
  const int npairs = 2 * Kp * Np / 2;
  const float* ws = a.params + e.w_self[v];
  const float* wa = a.params + e.w_agg[v];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < npairs; i += gridDim.x * blockDim.x) {
    // pair index -> (stage q, kk pair, c): consecutive threads = consecutive c (coalesced 16-byte granules do not matter here)
    const int c = i % Np, kp = (i / Np) % 8, q = i / (Np * 8);
    float val[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int kg = 16 * q + 2 * kp + u;
      const bool self = kg < Kp;
      const int k = self ? kg : kg - Kp;
      float w = 0.f;
      if (c < e.Nn && k < e.K) {
        const float* src = self ? ws : wa;
        w = e.transposed ? src[(int64_t)k * e.ld + c] : src[(int64_t)c * e.ld + k];
      }
      val[u] = w;
    }
    uint32_t hi2, lo2;
    split_f16x2(val[0], val[1], hi2, lo2);
    const int kk = 2 * kp;
    const int64_t byte = (int64_t)q * Np * 64 + (kk / 8) * (Np * 16) + (c / 8) * 128 + (c % 8) * 16 + (kk % 8) * 2;
    *(uint32_t*)(a.img + e.img_off + byte) = hi2;
    *(uint32_t*)(a.img + e.img_off + byte + Np * 32) = lo2;
  }
  if (blockIdx.x == 0 && e.bias[v] >= 0) {
    for (int c = threadIdx.x; c < Np; c += blockDim.x) a.bias_out[e.bias_off + c] = c < e.Nn ? a.params[e.bias[v] + c] : 0.f;
  }
}

__global__ void lw_prepare_kernel(const int* __restrict__ rowptr, int64_t N, float* __restrict__ degf, unsigned* zero,
                                  int zero_words) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < N) {
    const int d = rowptr[i + 1] - rowptr[i];
    degf[i] = (float)(d > 0 ? d : 1);
    degf[N + i] = 1.0f / (float)(d > 0 ? d : 1);
  }
  if (i < zero_words) zero[i] = 0u;
}

// ---- first layer ------------------------------------------------------------------------------------------------
constexpr int kL0Threads = 256, kL0Rows = 64;
__global__ void __launch_bounds__(kL0Threads) lw_layer0_kernel(const __grid_constant__ LwLayer0Args a) {
  __shared__ double red[kL0Threads / 32][2];
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int64_t row0 = (int64_t)blockIdx.x * kL0Rows;
  // phase 1, one thread per row: self features and the neighbour mean (independent latency chains, not 8 rows per warp in turn)
  __shared__ float feat[kL0Rows][8];
  if (threadIdx.x < kL0Rows) {
    const int64_t gr = row0 + threadIdx.x;
    float self[4] = {0.f, 0.f, 0.f, 0.f}, mean[4] = {0.f, 0.f, 0.f, 0.f};
    if (gr < a.N) {
      const int b = a.rowptr[gr], e = a.rowptr[gr + 1];
      for (int k = 0; k < a.Cin; ++k) self[k] = a.x[gr * a.ldx + a.xoff + k];
      for (int p = b; p < e; ++p) {
        const int64_t j = a.col[p];
        for (int k = 0; k < a.Cin; ++k) mean[k] += a.x[j * a.ldx + a.xoff + k];
      }
      const float dg = a.degf[gr];
      for (int k = 0; k < a.Cin; ++k) mean[k] /= dg;
    }
    for (int k = 0; k < 4; ++k) { feat[threadIdx.x][k] = mean[k]; feat[threadIdx.x][4 + k] = self[k]; }
  }
  __syncthreads();
  float s1 = 0.f, s2 = 0.f;
  for (int r = wid; r < kL0Rows; r += kL0Threads / 32) {
    const int64_t gr = row0 + r;
    if (gr >= a.N) break;
    for (int c = lane; c < a.Hp; c += 32) {
      float v = 0.f;
      if (c < a.H) {
        v = a.bl[c];
        for (int k = 0; k < a.Cin; ++k) v = fmaf(a.wl[c * a.Cin + k], feat[r][k], v);
        for (int k = 0; k < a.Cin; ++k) v = fmaf(a.wr[c * a.Cin + k], feat[r][4 + k], v);
        if (a.relu) v = fmaxf(v, 0.f);
        s1 += v; s2 += v * v;
      }
      a.out[gr * a.ldo + c] = v;
    }
  }
  if (a.stats) {
    const double d1 = warp_sum_d((double)s1), d2 = warp_sum_d((double)s2);
    if (lane == 0) { red[wid][0] = d1; red[wid][1] = d2; }
    __syncthreads();
    if (threadIdx.x == 0) {
      double t1 = 0.0, t2 = 0.0;
      for (int w = 0; w < kL0Threads / 32; ++w) { t1 += red[w][0]; t2 += red[w][1]; }
      a.stats[2 * (int64_t)blockIdx.x] = t1;
      a.stats[2 * (int64_t)blockIdx.x + 1] = t2;
    }
  }
}

// ---- graph-mode LayerNorm (+ ReLU), element-wise ----------------------------------------------------------------------
constexpr int kNormThreads = 256, kNormRows = 64;
__global__ void __launch_bounds__(kNormThreads) lw_norm_kernel(const __grid_constant__ LwNormArgs a) {
  __shared__ double red[kNormThreads / 32][2];
  __shared__ float mu_s, sd_s;
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (wid == 0) {   // statistics from the producer's partial sums, fixed order
    double t1 = 0.0, t2 = 0.0;
    for (int i = lane; i < a.nparts; i += 32) { t1 += a.stats_in[2 * i]; t2 += a.stats_in[2 * i + 1]; }
    t1 = warp_sum_d(t1); t2 = warp_sum_d(t2);
    if (lane == 0) {
      const double M = (double)a.N * (double)a.H;
      const double mu = t1 / M;
      double var = t2 / M - mu * mu;
      var = var > 0.0 ? var : 0.0;
      mu_s = (float)mu;
      sd_s = (float)sqrt(var) + a.eps;
    }
  }
  __syncthreads();
  const float mu = mu_s, sd = sd_s;
  const int v = (a.x && a.x[2] != 1.0f) ? 1 : 0;
  const float* gamma = a.gamma[v];
  const float* beta = a.beta[v];
  const int64_t row0 = (int64_t)blockIdx.x * kNormRows;
  float s1 = 0.f, s2 = 0.f;
  // all loads of the warp's eight rows are issued before the first value is used (one memory latency per block, not eight)
  constexpr int kRW = kNormRows / (kNormThreads / 32), kCL = kLwMaxK / 32 + 1;
  float vin[kRW][kCL];
#pragma unroll
  for (int k = 0; k < kRW; ++k) {
    const int64_t gr = row0 + wid + k * (kNormThreads / 32);
#pragma unroll
    for (int j = 0; j < kCL; ++j) {
      const int c = lane + 32 * j;
      vin[k][j] = (gr < a.N && c < a.H) ? a.in[gr * a.ldi + c] : 0.f;
    }
  }
  float gm[kCL], bt[kCL];
#pragma unroll
  for (int j = 0; j < kCL; ++j) {
    const int c = lane + 32 * j;
    gm[j] = c < a.H ? gamma[c] : 0.f;
    bt[j] = c < a.H ? beta[c] : 0.f;
  }
#pragma unroll
  for (int k = 0; k < kRW; ++k) {
    const int64_t gr = row0 + wid + k * (kNormThreads / 32);
    if (gr >= a.N) break;
#pragma unroll
    for (int j = 0; j < kCL; ++j) {
      const int c = lane + 32 * j;
      if (c >= a.Hp) break;
      float y = 0.f;
      if (c < a.H) {
        y = ((vin[k][j] - mu) / sd) * gm[j] + bt[j];
        if (a.relu) y = fmaxf(y, 0.f);
        s1 += y; s2 += y * y;
      }
      a.out[gr * a.ldo + c] = y;
    }
  }
  if (a.stats_out) {
    const double d1 = warp_sum_d((double)s1), d2 = warp_sum_d((double)s2);
    if (lane == 0) { red[wid][0] = d1; red[wid][1] = d2; }
    __syncthreads();
    if (threadIdx.x == 0) {
      double t1 = 0.0, t2 = 0.0;
      for (int w = 0; w < kNormThreads / 32; ++w) { t1 += red[w][0]; t2 += red[w][1]; }
      a.stats_out[2 * (int64_t)blockIdx.x] = t1;
      a.stats_out[2 * (int64_t)blockIdx.x + 1] = t2;
    }
  }
}

template <typename K, typename A>
int launch_pdl(K kern, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const A& args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  HEX_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, args));
  return HEXGNN_OK;
}

}  // namespace

// Rows per block: 128, or -- when that needs more blocks than SMs -- the even split of the rows over k full waves of the SMs
// (>= 64 rows: lw_stat_parts(N) slots always suffice for the per-block statistics).
int lw_sage_rows(int64_t N) {
  const int sms = bwd_grid_ctas();
  const int64_t blocks = (N + kLwRows - 1) / kLwRows;
  if (blocks <= sms) return kLwRows;
  const int64_t k = (blocks + sms - 1) / sms;
  int64_t r = (N + sms * k - 1) / (sms * k);
  return (int)(r < 64 ? 64 : (r > kLwRows ? kLwRows : r));
}
int lw_sage_blocks(int64_t N) {
  const int r = lw_sage_rows(N);
  return (int)((N + r - 1) / r);
}

int launch_lw_sage(const LwSageArgs& a_in, cudaStream_t st) {
  LwSageArgs a = a_in;
  a.rows = lw_sage_rows(a.N);
  if (a.Kp < 16 || a.Kp > kLwMaxK || a.Kp % 16 || a.Np < 16 || a.Np > kLwMaxK || a.Np % 16 || a.ldx % 4 || a.ldo % 4 ||
      (a.mask_src && a.ldm % 4) || (a.agg_out && a.lda % 4)) {
    set_error("lw_sage: unsupported shape Kp=%d Np=%d ldx=%d ldo=%d", a.Kp, a.Np, a.ldx, a.ldo);
    return HEXGNN_E_UNSUPPORTED;
  }
  if (a.N <= 0) return HEXGNN_OK;
  const LwSmem L = lw_smem(a.Kp, a.Np);
  static int max_set = 0;   // idempotent attribute
  if (L.total > max_set) {
    HEX_CUDA_CHECK(cudaFuncSetAttribute(lw_sage_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
    max_set = 232448;
  }
  return launch_pdl(lw_sage_kernel, dim3((unsigned)lw_sage_blocks(a.N)), dim3(kLwThreads), (size_t)L.total, st, a);
}

int launch_lw_pack(const LwPackArgs& a, cudaStream_t st) {
  if (a.n <= 0) return HEXGNN_OK;
  if (a.n > kLwMaxPack) { set_error("lw_pack: too many entries"); return HEXGNN_E_BADARG; }
  lw_pack_kernel<<<dim3(16, (unsigned)a.n), 256, 0, st>>>(a);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

int launch_lw_prepare(const int* rowptr, int64_t N, float* degf, unsigned* zero, int zero_words, cudaStream_t st) {
  const int64_t n = N > zero_words ? N : zero_words;
  if (n <= 0) return HEXGNN_OK;
  lw_prepare_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(rowptr, N, degf, zero, zero_words);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

int debug_read_lw_phases(long long* host, int n) {
#ifdef HEXGNN_PHASES
  if (n > 8) n = 8;
  HEX_CUDA_CHECK(cudaMemcpyFromSymbol(host, g_lw_phase, sizeof(long long) * n));
  unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  HEX_CUDA_CHECK(cudaMemcpyToSymbol(g_lw_phase, z, sizeof(z)));
  return HEXGNN_OK;
#else
  (void)host; (void)n;
  set_error("debug_phases: the library was built without HEXGNN_PHASES");
  return HEXGNN_E_UNSUPPORTED;
#endif
}

int lw_layer0_blocks(int64_t N) { return (int)((N + kL0Rows - 1) / kL0Rows); }
int launch_lw_layer0(const LwLayer0Args& a, cudaStream_t st) {
  if (a.Cin < 1 || a.Cin > 4) { set_error("lw_layer0: C_in=%d", a.Cin); return HEXGNN_E_UNSUPPORTED; }
  if (a.N <= 0) return HEXGNN_OK;
  return launch_pdl(lw_layer0_kernel, dim3((unsigned)lw_layer0_blocks(a.N)), dim3(kL0Threads), 0, st, a);
}

int lw_norm_blocks(int64_t N) { return (int)((N + kNormRows - 1) / kNormRows); }
int launch_lw_norm(const LwNormArgs& a, cudaStream_t st) {
  if (a.N <= 0) return HEXGNN_OK;
  return launch_pdl(lw_norm_kernel, dim3((unsigned)lw_norm_blocks(a.N)), dim3(kNormThreads), 0, st, a);
}

}  // namespace hexgnn
