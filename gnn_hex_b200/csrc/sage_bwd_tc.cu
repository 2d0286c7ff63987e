// Fused backward of modern_two_headed on the tensor cores, part A: the dH chain.
//
// One persistent CTA per SM, one 128-row tile (whole graphs) at a time, same warp roles as the forward
// (sage_fwd_tc.cu).  Reverse mode per SAGE layer s (output h_{s+1}, input h_s), aggregate first:
//     dpre  = dH_{s+1} (.) (h_{s+1} > 0)                         ReLU mask from the stash image
//     dZl_j = sum_{i in N(j)} dpre_i / deg(i)                    the graph is symmetric: the adjoint of the
//                                                                mean-gather is again a CSR gather, no atomics
//     dH_s  = dpre * W_r + dZl * W_l                             two 3xTF32 tcgen05 GEMMs, A in TMEM, D in TMEM
// and [dZl | dpre] is written to the dZ stash (RM32 images) for part B (sage_dw_tc.cu), which contracts it
// with h_s over ALL rows of ALL tiles: dW_s = h_s^T [dZl | dpre] (the bias gradient rides on the ones-channel).
// This kernel also differentiates the dueling head and layer 0 (2 -> H) on the CUDA cores; their parameter
// gradients accumulate in shared memory and leave as one small per-CTA record, except the value MLP's first
// weight (H/2 x 4H): each graph emits [dt | pooled] and reduce_tc_kernel (pack_tc.cu) sums the outer products in
// graph order.  max/min pooling route their gradient to the FIRST row attaining the extremum (torch_scatter CPU
// semantics).  The ReLU masks come from the forward's sign-mask words (sage_tc.cuh): the chain never reads the
// activation images except the last layer's (the head input).
// Epilogue mappings: E1/E3 thread = (node row, column slice) for TMEM traffic; E2 (the CSR gather) warp = row,
// lane = 4 channels, so every shared-memory access is a contiguous row.
#include "sage_tc.cuh"

namespace hexgnn {
using namespace tc;

constexpr int kBwdStages = 8;

struct BwdTcArgs {
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
  const float* stash;     // forward stash (RM32 images + embedding copy)
  float* dz;              // dZ stash written here: [Lt-1][B][dz_blocks * 16 KB]
  float* d_embeds;
  float* d_x;             // [N,3] dL/dx (rl_loop/explain.py:26-36 saliency), or NULL
  TcPackView pv;
  float* small_rec;       // [tiles][gl.total - gl.gb]  (the gv0 range stays unwritten): one record per TILE, so the
                          // gradient bits do not depend on which CTA processed which tile (schedule-independent)
  float* head_rec;        // [B][tc_head_rec_floats]: per graph [dt (H2, padded) | pooled (4H)]
  int* tile_flag;         // [B]: set (release) once every dZ image of the tile is in memory -- the dW kernel's cue
  TdFuse td;              // hexgnn_train_step: forward-tile flags + the TD loss whose dL/dQ replaces d_nodes
};

template <int NJ>
struct BwdTcSmem {
  static constexpr int Hp = 16 * NJ;
  static constexpr int P = Hp + 4;
  static constexpr int ECAP = kEdgeCapPerRow * kTcRows;
  static constexpr int half_bytes = kTcKC * Hp * 2;      // fp16 hi | fp16 2^11 lo part of a (which, 16-c chunk) stage
  static constexpr int stage_bytes = 2 * half_bytes;
  static constexpr int off_ring = 0;
  static constexpr int off_dump = off_ring + kBwdStages * stage_bytes;           // float [128][P]
  static constexpr int off_agg = off_dump + (kTcRows + 1) * P * 4;               // float [128][P]  (dump row 128 = zeros)
  static constexpr int off_rows = off_agg + kTcRows * P * 4;                     // float [8][128]
  static constexpr int off_pool = off_rows + 8 * kTcRows * 4;                    // float [4*Hp]
  static constexpr int off_hid = off_pool + 4 * Hp * 4;                          // float [Hp]
  static constexpr int off_arg = off_hid + Hp * 4;                               // int [2*Hp]
  static constexpr int off_part = off_agg;                                       // float [4][6][Hp] row-part partials
                                                                                 // (head / layer 0 only: aliases agg)
  static constexpr int off_acc = off_arg + 2 * Hp * 4;                           // float [8*Hp + 16] CTA accumulators
  static constexpr int off_misc = off_acc + (8 * Hp + 16) * 4;                   // float [64]: [0..3] scalars, [4..36) per-warp max |dpre| of the last two layers
  static constexpr int off_rowptr = off_misc + 256;                              // int [128+4]
  static constexpr int off_gp = off_rowptr + (kTcRows + 4) * 4;                  // int [64+4]
  static constexpr int off_col = off_gp + (kMaxTileGraphs + 4) * 4;              // u16 [ECAP]
  static constexpr int off_gtab = (off_col + ECAP * 2 + 15) & ~15;               // gather table (sage_tc.cuh)
  static constexpr int off_bar = off_gtab + gather_tab_bytes();                  // u64 [2*stages + 6], u32 slot
  static constexpr int bytes = off_bar + (2 * kBwdStages + 6) * 8 + 16;
};

struct TileDimsB { int g0, g1, r0, n, e0, ne; bool ok; };
__device__ __forceinline__ TileDimsB tile_dims_b(const CsrView& v, int t, int ecap) {
  TileDimsB d;
  d.g0 = v.tile_g0[t];
  d.g1 = v.tile_g0[t + 1];
  d.r0 = v.gptr[d.g0];
  const int r1 = v.gptr[d.g1];
  d.n = r1 - d.r0;
  d.e0 = v.rowptr[d.r0];
  d.ne = v.rowptr[r1] - d.e0;
  d.ok = d.n >= 0 && d.n <= kTcRows && d.ne >= 0 && d.ne <= ecap && (d.g1 - d.g0) <= kMaxTileGraphs;
  return d;
}
__device__ __forceinline__ void epi_bar_b() { asm volatile("bar.sync 1, %0;" ::"n"(kTcEpiThreads) : "memory"); }
__device__ __forceinline__ float warp_sum_b(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Phase profile (cycles per CTA): [0] wait dH (MMA_l)  [1] E1 (mask, dpre)  [2] E2 gather  [3] wait MMA_r + A write
// [4] head + load (cumulative: [6] up to the tile load, [7] up to table build + tanh)  [5] layer 0  [8] total  [9] tiles
__device__ long long g_bwd_phase[148 * 16];

template <int NJ>
__global__ void __launch_bounds__(kTcThreads, 1) sage_bwd_tc_kernel(const __grid_constant__ BwdTcArgs a) {
  using S = BwdTcSmem<NJ>;
  constexpr int Hp = S::Hp, P = S::P;
  constexpr int NG = 2 * NJ;
  constexpr int NQ = Hp / kTcKC;
  constexpr int MAXG = (NG + kTcSlices - 1) / kTcSlices;
  constexpr int NV = Hp / 4, PV = P / 4;
  // TMEM columns: A (fp16 pairs; per 16-c chunk 8 columns of 2^11 lo, then 8 columns of hi) | D | Ds: the all-fp16 3-product
  // scheme of sage_tc.cuh -- D takes the large products, Ds the 2^11-scaled small ones (the tensor core TRUNCATES its fp32
  // accumulator after every MMA; only the 2 x Hp / 16 large products truncate the large sum), dH = D + 2^-11 Ds.
  // Loss gradients are ~1e-6 .. 1e-3, far below fp16's normal range, so the A operand of a (tile, layer) is scaled by a power of
  // two 2^S chosen from the tile's max |dpre| (to ~2^9: dZl, a sum over neighbours, has 2^6 head-room to fp16's 65504) and dH is
  // scaled back by 2^-S when it is read: exact, local to the two GEMMs, every other quantity stays as it was.
  constexpr uint32_t TM_A = 0, TM_D = Hp, TM_DS = 2 * Hp;
  constexpr uint32_t TM_COLS = (3 * Hp <= 64) ? 64 : (3 * Hp <= 128) ? 128 : (3 * Hp <= 256) ? 256 : 512;
  extern __shared__ __align__(128) uint8_t smem[];
  float* dump = (float*)(smem + S::off_dump);
  float* agg = (float*)(smem + S::off_agg);
  float* partb = (float*)(smem + S::off_part);
  float* accs = (float*)(smem + S::off_acc);
  float* rows = (float*)(smem + S::off_rows);
  float *inv_s = rows, *x0_s = rows + 128, *x1_s = rows + 256, *m0_s = rows + 384, *m1_s = rows + 512,
        *adv_s = rows + 640, *th_s = rows + 768, *dq_s = rows + 896;
  float* pool = (float*)(smem + S::off_pool);
  float* hid = (float*)(smem + S::off_hid);
  int* arg_s = (int*)(smem + S::off_arg);
  float* misc = (float*)(smem + S::off_misc);
  int* rowptr_s = (int*)(smem + S::off_rowptr);
  int* gp_s = (int*)(smem + S::off_gp);
  unsigned short* col_s = (unsigned short*)(smem + S::off_col);
  const GatherTab gtab = gather_tab_view(smem + S::off_gtab);
  uint64_t* bar_full = (uint64_t*)(smem + S::off_bar);
  uint64_t* bar_empty = bar_full + kBwdStages;
  uint64_t* bar_a = bar_empty + kBwdStages;   // [4]: column quarter gi of the A operand is in TMEM
  uint64_t* bar_r = bar_a + 4;
  uint64_t* bar_d = bar_r + 1;
  uint32_t* tmem_slot = (uint32_t*)(bar_d + 1);

  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int H = a.d.H, H2 = a.d.H2, Lt = a.d.Lt;
  const int ntiles = *a.csr.ntiles;
  const long long t_kernel0 = PH_NOW();
  long long ph0 = 0, ph1 = 0, ph2 = 0, ph3 = 0, ph4 = 0, ph5 = 0, ph6 = 0, ph7 = 0, ntl = 0;
  long long hq0 = 0, hq1 = 0, hq2 = 0, hq3 = 0, hq4 = 0, hq5 = 0;   // head sub-phases (per graph loop)

  if (tid == 0) {
    for (int i = 0; i < kBwdStages; ++i) { mbar_init(&bar_full[i], 1); mbar_init(&bar_empty[i], 1); }
    for (int i = 0; i < 4; ++i) mbar_init(&bar_a[i], kTcEpiWarps);
    mbar_init(bar_r, 1);
    mbar_init(bar_d, 1);
    fence_mbar_init();
  }
  if (wid == kTcEpiWarps + 1) tmem_alloc(tmem_slot, TM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = *tmem_slot;
  // Every CTA of this grid is resident by now (persistent grid <= SM count): let the weight-gradient GEMM's CTAs take
  // the SMs of CTAs that run out of tiles; they consume tiles as the flags below are released.
  pdl_launch_dependents();

  if (wid == kTcEpiWarps) {
    // ================= producer: backward weight images -> ring =================
    if (lane == 0) {
      uint32_t it = 0;
      for (int rnd = 0; rnd * (int)gridDim.x < ntiles; ++rnd) {
      const int ti = tile_position(rnd, blockIdx.x, gridDim.x, a.td.fwd_flag == nullptr);
      if (ti >= ntiles) continue;
      const int t = a.csr.tile_order[ti];
        if (!tile_dims_b(a.csr, t, S::ECAP).ok) continue;
        for (int s = Lt - 1; s >= 1; --s) {
          const uint8_t* src = (const uint8_t*)a.pv.wimgT + (size_t)(s - 1) * 2 * NQ * S::stage_bytes;
          for (int q = 0; q < 2 * NQ; ++q, ++it) {   // which 0 (lin_r) chunks, then which 1 (lin_l) chunks
            const int st = it % kBwdStages;
            mbar_wait(&bar_empty[st], ((it / kBwdStages) & 1) ^ 1);
            mbar_expect_tx(&bar_full[st], S::stage_bytes);
            bulk_g2s(smem + S::off_ring + st * S::stage_bytes, src + (size_t)q * S::stage_bytes, S::stage_bytes,
                     &bar_full[st]);
          }
        }
      }
    }
  } else if (wid == kTcEpiWarps + 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      constexpr uint32_t idesc16 = make_idesc_f16(kTcRows, Hp);
      constexpr uint32_t lbo = Hp * 16, sbo = 128;
      uint32_t it = 0, na = 0;
      for (int rnd = 0; rnd * (int)gridDim.x < ntiles; ++rnd) {
      const int ti = tile_position(rnd, blockIdx.x, gridDim.x, a.td.fwd_flag == nullptr);
      if (ti >= ntiles) continue;
      const int t = a.csr.tile_order[ti];
        if (!tile_dims_b(a.csr, t, S::ECAP).ok) continue;
        for (int s = Lt - 1; s >= 1; --s) {
          for (int which = 0; which < 2; ++which) {
            for (int q = 0; q < NQ; ++q, ++it) {
              if ((q & 1) == 0) {
                // granules 2q .. 2q+3 of the A operand (dpre, then dZl) are in TMEM; the epilogue releases it quarter
                // by quarter, so these MMAs run under the rest of its E1 / E3
                mbar_wait(&bar_a[q >> 1], na & 1);
                tc_fence_after();
              }
              const int st = it % kBwdStages;
              mbar_wait(&bar_full[st], (it / kBwdStages) & 1);
              tc_fence_after();
              const uint32_t b0 = smem_u32(smem + S::off_ring + st * S::stage_bytes);
              const uint64_t bhi = make_smem_desc(b0, lbo, sbo), blo = make_smem_desc(b0 + S::half_bytes, lbo, sbo);
              const uint32_t alo = tbase + TM_A + q * 16, ahi = alo + 8;
              mma_f16_ts(tbase + TM_DS, alo, bhi, idesc16, (which | q) != 0);
              mma_f16_ts(tbase + TM_DS, ahi, blo, idesc16, true);
              mma_f16_ts(tbase + TM_D, ahi, bhi, idesc16, (which | q) != 0);
              mma_commit(&bar_empty[st]);
            }
            mma_commit(which == 0 ? bar_r : bar_d);
            ++na;
          }
        }
      }
    }
  } else {
    // ================= epilogue warps =================
    const int q4 = wid & 3, slice = wid >> 2;
    const int row = q4 * 32 + lane;
    const uint32_t trow = tbase + ((uint32_t)(q4 * 32) << 16);
    const float flag = a.x[2];
    const int head = (flag == 1.0f) ? 0 : 1;
    const float* Pm = a.params;
    const int64_t recn = a.gl.total - a.gl.gb;
    constexpr int A_GB0 = 0, A_G0 = Hp, A_GLIN = 5 * Hp, A_GLINB = 6 * Hp, A_GV0B = 6 * Hp + 4, A_GV1 = 7 * Hp + 4,
                  A_GV1B = 8 * Hp + 4, A_N = 8 * Hp + 16;
    // per-tile accumulators of the small parameter gradients (every element has ONE owner thread), flushed to the
    // tile's record and cleared at the end of the tile
    auto flush_tile_record = [&](int t) {
      float* rec = a.small_rec + (int64_t)t * recn - a.gl.gb;   // index with GradLayout offsets (>= gl.gb)
      for (int i = tid; i < Hp; i += kTcEpiThreads) { rec[a.gl.gb + i] = accs[A_GB0 + i]; rec[a.gl.glin + i] = accs[A_GLIN + i]; }
      for (int i = tid; i < 4 * Hp; i += kTcEpiThreads) rec[a.gl.g0 + i] = accs[A_G0 + i];
      for (int i = tid; i < H2; i += kTcEpiThreads) { rec[a.gl.gv0b + i] = accs[A_GV0B + i]; rec[a.gl.gv1 + i] = accs[A_GV1 + i]; }
      if (tid == 0) { rec[a.gl.glinb] = accs[A_GLINB]; rec[a.gl.gv1b] = accs[A_GV1B]; }
      epi_bar_b();
      for (int i = tid; i < A_N; i += kTcEpiThreads) accs[i] = 0.f;
    };
    for (int i = tid; i < A_N; i += kTcEpiThreads) accs[i] = 0.f;
    for (int i = tid; i < P; i += kTcEpiThreads) dump[kTcRows * P + i] = 0.f;   // the gather's zero row
    const float* wa = Pm + a.pm.lin_w[head];
    const float ba = Pm[a.pm.lin_b[head]];
    const float* w0 = Pm + a.pm.v0_w[head];
    const float* b0 = Pm + a.pm.v0_b[head];
    const float* w1 = Pm + a.pm.v1_w[head];
    const float b1 = Pm[a.pm.v1_b[head]];
    const size_t act_img = (size_t)tc_act_blocks(Hp) * kImgBlockBytes;
    const size_t dz_img = (size_t)tc_dz_blocks(Hp) * kImgBlockBytes;
    const uint32_t* mask_base = (const uint32_t*)(a.stash + tc_stash_mask_off_floats(a.d, a.N, a.B));
    const size_t mask_layer = (size_t)a.B * kTcSlices * kTcRows;
    const int hrec_n = tc_head_rec_floats(a.d), H2p = (H2 + 3) & ~3;
    const int part = tid >> 7, pc = tid & 127;      // row quarter / channel of the column-parallel loops
    uint32_t nr = 0, nd = 0;
    int range_bad = 0;
    float unscale = 1.f;                    // 2^-S of the scale the A operands of the GEMMs behind the current D / Ds carried
    // one 8-channel granule of the A operand, scaled by sc = 2^S: fp16 hi and 2^11-scaled lo halves, 4 packed columns each, into
    // the granule's half of its 16-c chunk ([lo(2q) lo(2q+1) | hi(2q) hi(2q+1)])
    auto store_A = [&](int g, const float (&v)[8], float sc) {
      uint32_t hi4[4], lo4[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float v0 = v[2 * j] * sc, v1 = v[2 * j + 1] * sc;
        split_f16x2(v0, v1, hi4[j], lo4[j]);
        if (fabsf(v0) >= 60000.f || fabsf(v1) >= 60000.f) range_bad = 1;
      }
      const uint32_t c = trow + TM_A + (g >> 1) * 16 + (g & 1) * 4;
      tmem_st4u(c, lo4);
      tmem_st4u(c + 8, hi4);
    };
    // dH of this thread's granules = (D + 2^-11 Ds) * unscale
    auto load_dH = [&](float (&dh)[MAXG][8]) {
#pragma unroll
      for (int gi = 0; gi < MAXG; ++gi) {
        const int g = slice + gi * kTcSlices;
        if (g < NG) tmem_ld8(trow + TM_D + g * 8, dh[gi]);
      }
      tmem_wait_ld();
      const float us = unscale, uss = unscale * (1.0f / 2048.0f);
#pragma unroll
      for (int gi = 0; gi < MAXG; gi += 2) {       // + the small-product accumulator, two granules per round trip
        float d0[8], d1[8];
        const int g0_ = slice + gi * kTcSlices, g1_ = slice + (gi + 1) * kTcSlices;
        if (g0_ < NG) tmem_ld8(trow + TM_DS + g0_ * 8, d0);
        if (gi + 1 < MAXG && g1_ < NG) tmem_ld8(trow + TM_DS + g1_ * 8, d1);
        tmem_wait_ld();
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          if (g0_ < NG) dh[gi][u] = fmaf(d0[u], uss, dh[gi][u] * us);
          if (gi + 1 < MAXG && g1_ < NG) dh[gi + 1][u] = fmaf(d1[u], uss, dh[gi + 1][u] * us);
        }
      }
    };
    for (int rnd = 0; rnd * (int)gridDim.x < ntiles; ++rnd) {
      const int ti = tile_position(rnd, blockIdx.x, gridDim.x, a.td.fwd_flag == nullptr);
      if (ti >= ntiles) continue;
      const int t = a.csr.tile_order[ti];
      const TileDimsB td = tile_dims_b(a.csr, t, S::ECAP);
      const int n = td.n, r0 = td.r0, g0 = td.g0, g1 = td.g1;
      if (a.td.fwd_flag && tid == 0) flag_wait(a.td.fwd_flag + t);   // acquire; the barrier extends it to the CTA
      epi_bar_b();
      long long tq = PH_NOW();
      ++ntl;
      if (!td.ok) {
        if (tid == 0) { accs[A_GB0] = __int_as_float(0x7fc00000); atomicOr(a.csr.status, HEXGNN_ST_GRAPH_TOO_LARGE); }
        for (int i = tid; i < (g1 - g0) * hrec_n; i += kTcEpiThreads) a.head_rec[(int64_t)g0 * hrec_n + i] = 0.f;
        epi_bar_b();
        flush_tile_record(t);               // poisoned record
        if (tid == 0) flag_release(a.tile_flag + t, 1);
        continue;
      }
      // ---- tile CSR + features -> shared ------------------------------------------------------------
      {
        int bad = 0;
        for (int i = tid; i <= kTcRows; i += kTcEpiThreads) rowptr_s[i] = (i <= n ? a.csr.rowptr[r0 + i] : a.csr.rowptr[r0 + n]) - td.e0;
        for (int g = tid; g <= g1 - g0; g += kTcEpiThreads) gp_s[g] = a.csr.gptr[g0 + g] - r0;
        for (int e = tid; e < td.ne; e += kTcEpiThreads) {
          int c = a.csr.col[td.e0 + e] - r0;
          if (c < 0 || c >= n) { bad |= HEXGNN_ST_EDGE_CROSSES; c = 0; }
          col_s[e] = (unsigned short)c;
        }
        for (int i = tid; i < kTcRows; i += kTcEpiThreads) {
          float v0 = 0.f, v1 = 0.f;
          if (i < n) { const float* xr = a.x + (int64_t)(r0 + i) * 3; v0 = xr[0]; v1 = xr[1]; }
          x0_s[i] = v0; x1_s[i] = v1;
          dq_s[i] = (a.td.action || i >= n) ? 0.f : a.d_nodes[r0 + i];
        }
        if (bad) atomicOr(a.csr.status, bad);
      }
      const bool live = row < n;
      const int n32 = (n + 31) & ~31;
      const char* act_t = (const char*)a.stash + (size_t)t * act_img;         // + layer * B * act_img
      char* dz_t = (char*)a.dz + (size_t)t * dz_img;                           // + (s-1) * B * dz_img
      const uint32_t* mask_t = mask_base + ((size_t)t * kTcSlices + slice) * kTcRows + row;
      // ---- g = h_Lt -> dump (row major) -----------------------------------------------------------------
      {
        const char* img = act_t + (size_t)(Lt - 1) * a.B * act_img;
        for (int g = slice; g < NG; g += kTcSlices) {
          float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
          if (row < n32) { const float4* src = (const float4*)(img + rm32_granule_off(g, row)); v0 = src[0]; v1 = src[1]; }
          float4* dst = (float4*)(dump + row * P + g * 8);
          dst[0] = v0; dst[1] = v1;
        }
      }
      epi_bar_b();
      PH_ADD(ph6, tq);
      if (a.td.action) {
        // dL/dQ of the TD loss (td_loss.cu): zero except at row ptr[g] + action[g] of every graph
        const float invB = a.td.scale / (float)a.B;
        for (int g = tid; g < g1 - g0; g += kTcEpiThreads) {
          const int64_t act = a.td.action[g0 + g];
          const int i = gp_s[g] + (int)act;
          if (act >= 0 && i < gp_s[g + 1]) {
            const float e = a.td.q[r0 + i] - a.td.target[g0 + g];
            const float w = a.td.weight ? a.td.weight[g0 + g] : 1.f;
            float dl;
            if (a.td.kind == 0) dl = 2.f * e;
            else { const float ab = fabsf(e); dl = ab < 1.f ? e : (e > 0.f ? 1.f : -1.f); }
            dq_s[i] += w * dl * invB;
          }
        }
      }
      gather_build<NJ>(gtab, rowptr_s, col_s, n, tid);
      for (int i = tid; i < kTcRows; i += kTcEpiThreads) {
        const int b = rowptr_s[i], e = rowptr_s[i + 1];
        const int deg = e - b;
        const float inv = 1.0f / (float)(deg > 0 ? deg : 1);
        float s0 = 0.f, s1 = 0.f;
        for (int p = b; p < e; ++p) { const int j = col_s[p]; s0 += x0_s[j]; s1 += x1_s[j]; }
        inv_s[i] = inv; m0_s[i] = s0 * inv; m1_s[i] = s1 * inv;
      }
      // th_i = tanh(w_a . g_i + b_a): stashed by the forward
      for (int i = tid; i < kTcRows; i += kTcEpiThreads)
        th_s[i] = a.stash[tc_stash_th_off_floats(a.d, a.N, a.B) + (int64_t)t * kTcRows + i];
      epi_bar_b();
      PH_ADD(ph7, tq);
      const float inv = inv_s[row];
      // ---- head backward per graph: dump rows become dL/dg in place -----------------------------------------
      for (int g = 0; g < g1 - g0; ++g) {
        const int rs = gp_s[g], re = gp_s[g + 1];
        const int ng = re - rs;
        float* hrec = a.head_rec + (int64_t)(g0 + g) * hrec_n;
        const bool adv_only = (a.flags & HEXGNN_ADV_ONLY) != 0;
        if (ng <= 0 || adv_only) {
          for (int i = tid; i < hrec_n; i += kTcEpiThreads) hrec[i] = 0.f;
          if (ng <= 0) continue;
        }
        long long th = PH_NOW();
        float sumdq = 0.f;
        if (!adv_only) {
          if (wid == 0) {
            float u = 0.f;
            for (int i = rs + lane; i < re; i += 32) u += dq_s[i];
            u = warp_sum_b(u);
            if (lane == 0) misc[0] = u;
          }
          epi_bar_b();
          sumdq = misc[0];
        }
        for (int i = rs + tid; i < re; i += kTcEpiThreads) {
          const float th = th_s[i];
          const float dn = dq_s[i];
          adv_s[i] = (adv_only ? dn : dn - sumdq / (float)ng) * 2.0f * (1.0f - th * th);
        }
        epi_bar_b();
        PH_ADD(hq0, th); th = PH_NOW();
        // column-parallel pass over the graph's rows, 4 row quarters: pools, arg rows, d head.linear.weight
        if (pc < H) {
          const int chunk = (ng + 3) >> 2;
          const int ia = rs + part * chunk, ib = (ia + chunk < re) ? ia + chunk : re;
          float sum = 0.f, mx = -INFINITY, mn = INFINITY, gl_ = 0.f;
          int amx = rs, amn = rs;
          for (int i = ia; i < ib; ++i) {
            const float v = dump[i * P + pc];
            sum += v;
            gl_ = fmaf(adv_s[i], v, gl_);
            if (v > mx) { mx = v; amx = i; }
            if (v < mn) { mn = v; amn = i; }
          }
          float* pp = partb + (part * 6) * Hp + pc;
          pp[0] = sum; pp[Hp] = mx; pp[2 * Hp] = mn; pp[3 * Hp] = gl_;
          pp[4 * Hp] = __int_as_float(amx); pp[5 * Hp] = __int_as_float(amn);
        }
        epi_bar_b();
        if (tid < H) {
          float sum = 0.f, mx = -INFINITY, mn = INFINITY, gl_ = 0.f;
          int amx = rs, amn = rs;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float* pp = partb + (q * 6) * Hp + tid;
            sum += pp[0];
            gl_ += pp[3 * Hp];
            if (pp[Hp] > mx) { mx = pp[Hp]; amx = __float_as_int(pp[4 * Hp]); }
            if (pp[2 * Hp] < mn) { mn = pp[2 * Hp]; amn = __float_as_int(pp[5 * Hp]); }
          }
          pool[tid] = sum; pool[H + tid] = mx; pool[2 * H + tid] = mn; pool[3 * H + tid] = sum / (float)ng;
          arg_s[tid] = amx; arg_s[Hp + tid] = amn;
          accs[A_GLIN + tid] += gl_;
        } else if (wid == kTcEpiWarps - 1) {
          float u = 0.f;
          for (int i = rs + lane; i < re; i += 32) u += adv_s[i];
          u = warp_sum_b(u);
          if (lane == 0) accs[A_GLINB] += u;
        }
        epi_bar_b();
        PH_ADD(hq1, th); th = PH_NOW();
        if (!adv_only) {
          // hidden activations of the value MLP: stashed by the forward (same bits as recomputing them)
          if (tid < H2) hid[tid] = a.stash[tc_stash_hid_off_floats(a.d, a.N, a.B) + (int64_t)(g0 + g) * H2p + tid];
          epi_bar_b();
          PH_ADD(hq2, th); th = PH_NOW();
          if (wid == kTcEpiWarps - 1) {
            float u = 0.f;
            for (int k = lane; k < H2; k += 32) u = fmaf(w1[k], hid[k], u);
            u = warp_sum_b(u);
            if (lane == 0) {
              const float v = tanhf(u + b1);
              const float dv = (a.flags & HEXGNN_SEPARATE) ? (a.d_value ? a.d_value[g0 + g] : 0.f) : sumdq;
              misc[1] = dv * (1.0f - v * v);
            }
          }
          epi_bar_b();
          const float ds = misc[1];
          if (tid < H2) {
            const float hk = hid[tid];
            accs[A_GV1 + tid] += ds * hk;
            const float dt = hk > 0.f ? ds * w1[tid] : 0.f;
            accs[A_GV0B + tid] += dt;
            hid[tid] = dt;
          }
          if (tid == 0) accs[A_GV1B] += ds;
          epi_bar_b();
          PH_ADD(hq3, th); th = PH_NOW();
          // graph record [dt | pooled]: the reduction kernel sums dt (x) pooled over graphs in graph order
          for (int i = tid; i < hrec_n; i += kTcEpiThreads)
            hrec[i] = i < H2p ? (i < H2 ? hid[i] : 0.f) : (i - H2p < 4 * H ? pool[i - H2p] : 0.f);
          float dp = 0.f;
          if (tid < 4 * H) {
            // 14 independent L2 loads in flight per batch (the weights are 97 KB: they do not live in L1); k ascending
            for (int k = 0; k < H2; k += 14) {
              float wv[14];
#pragma unroll
              for (int u = 0; u < 14; ++u) wv[u] = (k + u < H2) ? w0[(int64_t)(k + u) * 4 * H + tid] : 0.f;
#pragma unroll
              for (int u = 0; u < 14; ++u) dp = fmaf(wv[u], (k + u < H2) ? hid[k + u] : 0.f, dp);
            }
          }
          epi_bar_b();  // pooled values fully consumed before they become their own gradient
          if (tid < 4 * H) pool[tid] = dp;
          epi_bar_b();
          PH_ADD(hq4, th); th = PH_NOW();
        }
        // dump rows become dL/dg: thread = (row quarter, channel), the per-channel terms live in registers
        if (pc < Hp) {
          const int chunk = (ng + 3) >> 2;
          const int ia = rs + part * chunk, ib = (ia + chunk < re) ? ia + chunk : re;
          const bool cv = pc < H;
          float wa_c = 0.f, base = 0.f, vmx = 0.f, vmn = 0.f;
          int amx = -1, amn = -1;
          if (cv) {
            wa_c = wa[pc];
            if (!adv_only) {
              base = pool[pc] + pool[3 * H + pc] / (float)ng;
              amx = arg_s[pc]; amn = arg_s[Hp + pc];
              vmx = pool[H + pc]; vmn = pool[2 * H + pc];
            }
          }
          for (int i = ia; i < ib; ++i) {
            float v = 0.f;
            if (cv) {
              v = adv_s[i] * wa_c;
              if (!adv_only) {
                v += base;
                if (amx == i) v += vmx;
                if (amn == i) v += vmn;
              }
            }
            dump[i * P + pc] = v;
          }
        }
        epi_bar_b();
        PH_ADD(hq5, th);
      }
      PH_ADD(ph4, tq);
      // ---- SAGE layers Lt-1 .. 1 on the tensor cores ---------------------------------------------------------
      for (int s = Lt - 1; s >= 1; --s) {
        const bool first = (s == Lt - 1);
        tq = PH_NOW();
        const uint32_t mw = mask_t[(size_t)s * mask_layer];   // signs of h_{s+1}; in flight during the wait below
        float dh[MAXG][8];
        if (first) {
#pragma unroll
          for (int gi = 0; gi < MAXG; ++gi) {
            const int g = slice + gi * kTcSlices;
            if (g < NG) {
              const float4* src = (const float4*)(dump + row * P + g * 8);
              const float4 v0 = src[0], v1 = src[1];
              dh[gi][0] = v0.x; dh[gi][1] = v0.y; dh[gi][2] = v0.z; dh[gi][3] = v0.w;
              dh[gi][4] = v1.x; dh[gi][5] = v1.y; dh[gi][6] = v1.z; dh[gi][7] = v1.w;
            }
          }
        } else {
          mbar_wait(bar_d, nd & 1);   // dH_{s+1} complete in TMEM
          ++nd;
          tc_fence_after();
          PH_ADD(ph0, tq);
          tq = PH_NOW();
          load_dH(dh);
        }
        // E1: ReLU mask, dpre -> TMEM (A of dpre*W_r), dump (pre-scaled by 1/deg for the gather), dZ stash.  The power-of-two
        // scale 2^S of this layer's two GEMMs puts the tile's max |dpre| near 2^6: measured exactly for the first layer (one
        // extra barrier per tile), afterwards taken from the PREVIOUS layer's max (published by every warp, read a layer
        // later: no barrier) -- fp16 leaves 2^10 of head-room above 2^6, i.e. the gradient of a tile may grow 2^4 from one
        // layer to the next AND dZl (a sum over neighbours) may be 2^6 larger than dpre before anything is flagged.
        float ascale;
        {
          char* zimg = dz_t + (size_t)(s - 1) * a.B * dz_img;
          float mx = 0.f;
#pragma unroll
          for (int gi = 0; gi < MAXG; ++gi) {
            const int g = slice + gi * kTcSlices;
            if (g < NG) {
              if (s == a.d.L - 1 && a.d_embeds && live) {
                float4* de = (float4*)(a.d_embeds + (int64_t)(r0 + row) * Hp + g * 8);
                de[0] = make_float4(dh[gi][0], dh[gi][1], dh[gi][2], dh[gi][3]);
                de[1] = make_float4(dh[gi][4], dh[gi][5], dh[gi][6], dh[gi][7]);
              }
#pragma unroll
              for (int u = 0; u < 8; ++u) {
                const float r = ((mw >> (gi * 8 + u)) & 1u) ? dh[gi][u] : 0.f;
                dh[gi][u] = r;
                mx = fmaxf(mx, fabsf(r));
              }
            }
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
          float* mxbuf = misc + 4 + 16 * (s & 1);       // this layer's per-warp maxima; the other half holds the previous layer's
          if (lane == 0) mxbuf[wid] = mx;
          if (first) epi_bar_b();                       // first layer of the tile: its own maximum
          mx = (first ? mxbuf : misc + 4 + 16 * ((s + 1) & 1))[lane & 15];
#pragma unroll
          for (int o = 8; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
          // 2^S with max * 2^S in [2^6, 2^7); a tile without gradient (or a non-finite one) keeps S = 0
          int e = (int)((__float_as_uint(mx) >> 23) & 0xffu);
          if (e < 10 || e == 255) e = 127 + 6;
          ascale = __uint_as_float((uint32_t)(127 + 6 + 127 - e) << 23);
          unscale = __uint_as_float((uint32_t)(e - 6) << 23);
#pragma unroll
          for (int gi = 0; gi < MAXG; ++gi) {
            const int g = slice + gi * kTcSlices;
            if (g < NG) {
              store_A(g, dh[gi], ascale);
              float4* dd = (float4*)(dump + row * P + g * 8);
              dd[0] = make_float4(dh[gi][0] * inv, dh[gi][1] * inv, dh[gi][2] * inv, dh[gi][3] * inv);
              dd[1] = make_float4(dh[gi][4] * inv, dh[gi][5] * inv, dh[gi][6] * inv, dh[gi][7] * inv);
              if (row < n32) {
                st_global_v8(zimg + rm32_granule_off(NG + g, row), dh[gi][0], dh[gi][1], dh[gi][2], dh[gi][3], dh[gi][4],
                             dh[gi][5], dh[gi][6], dh[gi][7]);
              }
            }
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar_a[gi]);   // -> MMA: dH_s = dpre * W_r, column quarter gi
          }
        }
        epi_bar_b();                 // dump (dpre / deg) complete
        PH_ADD(ph1, tq);
        tq = PH_NOW();
        // E2: dZl_j = sum_{i in N(j)} dpre_i / deg(i) (tile_gather, sage_tc.cuh); runs under the dpre * W_r MMAs.
        tile_gather<NJ>(gtab, dump, agg, n, wid, lane);
        epi_bar_b();
        PH_ADD(ph2, tq);
        tq = PH_NOW();
        // E3: dZl -> dZ stash, then (once dpre * W_r has read the A columns) -> TMEM as the next A operand
        {
          char* zimg = dz_t + (size_t)(s - 1) * a.B * dz_img;
          float acc[MAXG][8];
#pragma unroll
          for (int gi = 0; gi < MAXG; ++gi) {
            const int g = slice + gi * kTcSlices;
            if (g < NG) {
              float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
              if (live) { const float4* src = (const float4*)(agg + row * P + g * 8); v0 = src[0]; v1 = src[1]; }
              acc[gi][0] = v0.x; acc[gi][1] = v0.y; acc[gi][2] = v0.z; acc[gi][3] = v0.w;
              acc[gi][4] = v1.x; acc[gi][5] = v1.y; acc[gi][6] = v1.z; acc[gi][7] = v1.w;
              if (row < n32) {
                st_global_v8(zimg + rm32_granule_off(g, row), v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w);
              }
            }
          }
          mbar_wait(bar_r, nr & 1);      // dpre * W_r has finished reading the A columns
          ++nr;
          tc_fence_after();
#pragma unroll
          for (int gi = 0; gi < MAXG; ++gi) {
            const int g = slice + gi * kTcSlices;
            if (g < NG) store_A(g, acc[gi], ascale);
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar_a[gi]);   // -> MMA: dH_s += dZl * W_l, column quarter gi
          }
        }
        PH_ADD(ph3, tq);
      }
      // ---- layer 0 (2 -> H): dpre_0 = dH_1 (.) (h_1 > 0); rank-2 weight gradients on the CUDA cores -------------------
      tq = PH_NOW();
      {
        const uint32_t mw = mask_t[0];
        float dh[MAXG][8];
        if (Lt == 1) {
#pragma unroll
          for (int gi = 0; gi < MAXG; ++gi) {
            const int g = slice + gi * kTcSlices;
            if (g < NG) {
              const float4* src = (const float4*)(dump + row * P + g * 8);
              const float4 v0 = src[0], v1 = src[1];
              dh[gi][0] = v0.x; dh[gi][1] = v0.y; dh[gi][2] = v0.z; dh[gi][3] = v0.w;
              dh[gi][4] = v1.x; dh[gi][5] = v1.y; dh[gi][6] = v1.z; dh[gi][7] = v1.w;
            }
          }
        } else {
          mbar_wait(bar_d, nd & 1);
          ++nd;
          tc_fence_after();
          load_dH(dh);
        }
#pragma unroll
        for (int gi = 0; gi < MAXG; ++gi) {
          const int g = slice + gi * kTcSlices;
          if (g < NG) {
            if (a.d.L == 1 && a.d_embeds && live) {
              float4* de = (float4*)(a.d_embeds + (int64_t)(r0 + row) * Hp + g * 8);
              de[0] = make_float4(dh[gi][0], dh[gi][1], dh[gi][2], dh[gi][3]);
              de[1] = make_float4(dh[gi][4], dh[gi][5], dh[gi][6], dh[gi][7]);
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) dh[gi][u] = ((mw >> (gi * 8 + u)) & 1u) ? dh[gi][u] : 0.f;
            float4* dd = (float4*)(dump + row * P + g * 8);
            dd[0] = make_float4(dh[gi][0], dh[gi][1], dh[gi][2], dh[gi][3]);
            dd[1] = make_float4(dh[gi][4], dh[gi][5], dh[gi][6], dh[gi][7]);
          }
        }
        tc_fence_before();
        epi_bar_b();
        // column-parallel over 4 row quarters, combined in quarter order (fixed summation order)
        if (pc < Hp) {
          const int chunk = (n + 3) >> 2;
          const int ia = part * chunk, ib = (ia + chunk < n) ? ia + chunk : n;
          float ub = 0.f, u0 = 0.f, u1 = 0.f, u2 = 0.f, u3 = 0.f;
          for (int i = ia; i < ib; ++i) {
            const float dpv = dump[i * P + pc];
            ub += dpv;
            u0 = fmaf(dpv, m0_s[i], u0);
            u1 = fmaf(dpv, m1_s[i], u1);
            u2 = fmaf(dpv, x0_s[i], u2);
            u3 = fmaf(dpv, x1_s[i], u3);
          }
          float* pp = partb + (part * 6) * Hp + pc;
          pp[0] = ub; pp[Hp] = u0; pp[2 * Hp] = u1; pp[3 * Hp] = u2; pp[4 * Hp] = u3;
        }
        epi_bar_b();
        for (int i = tid; i < 5 * Hp; i += kTcEpiThreads) {   // i = k*Hp + c: k 0 -> gb0, 1..4 -> g0 rows
          float u = 0.f;
#pragma unroll
          for (int q = 0; q < 4; ++q) u += partb[(q * 6) * Hp + i];
          accs[i] += u;                                        // A_GB0 = 0, A_G0 = Hp: contiguous
        }
        if (a.d_x) {
          // dL/dx of the tile (input saliency, rl_loop/explain.py:26-36): with dpre_0 in `dump`,
          //   r_i = W_r0^T dpre_0[i],  u_i = W_l0^T dpre_0[i] / deg(i),  dx_i = r_i + sum_{j in N(i)} u_j
          // (the graph is symmetric: the adjoint of the mean-gather is a gather).  x[:,2] is the head selector: zero.
          // 4 threads per row, one of (u_0, u_1, r_0, r_1) each; the x / mean staging rows are free by now.
          const int i = tid >> 2, k = tid & 3;
          const float* w = a.params + (k < 2 ? sage_wl(a.pm, 0, 0) + k : sage_wr(a.pm, 0, 0, H) + (k - 2));
          float sacc = 0.f;
          if (i < n) {
            const float* dp = dump + i * P;
            for (int c = 0; c < H; ++c) sacc = fmaf(dp[c], __ldg(w + 2 * c), sacc);
            if (k < 2) sacc *= inv_s[i];
          }
          float* dst = k == 0 ? m0_s : k == 1 ? m1_s : k == 2 ? x0_s : x1_s;
          dst[i] = sacc;
          epi_bar_b();
          if (tid < 2 * kTcRows) {
            const int r = tid >> 1, q = tid & 1;
            if (r < n) {
              const float* us = q ? m1_s : m0_s;
              float v = (q ? x1_s : x0_s)[r];
              for (int p = rowptr_s[r]; p < rowptr_s[r + 1]; ++p) v += us[col_s[p]];
              a.d_x[(int64_t)(r0 + r) * 3 + q] = v;
            }
          } else if (tid < 3 * kTcRows) {
            const int r = tid - 2 * kTcRows;
            if (r < n) a.d_x[(int64_t)(r0 + r) * 3 + 2] = 0.f;
          }
        }
      }
      PH_ADD(ph5, tq);
      // an operand outside fp16's range (a gradient that grew > 2^10 between two layers): flag the batch and poison the
      // tile's record, like a tile that does not fit -- the fused optimizer skips a step whose gradient is not finite
      if (range_bad) { atomicOr(a.csr.status, HEXGNN_ST_FP16_RANGE); accs[A_GB0] = __int_as_float(0x7fc00000); range_bad = 0; }
      // ---- hand the tile to the dW GEMM: all dZ stores of all epilogue threads, then one release ----------------------
      fence_proxy_async_global();          // the consumer reads with bulk copies (async proxy)
      epi_bar_b();
      if (tid == 0) { __threadfence(); flag_release(a.tile_flag + t, 1); }
      flush_tile_record(t);
    }
    // the dW kernel's last ticket waits for every chain CTA's record, so "dW complete" implies "chain complete"
    epi_bar_b();
    if (tid == 0) { __threadfence(); atomicAdd(a.tile_flag - kSyncTileOff + 1, 1); }
    if (tid == 0 && blockIdx.x < 148) {
      long long* o = g_bwd_phase + blockIdx.x * 16;
      o[0] = ph0; o[1] = ph1; o[2] = ph2; o[3] = ph3; o[4] = ph4; o[5] = ph5; o[6] = ph6; o[7] = ph7; o[8] = PH_NOW() - t_kernel0; o[9] = ntl;
      o[10] = hq0; o[11] = hq1; o[12] = hq2; o[13] = hq3; o[14] = hq4; o[15] = hq5;
    }
  }
  // ---- teardown ----------------------------------------------------------------------------------------------------------
  tc_fence_before();
  __syncthreads();
  if (wid == kTcEpiWarps + 1) tmem_dealloc(tbase, TM_COLS);
}

int debug_read_bwd_phases(long long* host, int n) {
  if (n > 148 * 16) n = 148 * 16;
  HEX_CUDA_CHECK(cudaMemcpyFromSymbol(host, g_bwd_phase, sizeof(long long) * n));
  return HEXGNN_OK;
}

template <int NJ>
static int launch_bwd_tc_t(const BwdTcArgs& a, int grid, cudaStream_t st) {
  using S = BwdTcSmem<NJ>;
  static_assert(S::bytes <= 227 * 1024, "chain kernel shared memory");
  auto kern = sage_bwd_tc_kernel<NJ>;
  HEX_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::bytes));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kTcThreads);
  cfg.dynamicSmemBytes = S::bytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;   // train_step: may start under the forward kernel
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = a.td.fwd_flag ? 1 : 0;
  HEX_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, a));
  return HEXGNN_OK;
}

int launch_backward_tc_chain(const Dims& d, const ParamMap& pm, const float* params, const float* x, int64_t N, int64_t E,
                             int64_t B, const void* csr, int flags, const float* d_nodes, const float* d_value,
                             const float* stash, float* dz, float* d_embeds, TcPackView pv, float* small_rec,
                             float* head_rec, int* sync, int grid, cudaStream_t st, const TdFuse* fuse, float* d_x) {
  BwdTcArgs a;
  a.d = d; a.pm = pm; grad_layout(d, &a.gl);
  a.params = params; a.x = x; a.N = N; a.E = E; a.B = B;
  a.csr = csr_view(const_cast<void*>(csr), N, E, B);
  a.flags = flags; a.d_nodes = d_nodes; a.d_value = d_value; a.stash = stash; a.dz = dz; a.d_embeds = d_embeds;
  a.d_x = d_x;
  a.pv = pv; a.small_rec = small_rec; a.head_rec = head_rec; a.tile_flag = sync + kSyncTileOff;
  if (fuse) a.td = *fuse; else a.td = TdFuse{nullptr, nullptr, nullptr, nullptr, nullptr, 0, 0.f};
  switch (d.NJ) {
    case 1: return launch_bwd_tc_t<1>(a, grid, st);
    case 2: return launch_bwd_tc_t<2>(a, grid, st);
    case 3: return launch_bwd_tc_t<3>(a, grid, st);
    case 4: return launch_bwd_tc_t<4>(a, grid, st);
    case 5: return launch_bwd_tc_t<5>(a, grid, st);
    case 6: return launch_bwd_tc_t<6>(a, grid, st);
    case 7: return launch_bwd_tc_t<7>(a, grid, st);
    default: break;
  }
  set_error("backward(tc): unsupported hidden=%d (padded %d)", d.H, d.Hp);
  return HEXGNN_E_UNSUPPORTED;
}

}  // namespace hexgnn
