// Fused forward of modern_two_headed on the 5th-gen tensor cores (tcgen05 + TMEM), see sage_tc.cuh.
//
// One persistent CTA per SM, one 128-row tile (whole graphs) at a time, three warp roles:
//   warps 0-15 epilogue.  Phases 1/3: thread = (node row, column slice), TMEM <-> registers.
//              Phase 2 (CSR gather): warp = node row, lane = 4 channels -- every shared-memory access is a
//              contiguous row (no bank conflicts, no degree divergence inside a warp).
//   warp  16   producer: streams the 3xTF32 weight images of every layer through a shared-memory ring
//              with 1-D bulk copies (UBLKCP) completing on mbarriers; runs ahead across layers/tiles
//   warp  17   MMA issuer: one thread issues tcgen05.mma (A = h in TMEM, B = ring stage, D = [Zl | Zr] in TMEM)
// Layer 0 (2 -> H) is a rank-2 update on the CUDA cores; the dueling head (GN0/models.py:368-384,
// 567-584) runs in the same CTA on the final activations.
#include "sage_tc.cuh"

namespace hexgnn {
using namespace tc;

struct FwdTcArgs {
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
  TcPackView pv;
  int* tile_flag;   // hexgnn_train_step: released per finished tile (the backward chain starts under the last round)
};

template <int NJ>
struct FwdTcSmem {
  static constexpr int Hp = 16 * NJ;
  static constexpr int P = Hp + 4;                       // row pitch (floats): float4 rows conflict-free
  static constexpr int ECAP = kEdgeCapPerRow * kTcRows;
  static constexpr int stages = NJ <= 7 ? 10 : 6;        // ring depth
  static constexpr int N2 = 32 * NJ;
  static constexpr int half_bytes = kTcKF * Hp * 2;      // one part (fp16 hi | fp16 2^11 lo) of a (lin_l | lin_r, 16-k chunk) stage
  static constexpr int stage_bytes = 2 * half_bytes;
  static constexpr int off_ring = 0;
  static constexpr int off_dump = off_ring + stages * stage_bytes;               // float [128][P]  Zl rows
  static constexpr int off_agg = off_dump + (kTcRows + 1) * P * 4;               // float [128][P]  sum_j Zl_j
                                                                                 // (dump row 128 = zeros)
  static constexpr int off_rows = off_agg + kTcRows * P * 4;                     // float [6][128]: inv,x0,x1,m0,m1,adv
  static constexpr int off_pool = off_rows + 6 * kTcRows * 4;                    // float [4*Hp]
  static constexpr int off_hid = off_pool + 4 * Hp * 4;                          // float [Hp]
  static constexpr int off_misc = off_hid + Hp * 4;                              // float [16]
  static constexpr int off_rowptr = off_misc + 64;                               // int [128+4]
  static constexpr int off_gp = off_rowptr + (kTcRows + 4) * 4;                  // int [64+4]
  static constexpr int off_col = off_gp + (kMaxTileGraphs + 4) * 4;              // u16 [ECAP]
  static constexpr int off_gtab = (off_col + ECAP * 2 + 15) & ~15;               // gather table (sage_tc.cuh)
  static constexpr int off_bar = off_gtab + gather_tab_bytes();                  // u64 [2*stages + 8], u32 tmem slot
  static constexpr int off_bias = off_bar + (2 * stages + 8) * 8 + 16;           // float [Lt][Hp] (runtime sized)
  static size_t bytes(int Lt) { return (size_t)off_bias + (size_t)Lt * Hp * 4; }
};

// Phase profile (cycles, summed over the tiles of each CTA; read back by hexgnn_debug_phases):
//   [0] epilogue: wait for Zl  [1] phase 1 (Zl dump)   [2] phase 2 (gather)   [3] tile load + layer 0
//   [4] head                   [5] MMA warp: wait for h   [6] MMA warp: wait for weights   [7] MMA warp: issue + commit
//   [8] total kernel cycles of the CTA   [9] tiles   [10] wait for Zr   [11] phase 3 (combine + write-back)
//   [12] phase 3 up to the Zr TMEM load   [13] phase 3 up to the last store issued (cumulative inside [11])
__device__ long long g_fwd_phase[148 * 16];

struct TileDims { int g0, g1, r0, n, e0, ne; bool ok; };
__device__ __forceinline__ TileDims tile_dims(const CsrView& v, int t, int ecap) {
  TileDims d;
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

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(kTcEpiThreads) : "memory"); }

template <int NJ>
__global__ void __launch_bounds__(kTcThreads, 1) sage_fwd_tc_kernel(const __grid_constant__ FwdTcArgs a) {
  using S = FwdTcSmem<NJ>;
  constexpr int Hp = S::Hp, P = S::P, kStages = S::stages;
  constexpr int NG = 2 * NJ;                 // 8-column granules per row
  constexpr int NQ = Hp / kTcKF;             // k chunks (ring stages) per layer
  constexpr int MAXG = (NG + kTcSlices - 1) / kTcSlices;   // granules per epilogue thread
  // TMEM columns: A (fp16 pairs; per 16-k chunk 8 columns of 2^11 lo, then 8 columns of hi) | D | Ds -- the all-fp16 3-product
  // scheme of sage_tc.cuh.  A layer is TWO N = Hp GEMMs on the same A operand: Zl = h W_l^T first, and Zr = h W_r^T while the
  // epilogue gathers Zl; D takes the large products, Ds the (2^11-scaled) small ones, Z = D + 2^-11 Ds in the epilogue.
  constexpr uint32_t TM_A = 0, TM_D = Hp, TM_DS = 2 * Hp;
  constexpr float kDsScale = 1.0f / 2048.0f;
  constexpr uint32_t TM_COLS = (3 * Hp <= 64) ? 64 : (3 * Hp <= 128) ? 128 : (3 * Hp <= 256) ? 256 : 512;
  extern __shared__ __align__(128) uint8_t smem[];
  float* dump = (float*)(smem + S::off_dump);
  float* agg = (float*)(smem + S::off_agg);
  float* rows = (float*)(smem + S::off_rows);
  float* inv_s = rows, *x0_s = rows + 128, *x1_s = rows + 256, *m0_s = rows + 384, *m1_s = rows + 512, *adv_s = rows + 640;
  float* pool = (float*)(smem + S::off_pool);
  float* hid = (float*)(smem + S::off_hid);
  float* misc = (float*)(smem + S::off_misc);
  int* rowptr_s = (int*)(smem + S::off_rowptr);
  int* gp_s = (int*)(smem + S::off_gp);
  unsigned short* col_s = (unsigned short*)(smem + S::off_col);
  const GatherTab gtab = gather_tab_view(smem + S::off_gtab);
  uint64_t* bar_full = (uint64_t*)(smem + S::off_bar);
  uint64_t* bar_empty = bar_full + kStages;
  uint64_t* bar_z = bar_empty + kStages;     // MMA -> epilogue: Zl complete in D / Ds
  uint64_t* bar_r = bar_z + 1;               // MMA -> epilogue: Zr complete in D / Ds
  uint64_t* bar_df = bar_r + 1;              // epilogue -> MMA: Zl has been drained, the Zr GEMM may overwrite D / Ds
  uint64_t* bar_a = bar_df + 1;              // [4]: column quarter gi of the next A operand is in TMEM
  uint32_t* tmem_slot = (uint32_t*)(bar_a + 4);
  float* bias_s = (float*)(smem + S::off_bias);

  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int H = a.d.H, Lt = a.d.Lt;
  const int ntiles = *a.csr.ntiles;
  const long long t_kernel0 = PH_NOW();
  long long ph0 = 0, ph1 = 0, ph2 = 0, ph3 = 0, ph4 = 0, ph5 = 0, ph6 = 0, ph7 = 0, ph10 = 0, ph11 = 0, ph12 = 0, ph13 = 0, ntl = 0;

  if (tid == 0) {
    for (int i = 0; i < kStages; ++i) { mbar_init(&bar_full[i], 1); mbar_init(&bar_empty[i], 1); }
    mbar_init(bar_z, 1);
    mbar_init(bar_r, 1);
    mbar_init(bar_df, kTcEpiWarps);
    for (int i = 0; i < 4; ++i) mbar_init(&bar_a[i], kTcEpiWarps);
    fence_mbar_init();
  }
  if (wid == kTcEpiWarps + 1) tmem_alloc(tmem_slot, TM_COLS);
  for (int i = tid; i < Lt * Hp; i += kTcThreads) bias_s[i] = a.pv.bias[i];
  for (int i = tid; i < P; i += kTcThreads) dump[kTcRows * P + i] = 0.f;   // the gather's zero row
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = *tmem_slot;
  pdl_launch_dependents();   // hexgnn_train_step: the chain kernel's CTAs may take the SMs this grid's last round leaves idle

  if (wid == kTcEpiWarps) {
    // ================= producer: weight images -> ring =================
    if (lane == 0) {
      uint32_t it = 0;
      for (int rnd = 0; rnd * (int)gridDim.x < ntiles; ++rnd) {
      const int ti = tile_position(rnd, blockIdx.x, gridDim.x, a.tile_flag == nullptr);
      if (ti >= ntiles) continue;
      const int t = a.csr.tile_order[ti];
        if (!tile_dims(a.csr, t, S::ECAP).ok) continue;
        for (int s = 1; s < Lt; ++s) {
          const uint8_t* src = (const uint8_t*)a.pv.wimg + (size_t)(s - 1) * 2 * NQ * S::stage_bytes;
          for (int q = 0; q < 2 * NQ; ++q, ++it) {   // the lin_l chunks, then the lin_r chunks
            const int st = it % kStages;
            mbar_wait(&bar_empty[st], ((it / kStages) & 1) ^ 1);
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
      const bool x1 = (a.flags & HEXGNN_TF32X1) != 0;
      uint32_t it = 0, na = 0;
      for (int rnd = 0; rnd * (int)gridDim.x < ntiles; ++rnd) {
      const int ti = tile_position(rnd, blockIdx.x, gridDim.x, a.tile_flag == nullptr);
      if (ti >= ntiles) continue;
      const int t = a.csr.tile_order[ti];
        if (!tile_dims(a.csr, t, S::ECAP).ok) continue;
        for (int s = 1; s < Lt; ++s) {
          long long tq;
          for (int which = 0; which < 2; ++which) {        // Zl = h W_l^T, then Zr = h W_r^T (same A operand)
            for (int q = 0; q < NQ; ++q, ++it) {           // 16-k chunks = granules 2q, 2q + 1 of h
              if (which == 0 && (q % 2) == 0) {
                // granules [2q, 2q + 4) of h are in TMEM; the epilogue releases h quarter by quarter (and has read Zr of the
                // previous layer before the first quarter), so these MMAs run under the rest of its phase 3
                tq = PH_NOW();
                mbar_wait(&bar_a[q / 2], na & 1);
                PH_ADD(ph5, tq);
                tc_fence_after();
              }
              if (which == 1 && q == 0) {                    // Zl has been drained from D / Ds
                tq = PH_NOW();
                mbar_wait(bar_df, na & 1);
                PH_ADD(ph5, tq);
                tc_fence_after();
              }
              const int st = it % kStages;
              tq = PH_NOW();
              mbar_wait(&bar_full[st], (it / kStages) & 1);
              PH_ADD(ph6, tq);
              tq = PH_NOW();
              tc_fence_after();
              const uint32_t b0 = smem_u32(smem + S::off_ring + st * S::stage_bytes);
              const uint64_t bhi = make_smem_desc(b0, lbo, sbo), blo = make_smem_desc(b0 + S::half_bytes, lbo, sbo);
              const uint32_t alo = tbase + TM_A + q * 16, ahi = alo + 8;
              if (!x1) {   // (HEXGNN_TF32X1, opt-in inference: the large product alone)
                mma_f16_ts(tbase + TM_DS, alo, bhi, idesc16, q != 0);
                mma_f16_ts(tbase + TM_DS, ahi, blo, idesc16, true);
              }
              mma_f16_ts(tbase + TM_D, ahi, bhi, idesc16, q != 0);
              mma_commit(&bar_empty[st]);   // stage free once these MMAs have read it
              PH_ADD(ph7, tq);
            }
            mma_commit(which == 0 ? bar_z : bar_r);
          }
          ++na;
        }
      }
      if (blockIdx.x < 148) {
        long long* o = g_fwd_phase + blockIdx.x * 16;
        o[5] = ph5; o[6] = ph6; o[7] = ph7;
      }
    }
  } else {
    // ================= epilogue warps =================
    const int q4 = wid & 3, slice = wid >> 2;           // TMEM lane quarter, column slice
    const int row = q4 * 32 + lane;
    const uint32_t trow = tbase + ((uint32_t)(q4 * 32) << 16);
    const float flag = a.x[2];
    const int head = (flag == 1.0f) ? 0 : 1;
    const float* Pm = a.params;
    const bool x1e = (a.flags & HEXGNN_TF32X1) != 0;   // single-pass inference: Ds is not written
    int range_bad = 0;
    // one 8-channel granule of the next A operand: fp16 hi and 2^11-scaled lo halves, 4 packed columns each, into the
    // granule's half of its 16-k chunk ([lo(2q) lo(2q+1) | hi(2q) hi(2q+1)])
    auto store_A = [&](int g, const float (&v)[8]) {
      uint32_t hi4[4], lo4[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        split_f16x2(v[2 * j], v[2 * j + 1], hi4[j], lo4[j]);
        if (v[2 * j] >= 60000.f || v[2 * j + 1] >= 60000.f) range_bad = 1;     // (activations are ReLU outputs: >= 0)
      }
      const uint32_t c = trow + TM_A + (g >> 1) * 16 + (g & 1) * 4;
      tmem_st4u(c, lo4);
      tmem_st4u(c + 8, hi4);
    };
    uint32_t nd = 0;
    for (int rnd = 0; rnd * (int)gridDim.x < ntiles; ++rnd) {
      const int ti = tile_position(rnd, blockIdx.x, gridDim.x, a.tile_flag == nullptr);
      if (ti >= ntiles) continue;
      const int t = a.csr.tile_order[ti];
      const TileDims td = tile_dims(a.csr, t, S::ECAP);
      const int n = td.n, r0 = td.r0, g0 = td.g0, g1 = td.g1;
      epi_bar();   // previous tile fully consumed
      long long tq = PH_NOW();
      ++ntl;
      if (!td.ok) {
        const int r1 = a.csr.gptr[g1];
        const float qnan = __int_as_float(0x7fc00000);
        for (int i = r0 + tid; i < r1; i += kTcEpiThreads) a.out_nodes[i] = qnan;
        if (a.out_value) for (int g = g0 + tid; g < g1; g += kTcEpiThreads) a.out_value[g] = qnan;
        if (tid == 0) atomicOr(a.csr.status, HEXGNN_ST_GRAPH_TOO_LARGE);
        if (a.tile_flag) { epi_bar(); if (tid == 0) { __threadfence(); flag_release(a.tile_flag + t, 1); } }
        continue;
      }
      // ---- tile CSR + features -> shared ------------------------------------------------------------
      {
        int bad = 0;
        if (tid == 0) ((int*)misc)[8] = 0;      // "an activation left the fp16 range" flag of this tile
        for (int i = tid; i <= kTcRows; i += kTcEpiThreads) rowptr_s[i] = (i <= n ? a.csr.rowptr[r0 + i] : a.csr.rowptr[r0 + n]) - td.e0;
        for (int g = tid; g <= g1 - g0; g += kTcEpiThreads) gp_s[g] = a.csr.gptr[g0 + g] - r0;
        for (int e = tid; e < td.ne; e += kTcEpiThreads) {
          int c = a.csr.col[td.e0 + e] - r0;
          if (c < 0 || c >= n) { bad |= HEXGNN_ST_EDGE_CROSSES; c = 0; }
          col_s[e] = (unsigned short)c;
        }
        for (int i = tid; i < kTcRows; i += kTcEpiThreads) {
          float v0 = 0.f, v1 = 0.f;
          if (i < n) {
            const float* xr = a.x + (int64_t)(r0 + i) * 3;
            v0 = xr[0]; v1 = xr[1];
            if (xr[2] != flag) bad |= HEXGNN_ST_FLAG_MIXED;
          }
          x0_s[i] = v0; x1_s[i] = v1;
        }
        if (bad) atomicOr(a.csr.status, bad);
        epi_bar();
        gather_build<NJ>(gtab, rowptr_s, col_s, n, tid);
        for (int i = tid; i < kTcRows; i += kTcEpiThreads) {
          const int b = rowptr_s[i], e = rowptr_s[i + 1];
          const int deg = e - b;
          const float inv = 1.0f / (float)(deg > 0 ? deg : 1);
          float s0 = 0.f, s1 = 0.f;
          for (int p = b; p < e; ++p) { const int j = col_s[p]; s0 += x0_s[j]; s1 += x1_s[j]; }
          inv_s[i] = inv; m0_s[i] = s0 * inv; m1_s[i] = s1 * inv;
        }
        epi_bar();
      }
      const bool live = row < n;
      const int n32 = (n + 31) & ~31;
      // training stash: RM32 image of this (layer, tile), see sage_tc.cuh
      const size_t img_bytes = (size_t)tc_act_blocks(Hp) * kImgBlockBytes;
      char* st_img = a.stash ? (char*)a.stash + ((size_t)0 * a.B + t) * img_bytes : nullptr;
      float* emb = a.stash ? a.stash + (size_t)Lt * a.B * (img_bytes / 4) : nullptr;
      // sign masks: word (layer, tile, slice, row), see sage_tc.cuh
      uint32_t* mask_t = a.stash ? (uint32_t*)(a.stash + tc_stash_mask_off_floats(a.d, a.N, a.B)) +
                                       ((size_t)t * kTcSlices + slice) * kTcRows + row : nullptr;
      const size_t mask_layer = (size_t)a.B * kTcSlices * kTcRows;
      const float inv = inv_s[row];
      // ---- layer 0 (2 -> H): rank-2 update, result -> TMEM as the first A operand --------------------------
      {
        const float* wl = Pm + sage_wl(a.pm, head, 0);
        const float* bl = Pm + sage_bl(a.pm, head, 0, H);
        const float* wr = Pm + sage_wr(a.pm, head, 0, H);
        const float xm0 = m0_s[row], xm1 = m1_s[row], xx0 = x0_s[row], xx1 = x1_s[row];
        uint32_t mw = 0;
        for (int g = slice, gi = 0; g < NG; g += kTcSlices, ++gi) {
          float v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int c = g * 8 + u;
            float r = 0.f;
            if (live && c < H)
              r = fmaxf(wl[2 * c] * xm0 + wl[2 * c + 1] * xm1 + bl[c] + wr[2 * c] * xx0 + wr[2 * c + 1] * xx1, 0.f);
            v[u] = r;
            if (r > 0.f) mw |= 1u << (gi * 8 + u);
          }
          store_A(g, v);
          if (st_img && row < n32) {
            st_global_v8(st_img + rm32_granule_off(g, row), v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7]);
          }
          if (emb && a.d.L == 1 && live) {
            float4* dst = (float4*)(emb + (int64_t)(r0 + row) * Hp + g * 8);
            dst[0] = make_float4(v[0], v[1], v[2], v[3]);
            dst[1] = make_float4(v[4], v[5], v[6], v[7]);
          }
        }
        if (st_img && slice == 0 && row < n32) {   // ones channel (bias gradient row of h^T dZ)
          st_global_v8(st_img + rm32_granule_off(NG, row), live ? 1.f : 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f);
        }
        if (mask_t) mask_t[0] = mw;
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
#pragma unroll
          for (int gi = 0; gi < MAXG; ++gi) mbar_arrive(&bar_a[gi]);
        }
      }
      PH_ADD(ph3, tq);
      // ---- layers 1 .. Lt-1 on the tensor cores ---------------------------------------------------------------
      for (int s = 1; s < Lt; ++s) {
        tq = PH_NOW();
        mbar_wait(bar_z, nd & 1);
        ++nd;
        PH_ADD(ph0, tq);
        tq = PH_NOW();
        tc_fence_after();
        // phase 1: Zl -> shared (row major); all TMEM loads in flight before the single wait
        {
          float v[MAXG][8];
#pragma unroll
          for (int gi = 0; gi < MAXG; ++gi) {
            const int g = slice + gi * kTcSlices;
            if (g < NG) tmem_ld8(trow + TM_D + g * 8, v[gi]);
          }
          tmem_wait_ld();
          if (!x1e) {
#pragma unroll
            for (int gi = 0; gi < MAXG; gi += 2) {       // + 2^-11 x the small-product accumulator, two granules per round trip
              float d0[8], d1[8];
              const int g0_ = slice + gi * kTcSlices, g1_ = slice + (gi + 1) * kTcSlices;
              if (g0_ < NG) tmem_ld8(trow + TM_DS + g0_ * 8, d0);
              if (gi + 1 < MAXG && g1_ < NG) tmem_ld8(trow + TM_DS + g1_ * 8, d1);
              tmem_wait_ld();
#pragma unroll
              for (int u = 0; u < 8; ++u) {
                if (g0_ < NG) v[gi][u] = fmaf(d0[u], kDsScale, v[gi][u]);
                if (gi + 1 < MAXG && g1_ < NG) v[gi + 1][u] = fmaf(d1[u], kDsScale, v[gi + 1][u]);
              }
            }
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_df);          // Zl is in registers: the Zr GEMM may start (it runs under phase 2)
#pragma unroll
          for (int gi = 0; gi < MAXG; ++gi) {
            const int g = slice + gi * kTcSlices;
            if (g < NG) {
              float4* dst = (float4*)(dump + row * P + g * 8);
              dst[0] = make_float4(v[gi][0], v[gi][1], v[gi][2], v[gi][3]);
              dst[1] = make_float4(v[gi][4], v[gi][5], v[gi][6], v[gi][7]);
            }
          }
        }
        epi_bar();
        PH_ADD(ph1, tq);
        tq = PH_NOW();
        // phase 2: agg_i = sum_{j -> i} Zl_j in CSR order (tile_gather, sage_tc.cuh)
        tile_gather<NJ>(gtab, dump, agg, n, wid, lane);
        epi_bar();
        PH_ADD(ph2, tq);
        tq = PH_NOW();
        // phase 3: h' = relu(agg_i / deg_i + Zr_i + b) -> TMEM (next A operand), stash image, sign mask
        const bool last = (s == Lt - 1);
        const float* bias = bias_s + s * Hp;
        char* st = (a.stash && row < n32) ? st_img + (size_t)s * a.B * img_bytes : nullptr;
        {
          float zr[MAXG][8];
          mbar_wait(bar_r, (nd - 1) & 1);             // Zr complete (its GEMM ran under the gather)
          tc_fence_after();
          PH_ADD(ph10, tq);
#pragma unroll
          for (int gi = 0; gi < MAXG; ++gi) {
            const int g = slice + gi * kTcSlices;
            if (g < NG) tmem_ld8(trow + TM_D + g * 8, zr[gi]);
          }
          tmem_wait_ld();
          if (!x1e) {
#pragma unroll
            for (int gi = 0; gi < MAXG; gi += 2) {
              float d0[8], d1[8];
              const int g0_ = slice + gi * kTcSlices, g1_ = slice + (gi + 1) * kTcSlices;
              if (g0_ < NG) tmem_ld8(trow + TM_DS + g0_ * 8, d0);
              if (gi + 1 < MAXG && g1_ < NG) tmem_ld8(trow + TM_DS + g1_ * 8, d1);
              tmem_wait_ld();
#pragma unroll
              for (int u = 0; u < 8; ++u) {
                if (g0_ < NG) zr[gi][u] = fmaf(d0[u], kDsScale, zr[gi][u]);
                if (gi + 1 < MAXG && g1_ < NG) zr[gi + 1][u] = fmaf(d1[u], kDsScale, zr[gi + 1][u]);
              }
            }
          }
          PH_ADD(ph12, tq);
          uint32_t mw = 0;
#pragma unroll
          for (int gi = 0; gi < MAXG; ++gi) {
            const int g = slice + gi * kTcSlices;
            if (g < NG) {
              float acc[8];
              if (live) {
                const float4* src = (const float4*)(agg + row * P + g * 8);
                const float4 v0 = src[0], v1 = src[1];
                acc[0] = v0.x; acc[1] = v0.y; acc[2] = v0.z; acc[3] = v0.w;
                acc[4] = v1.x; acc[5] = v1.y; acc[6] = v1.z; acc[7] = v1.w;
              }
#pragma unroll
              for (int u = 0; u < 8; ++u) {
                const float r = live ? fmaxf(fmaf(acc[u], inv, zr[gi][u] + bias[g * 8 + u]), 0.f) : 0.f;
                acc[u] = r;
                if (r > 0.f) mw |= 1u << (gi * 8 + u);
              }
              if (st) {
                st_global_v8(st + rm32_granule_off(g, row), acc[0], acc[1], acc[2], acc[3], acc[4], acc[5], acc[6], acc[7]);
              }
              if (emb && s == a.d.L - 1 && live) {
                float4* dst = (float4*)(emb + (int64_t)(r0 + row) * Hp + g * 8);
                dst[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
                dst[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
              }
              if (!last) {
                store_A(g, acc);
              } else {
                tmem_st8(trow + TM_D + g * 8, acc);   // park the exact fp32 activations in the drained Zl columns
              }
            }
            if (!last) {   // release column quarter gi of the next A operand to the MMA warp (one arrival per warp)
              tmem_wait_st();
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(&bar_a[gi]);
            }
          }
          if (mask_t) mask_t[(size_t)s * mask_layer] = mw;
        }
        if (st && slice == 0) {
          st_global_v8(st + rm32_granule_off(NG, row), live ? 1.f : 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f);
        }
        PH_ADD(ph13, tq);
        tmem_wait_st();
        PH_ADD(ph11, tq);
      }
      // an activation outside fp16's range: the products of this tile were not exact -- flag it and poison its outputs (NaN,
      // like a tile that does not fit), so that neither an arg-max nor an optimizer step is taken from them
      if (range_bad) { atomicOr(a.csr.status, HEXGNN_ST_FP16_RANGE); atomicOr((int*)misc + 8, 1); range_bad = 0; }
      tq = PH_NOW();
      // ---- final activations -> shared (row major) for the head ----------------------------------------------------
      epi_bar();   // every gather of the last layer has finished reading the dump
      const bool poisoned = ((volatile int*)misc)[8] != 0;
      for (int g = slice; g < NG; g += kTcSlices) {
        float v[8];
        tmem_ld8(trow + TM_D + g * 8, v);
        tmem_wait_ld();
        float4* dst = (float4*)(dump + row * P + g * 8);
        dst[0] = make_float4(v[0], v[1], v[2], v[3]);
        dst[1] = make_float4(v[4], v[5], v[6], v[7]);
      }
      tc_fence_before();
      epi_bar();
      // ---- advantage head: a_i = 2 tanh(w_a . g_i + b_a)  (models.py:374, 567-571) -----------------------------------
      {
        const float* wa = Pm + a.pm.lin_w[head];
        const float ba = Pm[a.pm.lin_b[head]];
        // thread = (row, quarter of the channels); inside a quarter the channel order is rotated by the lane so that the
        // 32 rows of a warp hit 32 different banks (row pitch Hp + 4 words); quarters combined in fixed order
        constexpr int CQ = Hp / 4;
        {
          const int r = tid & (kTcRows - 1), part = tid >> 7;
          float u = 0.f;
          if (r < n) {
            const float* gr = dump + r * P + part * CQ;
            const float* wq = wa + part * CQ;
            for (int j = 0; j < CQ; ++j) {
              const int c = (j + lane) % CQ;
              if (part * CQ + c < H) u = fmaf(wq[c], gr[c], u);
            }
          }
          agg[part * kTcRows + r] = u;           // agg is free here: partial sums [4][128]
        }
        epi_bar();
        if (tid < kTcRows) {
          float th = 0.f;
          if (tid < n) th = tanhf(((agg[tid] + agg[kTcRows + tid]) + (agg[2 * kTcRows + tid] + agg[3 * kTcRows + tid])) + ba);
          adv_s[tid] = 2.0f * th;
          if (a.stash) a.stash[tc_stash_th_off_floats(a.d, a.N, a.B) + (int64_t)t * kTcRows + tid] = th;   // for the backward
        }
        epi_bar();
      }
      if (a.flags & HEXGNN_ADV_ONLY) {
        for (int i = tid; i < n; i += kTcEpiThreads) a.out_nodes[r0 + i] = poisoned ? __int_as_float(0x7fc00000) : adv_s[i];
        PH_ADD(ph4, tq);
        continue;
      }
      // ---- value head per graph: pools (sum|max|min|mean) -> MLP -> tanh; dueling combine --------------------------------
      const float* w0 = Pm + a.pm.v0_w[head];
      const float* b0 = Pm + a.pm.v0_b[head];
      const float* w1 = Pm + a.pm.v1_w[head];
      const float b1 = Pm[a.pm.v1_b[head]];
      const int H2 = a.d.H2;
      for (int g = 0; g < g1 - g0; ++g) {
        const int rs = gp_s[g], re = gp_s[g + 1];
        const int ng = re - rs;
        if (tid < H) {
          float sum = 0.f, mx = 0.f, mn = 0.f;  // empty group -> 0 (torch_scatter semantics)
          if (ng > 0) { mx = -INFINITY; mn = INFINITY; }
          for (int i = rs; i < re; ++i) {
            const float v = dump[i * P + tid];
            sum += v; mx = fmaxf(mx, v); mn = fminf(mn, v);
          }
          pool[tid] = sum;
          pool[H + tid] = mx;
          pool[2 * H + tid] = mn;
          pool[3 * H + tid] = sum / (float)(ng > 0 ? ng : 1);
        }
        epi_bar();
        {
          // hid[k] = relu(w0[k] . pooled + b0[k]): warp = rows k = wid, wid + 16, ...; two rows per batch with all their
          // weight loads (L2) in flight before the first FMA; per row the same summation order as a plain lane-strided loop
          constexpr int JMAX = (4 * Hp + 31) / 32;
          float pl[JMAX];
#pragma unroll
          for (int j = 0; j < JMAX; ++j) { const int q = lane + 32 * j; pl[j] = q < 4 * H ? pool[q] : 0.f; }
          for (int k0 = wid; k0 < H2; k0 += 2 * kTcEpiWarps) {
            const int k1 = k0 + kTcEpiWarps;
            const float* r0p = w0 + (int64_t)k0 * 4 * H;
            const float* r1p = w0 + (int64_t)(k1 < H2 ? k1 : k0) * 4 * H;
            float wa0[JMAX], wa1[JMAX];
#pragma unroll
            for (int j = 0; j < JMAX; ++j) {
              const int q = lane + 32 * j;
              wa0[j] = q < 4 * H ? r0p[q] : 0.f;
              wa1[j] = q < 4 * H ? r1p[q] : 0.f;
            }
            float u0 = 0.f, u1 = 0.f;
#pragma unroll
            for (int j = 0; j < JMAX; ++j) { u0 = fmaf(wa0[j], pl[j], u0); u1 = fmaf(wa1[j], pl[j], u1); }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
              u0 += __shfl_xor_sync(0xffffffffu, u0, o);
              u1 += __shfl_xor_sync(0xffffffffu, u1, o);
            }
            if (lane == 0) {
              hid[k0] = fmaxf(u0 + b0[k0], 0.f);
              if (k1 < H2) hid[k1] = fmaxf(u1 + b0[k1], 0.f);
            }
          }
        }
        epi_bar();
        if (a.stash && tid < H2)   // kept for the backward (tc_stash_hid_off_floats)
          a.stash[tc_stash_hid_off_floats(a.d, a.N, a.B) + (int64_t)(g0 + g) * ((H2 + 3) & ~3) + tid] = hid[tid];
        if (wid == 0) {
          float u = 0.f;
          for (int k = lane; k < H2; k += 32) u = fmaf(w1[k], hid[k], u);
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) u += __shfl_xor_sync(0xffffffffu, u, o);
          if (lane == 0) misc[0] = tanhf(u + b1);
        } else if (wid == 1) {
          float u = 0.f;
          for (int i = rs + lane; i < re; i += 32) u += adv_s[i];
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) u += __shfl_xor_sync(0xffffffffu, u, o);
          if (lane == 0) misc[1] = u / (float)(ng > 0 ? ng : 1);
        }
        epi_bar();
        const float val = misc[0], amean = misc[1];
        if (a.flags & HEXGNN_SEPARATE) {
          for (int i = rs + tid; i < re; i += kTcEpiThreads) a.out_nodes[r0 + i] = poisoned ? __int_as_float(0x7fc00000) : adv_s[i] - amean;
        } else {
          for (int i = rs + tid; i < re; i += kTcEpiThreads) a.out_nodes[r0 + i] = poisoned ? __int_as_float(0x7fc00000) : val + (adv_s[i] - amean);
        }
        if (tid == 0 && a.out_value) a.out_value[g0 + g] = poisoned ? __int_as_float(0x7fc00000) : val;
        epi_bar();
      }
      PH_ADD(ph4, tq);
      if (a.tile_flag) {                   // Q, stash images and mask words of the tile are written: hand it over
        fence_proxy_async_global();
        epi_bar();
        if (tid == 0) { __threadfence(); flag_release(a.tile_flag + t, 1); }
      }
    }
    if (tid == 0 && blockIdx.x < 148) {
      long long* o = g_fwd_phase + blockIdx.x * 16;
      o[0] = ph0; o[1] = ph1; o[2] = ph2; o[3] = ph3; o[4] = ph4; o[8] = PH_NOW() - t_kernel0; o[9] = ntl;
      o[10] = ph10; o[11] = ph11; o[12] = ph12; o[13] = ph13;
    }
  }
  // ---- teardown ----------------------------------------------------------------------------------------------------------
  tc_fence_before();
  __syncthreads();
  if (wid == kTcEpiWarps + 1) tmem_dealloc(tbase, TM_COLS);
}

template <int NJ>
static int launch_fwd_tc_t(const FwdTcArgs& a, cudaStream_t st) {
  using S = FwdTcSmem<NJ>;
  auto kern = sage_fwd_tc_kernel<NJ>;
  const size_t smem = S::bytes(a.d.Lt);
  if (smem > 227 * 1024) { set_error("forward(tc): %zu bytes of shared memory needed", smem); return HEXGNN_E_UNSUPPORTED; }
  HEX_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int grid = (int)(a.B < 1 ? 1 : (a.B > bwd_grid_ctas() ? bwd_grid_ctas() : a.B));
  {
    ProfScope prof(0, st);
    kern<<<grid, kTcThreads, smem, st>>>(a);
  }
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

int debug_read_fwd_phases(long long* host, int n) {
  if (n > 148 * 16) n = 148 * 16;
  HEX_CUDA_CHECK(cudaMemcpyFromSymbol(host, g_fwd_phase, sizeof(long long) * n));
  return HEXGNN_OK;
}

int launch_forward_tc(const Dims& d, const ParamMap& pm, const float* params, const float* x, int64_t N, int64_t E,
                      int64_t B, const void* csr, int flags, float* out_nodes, float* out_value, float* stash,
                      TcPackView pv, cudaStream_t st, int* tile_flag) {
  FwdTcArgs a;
  a.tile_flag = tile_flag;
  a.d = d; a.pm = pm; a.params = params; a.x = x; a.N = N; a.E = E; a.B = B;
  a.csr = csr_view(const_cast<void*>(csr), N, E, B);
  a.flags = flags; a.out_nodes = out_nodes; a.out_value = out_value; a.stash = stash; a.pv = pv;
  switch (d.NJ) {
    case 1: return launch_fwd_tc_t<1>(a, st);
    case 2: return launch_fwd_tc_t<2>(a, st);
    case 3: return launch_fwd_tc_t<3>(a, st);
    case 4: return launch_fwd_tc_t<4>(a, st);
    case 5: return launch_fwd_tc_t<5>(a, st);
    case 6: return launch_fwd_tc_t<6>(a, st);
    case 7: return launch_fwd_tc_t<7>(a, st);
    case 8: return launch_fwd_tc_t<8>(a, st);
    default: break;
  }
  set_error("forward(tc): unsupported hidden=%d (padded %d)", d.H, d.Hp);
  return HEXGNN_E_UNSUPPORTED;
}

}  // namespace hexgnn
