// Fused forward of modern_two_headed on the 5th-gen tensor cores (tcgen05 + TMEM), see sage_tc.cuh.
//
// One persistent CTA per SM, TWO 128-row tiles (whole graphs) in flight per CTA, three warp roles:
//   warps 0-15 epilogue
//   warp  16   producer: streams the 3xTF32 weight images of every layer through a shared-memory ring
//              with 1-D bulk copies (UBLKCP) completing on mbarriers; runs ahead across layers / tiles
//   warp  17   MMA issuer: one thread issues tcgen05.mma (A = h in TMEM, B = ring stage, D = [Zl | Zr] in TMEM)
//
// Two tiles X, Y of a CTA advance layer by layer in lock step, and the tensor-core GEMM of one tile always runs under
// the CSR gather of the other.  The 512 TMEM columns hold ONE A region R1 (h hi | lo, 2 Hp columns) and ONE D region
// R2 ([Zl | Zr], 2 Hp columns); the tiles time-share them, which works because a tile needs tensor memory only in two
// short phases of its layer step:
//   P1  drain   thread = (row, column slice): R2 -> shared, Zl rows to bufL, Zr + bias rows to bufR      (frees R2)
//   P2  gather + combine (shared memory only): warp = 4 rows, lane = 4 channels; table-driven CSR gather of Zl rows
//       (sage_tc.cuh), h' = relu(agg / deg + Zr + b) written IN PLACE over the Zr row, stash image / sign-mask /
//       embedding stores straight from the registers.  The other tile's 42 MMAs run here.
//   P3  A write thread = (row, column slice): h' rows -> tf32 hi / lo split -> R1                         (needs R1 free)
// Epilogue order per layer s:  [wait MMA(X,s)]  P3(Y)  P1(X)  P2(X) || MMA(Y,s)   [wait MMA(Y,s)]  P3(X)  P1(Y)  P2(Y) || MMA(X,s+1)
// so every resource hand-over is a plain mbarrier: "A in R1" (P3 -> MMA), "R2 drained" (P1 -> MMA), "Z complete"
// (MMA -> P1).  A CTA with an odd number of tiles runs the last one alone (P1, P2, P3 back to back with its own MMA).
// Layer 0 (2 -> H) is a rank-2 update on the CUDA cores; the dueling head (GN0/models.py:368-384, 567-584) runs in the
// same CTA on the final activations, which P2 leaves in shared memory.
#include <stdlib.h>
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
  int stages;       // ring depth (what fits beside the runtime-sized bias table)
  int max_pair;     // 2: two tiles in flight per CTA; 1: one at a time (HEXGNN_FWD_PAIR=1, for A/B measurements)
  int pack_grid;    // 1: only ceil(tiles / max_pair) CTAs take tiles (every CTA gets a full group when tiles < 2 x SMs)
};

constexpr int kFwdMaxStages = 6;

template <int NJ>
struct FwdTcSmem {
  static constexpr int Hp = 16 * NJ;
  static constexpr int P = Hp + 4;                       // row pitch (floats): float4 rows conflict-free
  static constexpr int ECAP = kEdgeCapPerRow * kTcRows;
  static constexpr int N2 = 32 * NJ;
  static constexpr int half_bytes = kTcKF * N2 * 4;      // one tf32 half (hi or lo) of a k-chunk stage
  static constexpr int stage_bytes = 2 * half_bytes;
  static constexpr int off_bufL = 0;                                             // float [128 + 1][P]  Zl rows (row 128 = zeros)
  static constexpr int off_bufR = off_bufL + (kTcRows + 1) * P * 4;              // float [128][P]      Zr + b, then h' in place
  static constexpr int off_rows = off_bufR + kTcRows * P * 4;                    // float [5][128]: x0, x1, m0, m1, adv
  static constexpr int off_pool = off_rows + 5 * kTcRows * 4;                    // float [4*Hp]
  static constexpr int off_hid = off_pool + 4 * Hp * 4;                          // float [Hp]
  static constexpr int off_misc = off_hid + Hp * 4;                              // float [16]
  // per-tile slot (x 2): the tile's CSR and row table (whose row headers carry 1 / deg)
  static constexpr int slot_rowptr = 0;                                          // int [128 + 4]
  static constexpr int slot_gp = slot_rowptr + (kTcRows + 4) * 4;                // int [64 + 4]
  static constexpr int slot_col = slot_gp + (kMaxTileGraphs + 4) * 4;            // u16 [ECAP]
  static constexpr int slot_gtab = (slot_col + ECAP * 2 + 15) & ~15;             // row table (sage_tc.cuh)
  static constexpr int slot_bytes = (slot_gtab + row_tab_bytes() + 15) & ~15;
  static constexpr int off_slot = off_misc + 64;
  static constexpr int off_bar = off_slot + 2 * slot_bytes;                      // u64 [2 * kFwdMaxStages + 8], u32 tmem slot
  static constexpr int off_bias = off_bar + (2 * kFwdMaxStages + 8) * 8 + 16;    // float [Lt][Hp] (runtime sized)
  __host__ __device__ static size_t off_ring(int Lt) { return ((size_t)off_bias + (size_t)Lt * Hp * 4 + 127) & ~(size_t)127; }
  static size_t bytes(int Lt, int stages) { return off_ring(Lt) + (size_t)stages * stage_bytes; }
};

// Phase profile (cycles, summed over the tiles of each CTA; read back by hexgnn_debug_phases):
//   [0] epilogue: wait for Z   [1] P1 (drain)   [2] P2 (gather + combine)   [3] tile load + layer 0
//   [4] head                   [5] MMA warp: wait for A / R2   [6] MMA warp: wait for weights   [7] MMA warp: issue + commit
//   [8] total kernel cycles of the CTA   [9] tiles   [11] P3 (A write)
__device__ long long g_fwd_phase[148 * 16];

struct TileDims { int t, g0, g1, r0, n, e0, ne; bool ok; };
__device__ __forceinline__ TileDims tile_dims(const CsrView& v, int t, int ecap) {
  TileDims d;
  d.t = t;
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
// The CTA's next tile in schedule order (sage_tc.cuh tile_position), valid or not; false when the CTA has run out.
__device__ __forceinline__ bool fetch_tile(const CsrView& v, int ntiles, int G, bool zig, int ecap, int& rnd, TileDims& td) {
  if ((int)blockIdx.x >= G) return false;
  while (rnd * G < ntiles) {
    const int ti = tile_position(rnd, blockIdx.x, G, zig);
    ++rnd;
    if (ti >= ntiles) continue;
    td = tile_dims(v, v.tile_order[ti], ecap);
    return true;
  }
  return false;
}
// Up to `max_pair` VALID tiles; tiles that do not fit are reported through `bad` (the epilogue poisons their outputs).
template <typename BadFn>
__device__ __forceinline__ int fetch_group(const CsrView& v, int ntiles, int G, bool zig, int ecap, int max_pair, int& rnd,
                                           TileDims (&td)[2], BadFn bad) {
  int np = 0;
  while (np < max_pair) {
    TileDims t;
    if (!fetch_tile(v, ntiles, G, zig, ecap, rnd, t)) break;
    if (!t.ok) { bad(t); continue; }
    td[np++] = t;
  }
  return np;
}

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(kTcEpiThreads) : "memory"); }

template <int NJ>
__global__ void __launch_bounds__(kTcThreads, 1) sage_fwd_tc_kernel(const __grid_constant__ FwdTcArgs a) {
  using S = FwdTcSmem<NJ>;
  constexpr int Hp = S::Hp, P = S::P;
  constexpr int NG = 2 * NJ;                 // 8-column granules per row
  constexpr int NQ = Hp / kTcKF;             // k chunks (ring stages) per layer
  constexpr int NV = Hp / 4;                 // float4 per row
  constexpr int MAXG = (NG + kTcSlices - 1) / kTcSlices;   // granules per (row, slice) thread
  constexpr uint32_t TM_HI = 0, TM_LO = Hp, TM_D = 2 * Hp;     // TMEM columns: R1 = [hi | lo], R2 = [Zl | Zr]
  constexpr uint32_t TM_COLS = (4 * Hp <= 64) ? 64 : (4 * Hp <= 128) ? 128 : (4 * Hp <= 256) ? 256 : 512;
  static_assert(NV + 2 <= 32, "lanes NV, NV + 1 write the ones-channel granule");
  extern __shared__ __align__(128) uint8_t smem[];
  float* bufL = (float*)(smem + S::off_bufL);
  float* bufR = (float*)(smem + S::off_bufR);
  float* rows = (float*)(smem + S::off_rows);
  float *x0_s = rows, *x1_s = rows + 128, *m0_s = rows + 256, *m1_s = rows + 384, *adv_s = rows + 512;
  float* pool = (float*)(smem + S::off_pool);
  float* hid = (float*)(smem + S::off_hid);
  float* misc = (float*)(smem + S::off_misc);
  uint64_t* bar_full = (uint64_t*)(smem + S::off_bar);
  uint64_t* bar_empty = bar_full + kFwdMaxStages;
  uint64_t* bar_z = bar_empty + kFwdMaxStages;   // MMA -> epilogue: Z of the current job is complete in R2
  uint64_t* bar_d = bar_z + 1;                   // epilogue -> MMA: R2 has been drained
  uint64_t* bar_a = bar_d + 1;                   // [4] epilogue -> MMA: column quarter gi of the next A operand is in R1
  uint32_t* tmem_slot = (uint32_t*)(bar_a + 4);
  float* bias_s = (float*)(smem + S::off_bias);
  const int H = a.d.H, Lt = a.d.Lt;
  uint8_t* ring = smem + S::off_ring(Lt);
  const int kStages = a.stages;

  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int ntiles = *a.csr.ntiles;
  const bool zig = a.tile_flag == nullptr;
  // CTAs that take tiles: with fewer than max_pair x gridDim tiles only ceil(tiles / max_pair) CTAs work, each on a full
  // group -- a tile that runs alone pays its GEMMs un-overlapped, and the SMs left over go to the next kernel at once
  int G = gridDim.x;
  if (a.pack_grid) { const int want = (ntiles + a.max_pair - 1) / a.max_pair; G = want < G ? (want < 1 ? 1 : want) : G; }
  const long long t_kernel0 = PH_NOW();
  long long ph0 = 0, ph1 = 0, ph2 = 0, ph3 = 0, ph4 = 0, ph5 = 0, ph6 = 0, ph7 = 0, ph11 = 0, ntl = 0;

  if (tid == 0) {
    for (int i = 0; i < kStages; ++i) { mbar_init(&bar_full[i], 1); mbar_init(&bar_empty[i], 1); }
    mbar_init(bar_z, 1);
    mbar_init(bar_d, kTcEpiWarps);
    for (int i = 0; i < 4; ++i) mbar_init(&bar_a[i], kTcEpiWarps);
    fence_mbar_init();
  }
  if (wid == kTcEpiWarps + 1) tmem_alloc(tmem_slot, TM_COLS);
  for (int i = tid; i < Lt * Hp; i += kTcThreads) bias_s[i] = a.pv.bias[i];
  for (int i = tid; i < P; i += kTcThreads) bufL[kTcRows * P + i] = 0.f;   // the gather's zero row
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = *tmem_slot;
  pdl_launch_dependents();   // hexgnn_train_step: the chain kernel's CTAs may take the SMs this grid leaves idle

  if (wid == kTcEpiWarps) {
    // ================= producer: weight images -> ring =================
    if (lane == 0) {
      uint32_t it = 0;
      int rnd = 0;
      TileDims td[2];
      for (;;) {
        const int np = fetch_group(a.csr, ntiles, G, zig, S::ECAP, a.max_pair, rnd, td, [](const TileDims&) {});
        if (np == 0) break;
        for (int s = 1; s < Lt; ++s) {
          const uint8_t* src = (const uint8_t*)a.pv.wimg + (size_t)(s - 1) * NQ * S::stage_bytes;
          for (int k = 0; k < np; ++k) {
            for (int q = 0; q < NQ; ++q, ++it) {
              const int st = it % kStages;
              mbar_wait(&bar_empty[st], ((it / kStages) & 1) ^ 1);
              mbar_expect_tx(&bar_full[st], S::stage_bytes);
              bulk_g2s(ring + st * S::stage_bytes, src + (size_t)q * S::stage_bytes, S::stage_bytes, &bar_full[st]);
            }
          }
        }
      }
    }
  } else if (wid == kTcEpiWarps + 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_tf32(kTcRows, S::N2, false);
      constexpr uint32_t lbo = S::N2 * 16, sbo = 128;
      uint32_t it = 0, job = 0;
      int rnd = 0;
      TileDims td[2];
      for (;;) {
        const int np = fetch_group(a.csr, ntiles, G, zig, S::ECAP, a.max_pair, rnd, td, [](const TileDims&) {});
        if (np == 0) break;
        for (int s = 1; s < Lt; ++s) {
          for (int k = 0; k < np; ++k, ++job) {
            long long tq;
            for (int q = 0; q < NQ; ++q, ++it) {
              if ((q % kTcSlices) == 0) {
                // granules [q, q + kTcSlices) of h (k-steps q .. q+3) are in R1; for q = 0 also: R2 has been drained
                tq = PH_NOW();
                mbar_wait(&bar_a[q / kTcSlices], job & 1);
                if (q == 0) mbar_wait(bar_d, job & 1);
                PH_ADD(ph5, tq);
                tc_fence_after();
              }
              const int st = it % kStages;
              tq = PH_NOW();
              mbar_wait(&bar_full[st], (it / kStages) & 1);
              PH_ADD(ph6, tq);
              tq = PH_NOW();
              tc_fence_after();
              const uint32_t bhi = smem_u32(ring + st * S::stage_bytes);
              const uint64_t dhi = make_smem_desc(bhi, lbo, sbo);
              const uint64_t dlo = make_smem_desc(bhi + S::half_bytes, lbo, sbo);
              const uint32_t kcol = q * kTcKF;
              mma_tf32_ts(tbase + TM_D, tbase + TM_LO + kcol, dhi, idesc, q != 0);
              mma_tf32_ts(tbase + TM_D, tbase + TM_HI + kcol, dlo, idesc, true);
              mma_tf32_ts(tbase + TM_D, tbase + TM_HI + kcol, dhi, idesc, true);
              mma_commit(&bar_empty[st]);   // stage free once these MMAs have read it
              PH_ADD(ph7, tq);
            }
            mma_commit(bar_z);              // Z complete
          }
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
    const size_t img_bytes = (size_t)tc_act_blocks(Hp) * kImgBlockBytes;
    float* emb = a.stash ? a.stash + (size_t)Lt * a.B * (img_bytes / 4) : nullptr;
    uint4* mask_base = a.stash ? (uint4*)(a.stash + tc_stash_mask_off_floats(a.d, a.N, a.B)) : nullptr;
    const size_t mask_layer = (size_t)a.B * kTcRows;     // uint4 per layer
    uint32_t nd = 0;
    if (lane == 0) mbar_arrive(bar_d);                   // R2 starts out free

    auto slot_ptr = [&](int k) { return smem + S::off_slot + k * S::slot_bytes; };

    // --- a finished row of one (tile, layer): lane = channels [4 lane, 4 lane + 4) -> shared (in place over Zr), stash
    //     image, sign mask (4 ballot words per row: bit l of word c = h[4 l + c] > 0), embedding copy.  Lanes NV and NV + 1
    //     carry the granule behind the last channel: (1, 0, ...) = the ones channel of the stash image (the bias-gradient
    //     row of h^T dZ), so ONE 16-byte store per lane writes the whole row.  The per-(tile, layer) addresses are formed
    //     once (RowOut), the per-row part is a handful of integer ops.
    const uint32_t sL = smem_u32(bufL), sR = smem_u32(bufR);
    const int ln = lane < NV ? lane : 0;                   // idle lanes shadow lane 0 in shared-memory reads
    struct RowOut { char* img; uint4* mrow; float* erow; int gx; };
    auto row_out = [&](const TileDims& td, int s) {
      RowOut o;
      o.img = nullptr; o.mrow = nullptr; o.erow = nullptr; o.gx = 0;
      if (a.stash) {
        const int g = lane >> 1;                           // this lane's 8-channel granule (NG: the ones channel)
        o.img = (char*)a.stash + ((size_t)s * a.B + td.t) * img_bytes + (size_t)(g >> 2) * kImgBlockBytes + (lane & 1) * 16;
        o.gx = g & 3;
        o.mrow = mask_base + (size_t)s * mask_layer + (size_t)td.t * kTcRows;
        if (s == a.d.L - 1) o.erow = emb + (int64_t)td.r0 * Hp + ln * 4;
      }
      return o;
    };
    auto finish_row = [&](const RowOut& o, int r, float4 h) {
      if (lane < NV) sts128(sR + (uint32_t)(r * (P * 4) + lane * 16), h);
      if (o.img) {
        if (lane >= NV) h = make_float4(lane == NV ? 1.f : 0.f, 0.f, 0.f, 0.f);
        char* dst = o.img + (r >> 2) * 512 + (r & 3) * 128 + ((o.gx ^ (r & 3)) << 5);      // rm32_granule_off, row part
        if (lane < NV + 2) *(float4*)dst = h;
        const bool in = lane < NV;
        const uint32_t b0 = __ballot_sync(0xffffffffu, in && h.x > 0.f), b1 = __ballot_sync(0xffffffffu, in && h.y > 0.f);
        const uint32_t b2 = __ballot_sync(0xffffffffu, in && h.z > 0.f), b3 = __ballot_sync(0xffffffffu, in && h.w > 0.f);
        if (lane == 0) o.mrow[r] = make_uint4(b0, b1, b2, b3);
        if (o.erow && lane < NV) *(float4*)(o.erow + r * Hp) = h;
      }
    };
    // rows [n, ceil32(n)) of the stash image are zeros (the weight-gradient GEMM contracts whole 32-row chunks)
    auto finish_tail = [&](const RowOut& o, int n) {
      if (!o.img) return;
      const int n32 = (n + 31) & ~31;
      for (int r = n + wid; r < n32; r += kTcEpiWarps) {
        char* dst = o.img + (r >> 2) * 512 + (r & 3) * 128 + ((o.gx ^ (r & 3)) << 5);
        if (lane < NV + 2) *(float4*)dst = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    };

    // --- tile CSR + features -> slot k (+ the shared x / mean staging rows for layer 0) ----------------------------
    auto load_tile = [&](const TileDims& td, int k) {
      uint8_t* sl = slot_ptr(k);
      int* rowptr_s = (int*)(sl + S::slot_rowptr);
      int* gp_s = (int*)(sl + S::slot_gp);
      unsigned short* col_s = (unsigned short*)(sl + S::slot_col);
      const RowTab rtab = row_tab_view(sl + S::slot_gtab);
      const int n = td.n, r0 = td.r0;
      int bad = 0;
      for (int i = tid; i <= kTcRows; i += kTcEpiThreads) rowptr_s[i] = (i <= n ? a.csr.rowptr[r0 + i] : a.csr.rowptr[r0 + n]) - td.e0;
      for (int g = tid; g <= td.g1 - td.g0; g += kTcEpiThreads) gp_s[g] = a.csr.gptr[td.g0 + g] - r0;
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
      row_tab_build<NJ>(rtab, rowptr_s, col_s, n, tid);
      for (int i = tid; i < kTcRows; i += kTcEpiThreads) {
        const int b = rowptr_s[i], e = rowptr_s[i + 1];
        const int deg = e - b;
        const float inv = 1.0f / (float)(deg > 0 ? deg : 1);
        float s0 = 0.f, s1 = 0.f;
        for (int p = b; p < e; ++p) { const int j = col_s[p]; s0 += x0_s[j]; s1 += x1_s[j]; }
        m0_s[i] = s0 * inv; m1_s[i] = s1 * inv;
      }
      epi_bar();
    };

    // --- P3: h' rows (bufR) -> tf32 hi / lo -> R1, released to the MMA warp column quarter by column quarter.  Rows
    //     >= n of bufR are zero by construction (layer 0 clears them, P1 writes Zr = 0 * W without the bias): no select.
    auto write_A = [&]() {
#pragma unroll
      for (int gi = 0; gi < MAXG; ++gi) {
        const int g = slice + gi * kTcSlices;
        if (g < NG) {
          float v[8], hi[8], lo[8];
          const float4 v0 = lds128(sR + (uint32_t)(row * (P * 4) + g * 32)), v1 = lds128(sR + (uint32_t)(row * (P * 4) + g * 32 + 16));
          v[0] = v0.x; v[1] = v0.y; v[2] = v0.z; v[3] = v0.w; v[4] = v1.x; v[5] = v1.y; v[6] = v1.z; v[7] = v1.w;
#pragma unroll
          for (int u = 0; u < 8; ++u) split_tf32(v[u], hi[u], lo[u]);
          tmem_st8(trow + TM_HI + g * 8, hi);
          tmem_st8(trow + TM_LO + g * 8, lo);
        }
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_a[gi]);
      }
    };

    // --- layer 0 (2 -> H): rank-2 update, warp = rows, lane = 4 channels; then P3 --------------------------------
    auto layer0 = [&](const TileDims& td) {
      const float* wl = Pm + sage_wl(a.pm, head, 0);
      const float* bl = Pm + sage_bl(a.pm, head, 0, H);
      const float* wr = Pm + sage_wr(a.pm, head, 0, H);
      float wl0[4], wl1[4], wr0[4], wr1[4], bb[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int c = lane * 4 + j;
        const bool cv = lane < NV && c < H;
        wl0[j] = cv ? wl[2 * c] : 0.f; wl1[j] = cv ? wl[2 * c + 1] : 0.f;
        wr0[j] = cv ? wr[2 * c] : 0.f; wr1[j] = cv ? wr[2 * c + 1] : 0.f;
        bb[j] = cv ? bl[c] : 0.f;
      }
      const RowOut ro = row_out(td, 0);
      const int n = td.n;
      for (int r = wid; r < n; r += kTcEpiWarps) {
        const float xm0 = m0_s[r], xm1 = m1_s[r], xx0 = x0_s[r], xx1 = x1_s[r];
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) o[j] = fmaxf(wl0[j] * xm0 + wl1[j] * xm1 + bb[j] + wr0[j] * xx0 + wr1[j] * xx1, 0.f);
        finish_row(ro, r, make_float4(o[0], o[1], o[2], o[3]));
      }
      finish_tail(ro, n);
      for (int r = n + wid; r < kTcRows; r += kTcEpiWarps)     // dead rows stay zero through every layer (see write_A)
        if (lane < NV) sts128(sR + (uint32_t)(r * (P * 4) + lane * 16), make_float4(0.f, 0.f, 0.f, 0.f));
      epi_bar();
      write_A();
    };

    // --- P1: R2 -> shared (Zl rows -> bufL, Zr rows -> bufR; the bias joins in P2); frees R2 ---------------------------------
    auto drain = [&]() {
      float zl[MAXG][8], zr[MAXG][8];
#pragma unroll
      for (int gi = 0; gi < MAXG; ++gi) {
        const int g = slice + gi * kTcSlices;
        if (g < NG) tmem_ld8(trow + TM_D + g * 8, zl[gi]);
      }
      tmem_wait_ld();
#pragma unroll
      for (int gi = 0; gi < MAXG; ++gi) {
        const int g = slice + gi * kTcSlices;
        if (g < NG) tmem_ld8(trow + TM_D + Hp + g * 8, zr[gi]);
      }
#pragma unroll
      for (int gi = 0; gi < MAXG; ++gi) {
        const int g = slice + gi * kTcSlices;
        if (g < NG) {
          const uint32_t d = sL + (uint32_t)(row * (P * 4) + g * 32);
          sts128(d, make_float4(zl[gi][0], zl[gi][1], zl[gi][2], zl[gi][3]));
          sts128(d + 16, make_float4(zl[gi][4], zl[gi][5], zl[gi][6], zl[gi][7]));
        }
      }
      tmem_wait_ld();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_d);    // every tcgen05.ld of this warp has completed: R2 may be overwritten
#pragma unroll
      for (int gi = 0; gi < MAXG; ++gi) {
        const int g = slice + gi * kTcSlices;
        if (g < NG) {
          const uint32_t d = sR + (uint32_t)(row * (P * 4) + g * 32);
          sts128(d, make_float4(zr[gi][0], zr[gi][1], zr[gi][2], zr[gi][3]));
          sts128(d + 16, make_float4(zr[gi][4], zr[gi][5], zr[gi][6], zr[gi][7]));
        }
      }
    };

    // --- P2: agg_i = sum_{j -> i} Zl_j (row table of sage_tc.cuh: warp = one row, four independent accumulators over the
    //     neighbours k = 0, 1, 2, 3 mod 4, combined (s0 + s1) + (s2 + s3)); h'_i = relu(agg_i / deg_i + (Zr_i + b)).
    //     Software-pipelined over the warp's rows: the header, first-step offsets and first-step loads of the NEXT row are
    //     issued before the current row is finished (the loads are volatile asm: program order = issue order). ---------------
    auto gather_combine = [&](const TileDims& td, int k, int s) {
      uint8_t* sl = slot_ptr(k);
      const uint32_t tabA = smem_u32(sl + S::slot_gtab), hdrA = tabA + kRowTabMaxSteps * 16, firstA = hdrA + kTcRows * 16;
      const RowOut ro = row_out(td, s);
      const int n = td.n;
      const float4 bs = *(const float4*)(bias_s + s * Hp + ln * 4);
      const uint32_t sb = sL + ln * 16, zb = sR + ln * 16;
      int r = wid;
      if (r < n) {
        int4 hd = lds128i(hdrA + r * 16);
        {
          const int4 f = lds128i(firstA + r * 16);
          float4 a0 = lds128(sb + f.x), a1 = lds128(sb + f.y), a2 = lds128(sb + f.z), a3 = lds128(sb + f.w);
          float4 z = lds128(zb + r * (P * 4));
          for (;;) {
            const int rn = r + kTcEpiWarps;
            const int rc = rn < n ? rn : r;                        // last row: harmless re-read of itself
            const int4 hdn = lds128i(hdrA + rc * 16);
            const int4 fn = lds128i(firstA + rc * 16);
            for (int kk = hd.x; kk < hd.y; ++kk) {               // steps beyond the first (degree > 4)
              const int4 p = lds128i(tabA + kk * 16);
              const float4 v0 = lds128(sb + p.x), v1 = lds128(sb + p.y), v2 = lds128(sb + p.z), v3 = lds128(sb + p.w);
              fadd2x2(a0, v0); fadd2x2(a1, v1); fadd2x2(a2, v2); fadd2x2(a3, v3);
            }
            const float4 n0 = lds128(sb + fn.x), n1 = lds128(sb + fn.y), n2 = lds128(sb + fn.z), n3 = lds128(sb + fn.w);
            const float4 zn = lds128(zb + rc * (P * 4));
            fadd2x2(a0, a1); fadd2x2(a2, a3); fadd2x2(a0, a2);
            const float iv = __int_as_float(hd.z);
            z.x += bs.x; z.y += bs.y; z.z += bs.z; z.w += bs.w;
            float4 h;
            h.x = fmaxf(fmaf(a0.x, iv, z.x), 0.f); h.y = fmaxf(fmaf(a0.y, iv, z.y), 0.f);
            h.z = fmaxf(fmaf(a0.z, iv, z.z), 0.f); h.w = fmaxf(fmaf(a0.w, iv, z.w), 0.f);
            finish_row(ro, r, h);
            if (rn >= n) break;
            r = rn; hd = hdn; a0 = n0; a1 = n1; a2 = n2; a3 = n3; z = zn;
          }
        }
      }
      finish_tail(ro, n);
    };

    // --- dueling head on the final activations (rows of bufR); bufL is scratch by now ------------------------------------
    auto run_head = [&](const TileDims& td, int k) {
      uint8_t* sl = slot_ptr(k);
      const int* gp_s = (const int*)(sl + S::slot_gp);
      const int n = td.n, r0 = td.r0, g0 = td.g0, g1 = td.g1, t = td.t;
      float* part = bufL;                                   // [4][128] partial sums
      // advantage head: a_i = 2 tanh(w_a . g_i + b_a)  (models.py:374, 567-571)
      {
        const float* wa = Pm + a.pm.lin_w[head];
        const float ba = Pm[a.pm.lin_b[head]];
        // thread = (row, quarter of the channels); inside a quarter the channel order is rotated by the lane so that the
        // 32 rows of a warp hit 32 different banks (row pitch Hp + 4 words); quarters combined in fixed order
        constexpr int CQ = Hp / 4;
        {
          const int r = tid & (kTcRows - 1), pt = tid >> 7;
          float u = 0.f;
          if (r < n) {
            const float* gr = bufR + r * P + pt * CQ;
            const float* wq = wa + pt * CQ;
            for (int j = 0; j < CQ; ++j) {
              const int c = (j + lane) % CQ;
              if (pt * CQ + c < H) u = fmaf(wq[c], gr[c], u);
            }
          }
          part[pt * kTcRows + r] = u;
        }
        epi_bar();
        if (tid < kTcRows) {
          float th = 0.f;
          if (tid < n) th = tanhf(((part[tid] + part[kTcRows + tid]) + (part[2 * kTcRows + tid] + part[3 * kTcRows + tid])) + ba);
          adv_s[tid] = 2.0f * th;
          if (a.stash) a.stash[tc_stash_th_off_floats(a.d, a.N, a.B) + (int64_t)t * kTcRows + tid] = th;   // for the backward
        }
        epi_bar();
      }
      if (a.flags & HEXGNN_ADV_ONLY) {
        for (int i = tid; i < n; i += kTcEpiThreads) a.out_nodes[r0 + i] = adv_s[i];
        epi_bar();
        return;
      }
      // value head per graph: pools (sum|max|min|mean) -> MLP -> tanh; dueling combine
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
            const float v = bufR[i * P + tid];
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
          for (int k2 = lane; k2 < H2; k2 += 32) u = fmaf(w1[k2], hid[k2], u);
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
          for (int i = rs + tid; i < re; i += kTcEpiThreads) a.out_nodes[r0 + i] = adv_s[i] - amean;
        } else {
          for (int i = rs + tid; i < re; i += kTcEpiThreads) a.out_nodes[r0 + i] = val + (adv_s[i] - amean);
        }
        if (tid == 0 && a.out_value) a.out_value[g0 + g] = val;
        epi_bar();
      }
    };
    auto release_tile = [&](const TileDims& td) {
      if (!a.tile_flag) return;               // Q, stash images and mask words of the tile are written: hand it over
      fence_proxy_async_global();
      epi_bar();
      if (tid == 0) { __threadfence(); flag_release(a.tile_flag + td.t, 1); }
    };
    auto poison = [&](const TileDims& td) {
      const int r1 = a.csr.gptr[td.g1];
      const float qnan = __int_as_float(0x7fc00000);
      for (int i = td.r0 + tid; i < r1; i += kTcEpiThreads) a.out_nodes[i] = qnan;
      if (a.out_value) for (int g = td.g0 + tid; g < td.g1; g += kTcEpiThreads) a.out_value[g] = qnan;
      if (tid == 0) atomicOr(a.csr.status, HEXGNN_ST_GRAPH_TOO_LARGE);
      if (a.tile_flag) { epi_bar(); if (tid == 0) { __threadfence(); flag_release(a.tile_flag + td.t, 1); } }
    };

    int rnd = 0;
    TileDims td[2];
    for (;;) {
      const int np = fetch_group(a.csr, ntiles, G, zig, S::ECAP, a.max_pair, rnd, td, poison);
      if (np == 0) break;
      ntl += np;
      long long tq = PH_NOW();
      epi_bar();                              // previous group fully consumed (head reads of bufR / scratch rows)
      load_tile(td[0], 0);
      layer0(td[0]);                          // -> R1 (free: every MMA of the previous group has completed)
      if (np == 2) load_tile(td[1], 1);       // under MMA(X, 1)
      PH_ADD(ph3, tq);
      for (int s = 1; s < Lt; ++s) {
        const bool last = s == Lt - 1;
        for (int k = 0; k < np; ++k) {
          tq = PH_NOW();
          mbar_wait(bar_z, nd & 1);           // Z of (tile k, layer s) complete; R1 is free
          ++nd;
          tc_fence_after();
          PH_ADD(ph0, tq);
          if (np == 2) {                      // next A operand of the OTHER tile: its GEMM runs under this tile's P2
            tq = PH_NOW();
            if (k == 0) {
              if (s == 1) layer0(td[1]); else write_A();       // A of (Y, s)
            } else if (!last) {
              write_A();                                        // A of (X, s + 1)
            }
            PH_ADD(ph11, tq);
          }
          tq = PH_NOW();
          drain();
          epi_bar();
          PH_ADD(ph1, tq);
          tq = PH_NOW();
          gather_combine(td[k], k, s);
          epi_bar();
          PH_ADD(ph2, tq);
          if (last) {
            tq = PH_NOW();
            run_head(td[k], k);
            release_tile(td[k]);
            PH_ADD(ph4, tq);
          } else if (np == 1) {
            tq = PH_NOW();
            write_A();
            PH_ADD(ph11, tq);
          }
        }
      }
    }
    if (tid == 0 && blockIdx.x < 148) {
      long long* o = g_fwd_phase + blockIdx.x * 16;
      o[0] = ph0; o[1] = ph1; o[2] = ph2; o[3] = ph3; o[4] = ph4; o[8] = PH_NOW() - t_kernel0; o[9] = ntl;
      o[11] = ph11;
    }
  }
  // ---- teardown ----------------------------------------------------------------------------------------------------------
  tc_fence_before();
  __syncthreads();
  if (wid == kTcEpiWarps + 1) tmem_dealloc(tbase, TM_COLS);
}

static int fwd_max_pair() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("HEXGNN_FWD_PAIR");
    v = (e && e[0] == '1') ? 1 : 2;
  }
  return v;
}

static int fwd_pack_grid() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("HEXGNN_FWD_PACK_GRID");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v;
}

template <int NJ>
static int launch_fwd_tc_t(FwdTcArgs& a, cudaStream_t st) {
  using S = FwdTcSmem<NJ>;
  auto kern = sage_fwd_tc_kernel<NJ>;
  int stages = kFwdMaxStages;
  while (stages > 2 && S::bytes(a.d.Lt, stages) > 227 * 1024) --stages;
  const size_t smem = S::bytes(a.d.Lt, stages);
  if (smem > 227 * 1024) { set_error("forward(tc): %zu bytes of shared memory needed", smem); return HEXGNN_E_UNSUPPORTED; }
  a.stages = stages;
  a.max_pair = fwd_max_pair();
  a.pack_grid = fwd_pack_grid();
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
    default: break;
  }
  set_error("forward(tc): unsupported hidden=%d (padded %d)", d.H, d.Hp);
  return HEXGNN_E_UNSUPPORTED;
}

}  // namespace hexgnn
