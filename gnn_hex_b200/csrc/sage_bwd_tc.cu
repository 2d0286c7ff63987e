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
// gradients go to a small per-CTA record, reduced in fixed order by reduce_tc_kernel (pack_tc.cu).
// max/min pooling route their gradient to the FIRST row attaining the extremum (torch_scatter CPU semantics).
#include "sage_tc.cuh"

namespace hexgnn {
using namespace tc;

constexpr int kBwdStages = 6;

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
  TcPackView pv;
  float* small_rec;       // [gridDim.x][gl.total - gl.gb]
};

template <int NJ>
struct BwdTcSmem {
  static constexpr int Hp = 16 * NJ;
  static constexpr int P = Hp + 4;
  static constexpr int ECAP = kEdgeCapPerRow * kTcRows;
  static constexpr int half_bytes = kTcKC * Hp * 4;
  static constexpr int stage_bytes = 2 * half_bytes;
  static constexpr int off_ring = 0;
  static constexpr int off_dump = off_ring + kBwdStages * stage_bytes;           // float [128][P]
  static constexpr int off_rows = off_dump + kTcRows * P * 4;                    // float [7][128]
  static constexpr int off_pool = off_rows + 7 * kTcRows * 4;                    // float [4*Hp]
  static constexpr int off_hid = off_pool + 4 * Hp * 4;                          // float [Hp]
  static constexpr int off_arg = off_hid + Hp * 4;                               // int [2*Hp]
  static constexpr int off_misc = off_arg + 2 * Hp * 4;                          // float [16]
  static constexpr int off_rowptr = off_misc + 64;                               // int [128+4]
  static constexpr int off_gp = off_rowptr + (kTcRows + 4) * 4;                  // int [64+4]
  static constexpr int off_col = off_gp + (kMaxTileGraphs + 4) * 4;              // u16 [ECAP]
  static constexpr int off_bar = off_col + ECAP * 2;                             // u64 [2*stages + 3], u32 slot
  static constexpr int bytes = off_bar + (2 * kBwdStages + 3) * 8 + 16;
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
// [4] head + load  [5] layer 0  [8] total  [9] tiles
__device__ long long g_bwd_phase[148 * 16];

template <int NJ>
__global__ void __launch_bounds__(kTcThreads, 1) sage_bwd_tc_kernel(const __grid_constant__ BwdTcArgs a) {
  using S = BwdTcSmem<NJ>;
  constexpr int Hp = S::Hp, P = S::P;
  constexpr int NG = 2 * NJ;
  constexpr int NQ = Hp / kTcKC;
  constexpr int MAXG = (NG + kTcSlices - 1) / kTcSlices;
  constexpr uint32_t TM_HI = 0, TM_LO = Hp, TM_D = 2 * Hp;
  constexpr uint32_t TM_COLS = (3 * Hp <= 64) ? 64 : (3 * Hp <= 128) ? 128 : (3 * Hp <= 256) ? 256 : 512;
  extern __shared__ __align__(128) uint8_t smem[];
  float* dump = (float*)(smem + S::off_dump);
  float* rows = (float*)(smem + S::off_rows);
  float *inv_s = rows, *x0_s = rows + 128, *x1_s = rows + 256, *m0_s = rows + 384, *m1_s = rows + 512,
        *adv_s = rows + 640, *th_s = rows + 768;
  float* pool = (float*)(smem + S::off_pool);
  float* hid = (float*)(smem + S::off_hid);
  int* arg_s = (int*)(smem + S::off_arg);
  float* misc = (float*)(smem + S::off_misc);
  int* rowptr_s = (int*)(smem + S::off_rowptr);
  int* gp_s = (int*)(smem + S::off_gp);
  unsigned short* col_s = (unsigned short*)(smem + S::off_col);
  uint64_t* bar_full = (uint64_t*)(smem + S::off_bar);
  uint64_t* bar_empty = bar_full + kBwdStages;
  uint64_t* bar_a = bar_empty + kBwdStages;
  uint64_t* bar_r = bar_a + 1;
  uint64_t* bar_d = bar_r + 1;
  uint32_t* tmem_slot = (uint32_t*)(bar_d + 1);

  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int H = a.d.H, H2 = a.d.H2, Lt = a.d.Lt;
  const int ntiles = *a.csr.ntiles;
  const long long t_kernel0 = clock64();
  long long ph0 = 0, ph1 = 0, ph2 = 0, ph3 = 0, ph4 = 0, ph5 = 0, ntl = 0;

  if (tid == 0) {
    for (int i = 0; i < kBwdStages; ++i) { mbar_init(&bar_full[i], 1); mbar_init(&bar_empty[i], 1); }
    mbar_init(bar_a, kTcEpiThreads);
    mbar_init(bar_r, 1);
    mbar_init(bar_d, 1);
    fence_mbar_init();
  }
  if (wid == kTcEpiWarps + 1) tmem_alloc(tmem_slot, TM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = *tmem_slot;

  if (wid == kTcEpiWarps) {
    // ================= producer: backward weight images -> ring =================
    if (lane == 0) {
      uint32_t it = 0;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
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
      constexpr uint32_t idesc = make_idesc_tf32(kTcRows, Hp, false);
      constexpr uint32_t lbo = Hp * 16, sbo = 128;
      uint32_t it = 0, na = 0;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
        if (!tile_dims_b(a.csr, t, S::ECAP).ok) continue;
        for (int s = Lt - 1; s >= 1; --s) {
          for (int which = 0; which < 2; ++which) {
            mbar_wait(bar_a, na & 1);   // A operand (dpre, then dZl) is in TMEM
            ++na;
            tc_fence_after();
            for (int q = 0; q < NQ; ++q, ++it) {
              const int st = it % kBwdStages;
              mbar_wait(&bar_full[st], (it / kBwdStages) & 1);
              tc_fence_after();
              const uint32_t bhi = smem_u32(smem + S::off_ring + st * S::stage_bytes);
              const uint32_t blo = bhi + S::half_bytes;
#pragma unroll
              for (int kk = 0; kk < kTcKC / 8; ++kk) {
                const uint64_t dhi = make_smem_desc(bhi + kk * 2 * lbo, lbo, sbo);
                const uint64_t dlo = make_smem_desc(blo + kk * 2 * lbo, lbo, sbo);
                const uint32_t kcol = q * kTcKC + kk * 8;
                mma_tf32_ts(tbase + TM_D, tbase + TM_LO + kcol, dhi, idesc, (which | q | kk) != 0);
                mma_tf32_ts(tbase + TM_D, tbase + TM_HI + kcol, dlo, idesc, true);
                mma_tf32_ts(tbase + TM_D, tbase + TM_HI + kcol, dhi, idesc, true);
              }
              mma_commit(&bar_empty[st]);
            }
            mma_commit(which == 0 ? bar_r : bar_d);
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
    float* rec = a.small_rec + (int64_t)blockIdx.x * recn - a.gl.gb;   // index with GradLayout offsets (>= gl.gb)
    for (int64_t i = a.gl.gb + tid; i < a.gl.total; i += kTcEpiThreads) rec[i] = 0.f;
    const float* wa = Pm + a.pm.lin_w[head];
    const float ba = Pm[a.pm.lin_b[head]];
    const float* w0 = Pm + a.pm.v0_w[head];
    const float* b0 = Pm + a.pm.v0_b[head];
    const float* w1 = Pm + a.pm.v1_w[head];
    const float b1 = Pm[a.pm.v1_b[head]];
    const size_t act_img = (size_t)tc_act_blocks(Hp) * kImgBlockBytes;
    const size_t dz_img = (size_t)tc_dz_blocks(Hp) * kImgBlockBytes;
    uint32_t nr = 0, nd = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
      const TileDimsB td = tile_dims_b(a.csr, t, S::ECAP);
      const int n = td.n, r0 = td.r0, g0 = td.g0, g1 = td.g1;
      epi_bar_b();
      long long tq = clock64();
      ++ntl;
      if (!td.ok) {
        if (tid == 0) { rec[a.gl.gb] = __int_as_float(0x7fc00000); atomicOr(a.csr.status, HEXGNN_ST_GRAPH_TOO_LARGE); }
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
        }
        if (bad) atomicOr(a.csr.status, bad);
        epi_bar_b();
        for (int i = tid; i < kTcRows; i += kTcEpiThreads) {
          const int b = rowptr_s[i], e = rowptr_s[i + 1];
          const int deg = e - b;
          const float inv = 1.0f / (float)(deg > 0 ? deg : 1);
          float s0 = 0.f, s1 = 0.f;
          for (int p = b; p < e; ++p) { const int j = col_s[p]; s0 += x0_s[j]; s1 += x1_s[j]; }
          inv_s[i] = inv; m0_s[i] = s0 * inv; m1_s[i] = s1 * inv;
        }
        epi_bar_b();
      }
      const bool live = row < n;
      const int n32 = (n + 31) & ~31;
      const float inv = inv_s[row];
      const int eb = rowptr_s[row], ee = rowptr_s[row + 1];
      const char* act_t = (const char*)a.stash + (size_t)t * act_img;         // + layer * B * act_img
      char* dz_t = (char*)a.dz + (size_t)t * dz_img;                           // + (s-1) * B * dz_img
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
      if (tid < kTcRows && tid < n) {
        float u = 0.f;
        const float* gr = dump + tid * P;
        for (int c = 0; c < H; ++c) u = fmaf(wa[c], gr[c], u);
        th_s[tid] = tanhf(u + ba);
      }
      epi_bar_b();
      // ---- head backward per graph: dump rows become dL/dg in place -----------------------------------------
      for (int g = 0; g < g1 - g0; ++g) {
        const int rs = gp_s[g], re = gp_s[g + 1];
        const int ng = re - rs;
        if (ng <= 0) continue;
        const bool adv_only = (a.flags & HEXGNN_ADV_ONLY) != 0;
        float sumdq = 0.f;
        if (!adv_only) {
          if (wid == 0) {
            float u = 0.f;
            for (int i = rs + lane; i < re; i += 32) u += a.d_nodes[r0 + i];
            u = warp_sum_b(u);
            if (lane == 0) misc[0] = u;
          }
          if (tid < H) {
            float sum = 0.f, mx = -INFINITY, mn = INFINITY;
            int amx = rs, amn = rs;
            for (int i = rs; i < re; ++i) {
              const float v = dump[i * P + tid];
              sum += v;
              if (v > mx) { mx = v; amx = i; }
              if (v < mn) { mn = v; amn = i; }
            }
            pool[tid] = sum; pool[H + tid] = mx; pool[2 * H + tid] = mn; pool[3 * H + tid] = sum / (float)ng;
            arg_s[tid] = amx; arg_s[Hp + tid] = amn;
          }
          epi_bar_b();
          sumdq = misc[0];
        }
        for (int i = rs + tid; i < re; i += kTcEpiThreads) {
          const float th = th_s[i];
          const float dn = a.d_nodes[r0 + i];
          adv_s[i] = (adv_only ? dn : dn - sumdq / (float)ng) * 2.0f * (1.0f - th * th);
        }
        if (!adv_only) {
          for (int k = wid; k < H2; k += kTcEpiWarps) {
            float u = 0.f;
            const float* wrow = w0 + (int64_t)k * 4 * H;
            for (int q = lane; q < 4 * H; q += 32) u = fmaf(wrow[q], pool[q], u);
            u = warp_sum_b(u);
            if (lane == 0) hid[k] = fmaxf(u + b0[k], 0.f);
          }
        }
        epi_bar_b();
        // head.linear grads need g (still in dump) and du
        if (tid < H) {
          float u = 0.f;
          for (int i = rs; i < re; ++i) u = fmaf(adv_s[i], dump[i * P + tid], u);
          rec[a.gl.glin + tid] += u;
        } else if (tid == H) {
          float u = 0.f;
          for (int i = rs; i < re; ++i) u += adv_s[i];
          rec[a.gl.glinb] += u;
        }
        if (!adv_only) {
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
            rec[a.gl.gv1 + tid] += ds * hk;
            const float dt = hk > 0.f ? ds * w1[tid] : 0.f;
            rec[a.gl.gv0b + tid] += dt;
            hid[tid] = dt;
          }
          if (tid == 0) rec[a.gl.gv1b] += ds;
          epi_bar_b();
          for (int idx = tid; idx < H2 * 4 * H; idx += kTcEpiThreads) {
            const int k = idx / (4 * H), q = idx - k * 4 * H;
            rec[a.gl.gv0 + idx] += hid[k] * pool[q];
          }
          epi_bar_b();  // pool (P) fully consumed before it becomes dP
          float dp = 0.f;
          if (tid < 4 * H) {
            for (int k = 0; k < H2; ++k) dp = fmaf(w0[(int64_t)k * 4 * H + tid], hid[k], dp);
          }
          epi_bar_b();
          if (tid < 4 * H) pool[tid] = dp;
        }
        epi_bar_b();
        for (int idx = tid; idx < ng * Hp; idx += kTcEpiThreads) {
          const int i = rs + idx / Hp, c = idx % Hp;
          float v = 0.f;
          if (c < H) {
            v = adv_s[i] * wa[c];
            if (!adv_only) {
              v += pool[c] + pool[3 * H + c] / (float)ng;
              if (arg_s[c] == i) v += pool[H + c];
              if (arg_s[Hp + c] == i) v += pool[2 * H + c];
            }
          }
          dump[i * P + c] = v;
        }
        epi_bar_b();
      }
      ph4 += clock64() - tq;
      // ---- SAGE layers Lt-1 .. 1 on the tensor cores ---------------------------------------------------------
      for (int s = Lt - 1; s >= 1; --s) {
        const bool first = (s == Lt - 1);
        tq = clock64();
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
          epi_bar_b();   // everyone holds its dH before the dump is reused
        } else {
          mbar_wait(bar_d, nd & 1);   // dH_{s+1} complete in TMEM
          ++nd;
          tc_fence_after();
          ph0 += clock64() - tq;
          tq = clock64();
#pragma unroll
          for (int gi = 0; gi < MAXG; ++gi) {
            const int g = slice + gi * kTcSlices;
            if (g < NG) tmem_ld8(trow + TM_D + g * 8, dh[gi]);
          }
          tmem_wait_ld();
        }
        // E1: ReLU mask, dpre -> TMEM (A of dpre*W_r), dump (pre-scaled by 1/deg for the gather), dZ stash
        {
          const char* himg = act_t + (size_t)s * a.B * act_img;          // h_{s+1}
          char* zimg = dz_t + (size_t)(s - 1) * a.B * dz_img;
#pragma unroll
          for (int gi = 0; gi < MAXG; ++gi) {
            const int g = slice + gi * kTcSlices;
            if (g < NG) {
              if (s == a.d.L - 1 && a.d_embeds && live) {
                float4* de = (float4*)(a.d_embeds + (int64_t)(r0 + row) * Hp + g * 8);
                de[0] = make_float4(dh[gi][0], dh[gi][1], dh[gi][2], dh[gi][3]);
                de[1] = make_float4(dh[gi][4], dh[gi][5], dh[gi][6], dh[gi][7]);
              }
              float hv[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
              if (live) {
                const float4* src = (const float4*)(himg + rm32_granule_off(g, row));
                const float4 v0 = src[0], v1 = src[1];
                hv[0] = v0.x; hv[1] = v0.y; hv[2] = v0.z; hv[3] = v0.w; hv[4] = v1.x; hv[5] = v1.y; hv[6] = v1.z; hv[7] = v1.w;
              }
              float hi[8], lo[8], sc[8];
#pragma unroll
              for (int u = 0; u < 8; ++u) {
                const float r = hv[u] > 0.f ? dh[gi][u] : 0.f;
                dh[gi][u] = r;
                sc[u] = r * inv;
                split_tf32(r, hi[u], lo[u]);
              }
              tmem_st8(trow + TM_HI + g * 8, hi);
              tmem_st8(trow + TM_LO + g * 8, lo);
              float4* dd = (float4*)(dump + row * P + g * 8);
              dd[0] = make_float4(sc[0], sc[1], sc[2], sc[3]);
              dd[1] = make_float4(sc[4], sc[5], sc[6], sc[7]);
              if (row < n32) {
                float4* dst = (float4*)(zimg + rm32_granule_off(NG + g, row));
                dst[0] = make_float4(dh[gi][0], dh[gi][1], dh[gi][2], dh[gi][3]);
                dst[1] = make_float4(dh[gi][4], dh[gi][5], dh[gi][6], dh[gi][7]);
              }
            }
          }
          tmem_wait_st();
          tc_fence_before();
          mbar_arrive(bar_a);        // -> MMA: dH_s = dpre * W_r
        }
        epi_bar_b();                 // dump (dpre / deg) complete
        ph1 += clock64() - tq;
        tq = clock64();
        // E2: dZl_j = sum_{i in N(j)} dpre_i / deg(i)
        float acc[MAXG][8];
#pragma unroll
        for (int gi = 0; gi < MAXG; ++gi)
#pragma unroll
          for (int u = 0; u < 8; ++u) acc[gi][u] = 0.f;
        {
          int j = eb < ee ? (int)col_s[eb] : 0;
          for (int p = eb; p < ee; ++p) {
            const int jn = (p + 1 < ee) ? (int)col_s[p + 1] : 0;
            const float* src = dump + j * P + slice * 8;
#pragma unroll
            for (int gi = 0; gi < MAXG; ++gi) {
              if (slice + gi * kTcSlices < NG) {
                const float4 v0 = *(const float4*)(src + gi * kTcSlices * 8);
                const float4 v1 = *(const float4*)(src + gi * kTcSlices * 8 + 4);
                acc[gi][0] += v0.x; acc[gi][1] += v0.y; acc[gi][2] += v0.z; acc[gi][3] += v0.w;
                acc[gi][4] += v1.x; acc[gi][5] += v1.y; acc[gi][6] += v1.z; acc[gi][7] += v1.w;
              }
            }
            j = jn;
          }
        }
        {
          char* zimg = dz_t + (size_t)(s - 1) * a.B * dz_img;
          if (row < n32) {
#pragma unroll
            for (int gi = 0; gi < MAXG; ++gi) {
              const int g = slice + gi * kTcSlices;
              if (g < NG) {
                float4* dst = (float4*)(zimg + rm32_granule_off(g, row));
                dst[0] = make_float4(acc[gi][0], acc[gi][1], acc[gi][2], acc[gi][3]);
                dst[1] = make_float4(acc[gi][4], acc[gi][5], acc[gi][6], acc[gi][7]);
              }
            }
          }
        }
        ph2 += clock64() - tq;
        tq = clock64();
        mbar_wait(bar_r, nr & 1);      // dpre * W_r has finished reading the A columns
        ++nr;
        tc_fence_after();
#pragma unroll
        for (int gi = 0; gi < MAXG; ++gi) {
          const int g = slice + gi * kTcSlices;
          if (g < NG) {
            float hi[8], lo[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) split_tf32(acc[gi][u], hi[u], lo[u]);
            tmem_st8(trow + TM_HI + g * 8, hi);
            tmem_st8(trow + TM_LO + g * 8, lo);
          }
        }
        tmem_wait_st();
        tc_fence_before();
        mbar_arrive(bar_a);            // -> MMA: dH_s += dZl * W_l
        ph3 += clock64() - tq;
      }
      // ---- layer 0 (2 -> H): dpre_0 = dH_1 (.) (h_1 > 0); rank-2 weight gradients on the CUDA cores -------------------
      tq = clock64();
      {
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
          epi_bar_b();
        } else {
          mbar_wait(bar_d, nd & 1);
          ++nd;
          tc_fence_after();
#pragma unroll
          for (int gi = 0; gi < MAXG; ++gi) {
            const int g = slice + gi * kTcSlices;
            if (g < NG) tmem_ld8(trow + TM_D + g * 8, dh[gi]);
          }
          tmem_wait_ld();
        }
        const char* himg = act_t;   // h_1
#pragma unroll
        for (int gi = 0; gi < MAXG; ++gi) {
          const int g = slice + gi * kTcSlices;
          if (g < NG) {
            if (a.d.L == 1 && a.d_embeds && live) {
              float4* de = (float4*)(a.d_embeds + (int64_t)(r0 + row) * Hp + g * 8);
              de[0] = make_float4(dh[gi][0], dh[gi][1], dh[gi][2], dh[gi][3]);
              de[1] = make_float4(dh[gi][4], dh[gi][5], dh[gi][6], dh[gi][7]);
            }
            float hv[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (live) {
              const float4* src = (const float4*)(himg + rm32_granule_off(g, row));
              const float4 v0 = src[0], v1 = src[1];
              hv[0] = v0.x; hv[1] = v0.y; hv[2] = v0.z; hv[3] = v0.w; hv[4] = v1.x; hv[5] = v1.y; hv[6] = v1.z; hv[7] = v1.w;
            }
            float4* dd = (float4*)(dump + row * P + g * 8);
            dd[0] = make_float4(hv[0] > 0.f ? dh[gi][0] : 0.f, hv[1] > 0.f ? dh[gi][1] : 0.f, hv[2] > 0.f ? dh[gi][2] : 0.f,
                                hv[3] > 0.f ? dh[gi][3] : 0.f);
            dd[1] = make_float4(hv[4] > 0.f ? dh[gi][4] : 0.f, hv[5] > 0.f ? dh[gi][5] : 0.f, hv[6] > 0.f ? dh[gi][6] : 0.f,
                                hv[7] > 0.f ? dh[gi][7] : 0.f);
          }
        }
        tc_fence_before();
        epi_bar_b();
        if (tid < Hp) {
          float ub = 0.f, u0 = 0.f, u1 = 0.f, u2 = 0.f, u3 = 0.f;
          for (int i = 0; i < n; ++i) {
            const float dpv = dump[i * P + tid];
            ub += dpv;
            u0 = fmaf(dpv, m0_s[i], u0);
            u1 = fmaf(dpv, m1_s[i], u1);
            u2 = fmaf(dpv, x0_s[i], u2);
            u3 = fmaf(dpv, x1_s[i], u3);
          }
          rec[a.gl.gb + tid] += ub;
          rec[a.gl.g0 + 0 * Hp + tid] += u0;
          rec[a.gl.g0 + 1 * Hp + tid] += u1;
          rec[a.gl.g0 + 2 * Hp + tid] += u2;
          rec[a.gl.g0 + 3 * Hp + tid] += u3;
        }
      }
      ph5 += clock64() - tq;
    }
    if (tid == 0 && blockIdx.x < 148) {
      long long* o = g_bwd_phase + blockIdx.x * 16;
      o[0] = ph0; o[1] = ph1; o[2] = ph2; o[3] = ph3; o[4] = ph4; o[5] = ph5; o[8] = clock64() - t_kernel0; o[9] = ntl;
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
  auto kern = sage_bwd_tc_kernel<NJ>;
  HEX_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::bytes));
  {
    ProfScope prof(1, st);
    kern<<<grid, kTcThreads, S::bytes, st>>>(a);
  }
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

int launch_backward_tc_chain(const Dims& d, const ParamMap& pm, const float* params, const float* x, int64_t N, int64_t E,
                             int64_t B, const void* csr, int flags, const float* d_nodes, const float* d_value,
                             const float* stash, float* dz, float* d_embeds, TcPackView pv, float* small_rec, int grid,
                             cudaStream_t st) {
  BwdTcArgs a;
  a.d = d; a.pm = pm; grad_layout(d, &a.gl);
  a.params = params; a.x = x; a.N = N; a.E = E; a.B = B;
  a.csr = csr_view(const_cast<void*>(csr), N, E, B);
  a.flags = flags; a.d_nodes = d_nodes; a.d_value = d_value; a.stash = stash; a.dz = dz; a.d_embeds = d_embeds;
  a.pv = pv; a.small_rec = small_rec;
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
