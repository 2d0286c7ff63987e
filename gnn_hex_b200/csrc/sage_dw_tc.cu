// Fused backward on the tensor cores, part B: the weight gradients of SAGE layers 1..Lt-1,
//     dWcat_s [Hp(+1) x 2Hp] = [h_s | 1]^T  *  [dZl_s | dpre_s]          (contraction over ALL node rows)
// as a split-K tcgen05 GEMM.  Both operands are the row-major RM32 stash images (sage_tc.cuh) read as MN-major
// SWIZZLE_128B_BASE32B shared-memory operands (SS mode): no transposition anywhere.  Work item = (layer s, group g):
// the CTA walks the tiles t = g, g+G, ... and accumulates in ONE TMEM accumulator, so only G partial records per
// layer exist (G = 148 / (Lt-1)) instead of one per CTA of the chain kernel, and they are summed in fixed order.
// The kernel is launched as a programmatic dependent of the chain kernel: its CTAs become resident on the SMs of chain
// CTAs that have run out of tiles (148 persistent CTAs, e.g. 209 tiles: 87 SMs idle through the second round), take a
// ticket, and consume tiles as the chain releases their flags.  Items are ordered by ticket: phase A = the tiles of all
// complete chain rounds (finished before any chain CTA retires), phase B = the last round's tiles; the phase-B CTA of
// (layer, group) adds its accumulator onto the phase-A record in a fixed order, so results stay bit-reproducible.
// Warp roles: producer (2 KB bulk copies of 16-row k-chunks), 8 split warps (lo = a - trunc_tf32(a) copy; hi = the fp32 image,
// 3xTF32), MMA issuer, and the split warps double as the epilogue.
#include "sage_tc.cuh"

namespace hexgnn {
using namespace tc;

constexpr int kDwSplitWarps = 8;
constexpr int kDwSplitThreads = kDwSplitWarps * 32;
constexpr int kDwThreads = kDwSplitThreads + 64;
constexpr int kDwStages = 4;
constexpr int kDwChunkRows = 16;              // node rows (= MMA k) per ring stage
constexpr int kDwBlk = kDwChunkRows * 128;     // bytes of one 32-channel block of a chunk

struct DwArgs {
  Dims d;
  GradLayout gl;
  int64_t B;
  CsrView csr;
  const float* stash;   // activations (RM32 images)
  const float* dz;      // [dZl | dpre] (RM32 images)
  float* big_rec;       // [G][gl.g0]
  int G;
  int chain_grid;       // CTAs of the chain kernel (tiles of one chain round)
  int* sync;            // sage_tc.cuh: ticket | tile flags | item flags
};

template <int NJ>
__global__ void __launch_bounds__(kDwThreads, 1) sage_dw_tc_kernel(const __grid_constant__ DwArgs a) {
  constexpr int Hp = 16 * NJ, N2 = 32 * NJ;
  constexpr int ABLK = 4;                                   // A blocks addressed by the M = 128 descriptor
  constexpr int a_bytes = ABLK * kDwBlk, b_bytes = NJ * kDwBlk; // one row chunk of each operand (fp32)
  constexpr int half_bytes = a_bytes + b_bytes;             // hi half; the lo half follows
  constexpr int stage_bytes = 2 * half_bytes;
  // two accumulators: D (columns [0, N2)) takes the hi * hi products, Ds ([N2, 2 N2)) the lo * hi and hi * lo ones.  The tensor
  // core truncates its fp32 accumulator after every MMA (~2^-25 of the running sum each), and one item accumulates hundreds of
  // MMAs (every 8 node rows of its tiles: three); with the small products elsewhere only a third of them truncate the large
  // sum, and the epilogue adds D + Ds in round-to-nearest fp32.
  constexpr uint32_t TM_DS = N2;
  constexpr uint32_t TM_COLS = (2 * N2 <= 32) ? 32 : (2 * N2 <= 64) ? 64 : (2 * N2 <= 128) ? 128 : (2 * N2 <= 256) ? 256 : 512;
  extern __shared__ uint8_t smem_dyn[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar_full[kDwStages], bar_split[kDwStages], bar_empty[kDwStages], bar_done;
  __shared__ uint32_t tmem_slot;
  __shared__ int ticket_s;

  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) ticket_s = atomicAdd(a.sync, 1);
  __syncthreads();
  const int nitems = (int)gridDim.x >> 1;                  // (Lt-1) * G items per phase
  const int phase = ticket_s >= nitems, item = ticket_s - phase * nitems;
  const int s = 1 + item / a.G, grp = item % a.G;
  const int ntiles = *a.csr.ntiles;
  // positions in the chain's processing order (csr.tile_order): phase A = [0, TA) = all complete chain rounds but the
  // last; phase B = the last round [TA, ntiles)
  const int TA = ntiles > 0 ? ((ntiles - 1) / a.chain_grid) * a.chain_grid : 0;
  const int t_begin = (phase ? TA : 0) + grp, t_end = phase ? ntiles : TA;
  const int* tile_flag = a.sync + kSyncTileOff;
  int* item_flag = a.sync + kSyncTileOff + a.B + item;
  if (phase == 0 && TA == 0) return;                       // nothing before the last round: phase B writes the record
  const int CB = tc_act_blocks(Hp);
  const size_t act_img = (size_t)CB * kImgBlockBytes, dz_img = (size_t)NJ * kImgBlockBytes;

  if (tid == 0) {
    for (int i = 0; i < kDwStages; ++i) { mbar_init(&bar_full[i], 1); mbar_init(&bar_split[i], kDwSplitThreads); mbar_init(&bar_empty[i], 1); }
    mbar_init(&bar_done, 1);
    fence_mbar_init();
  }
  if (wid == kDwSplitWarps + 1) tmem_alloc(&tmem_slot, TM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_slot;

  auto tile_rows32 = [&](int t) {
    const int gA = a.csr.tile_g0[t], gB = a.csr.tile_g0[t + 1];
    const int n = a.csr.gptr[gB] - a.csr.gptr[gA];
    return (n < 0 || n > kTcRows) ? 0 : ((n + kDwChunkRows - 1) & ~(kDwChunkRows - 1));   // rows [n, ceil32(n)) are zeros
  };

  if (wid == kDwSplitWarps) {
    // ================= producer =================
    if (lane == 0) {
      uint32_t it = 0;
      for (int ti = t_begin; ti < t_end; ti += a.G) {
        const int t = a.csr.tile_order[ti];
        const int n32 = tile_rows32(t);
        if (n32 > 0) { flag_wait(tile_flag + t); fence_proxy_async_global(); }
        const char* act = (const char*)a.stash + ((size_t)(s - 1) * a.B + t) * act_img;
        const char* dzp = (const char*)a.dz + ((size_t)(s - 1) * a.B + t) * dz_img;
        for (int c0 = 0; c0 < n32; c0 += kDwChunkRows, ++it) {
          const int st = it % kDwStages;
          mbar_wait(&bar_empty[st], ((it / kDwStages) & 1) ^ 1);
          uint8_t* dstA = smem + st * stage_bytes;
          uint8_t* dstB = dstA + a_bytes;
          mbar_expect_tx(&bar_full[st], (uint32_t)(CB + NJ) * (uint32_t)kDwBlk);
          for (int b = 0; b < CB; ++b) bulk_g2s(dstA + b * kDwBlk, act + (size_t)b * kImgBlockBytes + (size_t)c0 * 128, kDwBlk, &bar_full[st]);
          for (int b = 0; b < NJ; ++b) bulk_g2s(dstB + b * kDwBlk, dzp + (size_t)b * kImgBlockBytes + (size_t)c0 * 128, kDwBlk, &bar_full[st]);
        }
      }
    }
  } else if (wid == kDwSplitWarps + 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_tf32(128, N2, true, true);
      uint32_t it = 0;
      for (int ti = t_begin; ti < t_end; ti += a.G) {
        const int t = a.csr.tile_order[ti];
        const int n32 = tile_rows32(t);
        for (int c0 = 0; c0 < n32; c0 += kDwChunkRows, ++it) {
          const int st = it % kDwStages;
          mbar_wait(&bar_split[st], (it / kDwStages) & 1);
          tc_fence_after();
          const uint32_t ahi = smem_u32(smem + st * stage_bytes), bhi = ahi + a_bytes;
          const uint32_t alo = ahi + half_bytes, blo = bhi + half_bytes;
#pragma unroll
          for (int ks = 0; ks < kDwChunkRows / 8; ++ks) {
            const uint64_t one = (uint64_t)1 << 61;   // SWIZZLE_128B_BASE32B
            const uint64_t dah = make_smem_desc(ahi + ks * 1024, kDwBlk, 512) | one;
            const uint64_t dal = make_smem_desc(alo + ks * 1024, kDwBlk, 512) | one;
            const uint64_t dbh = make_smem_desc(bhi + ks * 1024, kDwBlk, 512) | one;
            const uint64_t dbl = make_smem_desc(blo + ks * 1024, kDwBlk, 512) | one;
            mma_tf32_ss(tbase + TM_DS, dal, dbh, idesc, (it | ks) != 0);
            mma_tf32_ss(tbase + TM_DS, dah, dbl, idesc, true);
            mma_tf32_ss(tbase, dah, dbh, idesc, (it | ks) != 0);
          }
          mma_commit(&bar_empty[st]);
        }
      }
      mma_commit(&bar_done);
    }
  } else {
    // ================= split warps: fp32 -> (tf32 hi in place, tf32 lo copy) =================
    uint32_t it = 0;
    for (int ti = t_begin; ti < t_end; ti += a.G) {
        const int t = a.csr.tile_order[ti];
      const int n32 = tile_rows32(t);
      for (int c0 = 0; c0 < n32; c0 += kDwChunkRows, ++it) {
        const int st = it % kDwStages;
        mbar_wait(&bar_full[st], (it / kDwStages) & 1);
        float4* base = (float4*)(smem + st * stage_bytes);
        float4* lob = (float4*)(smem + st * stage_bytes + half_bytes);
        const int nA = CB * (kDwBlk / 16), nB = NJ * (kDwBlk / 16);       // float4 per region
        for (int i = tid; i < nA + nB; i += kDwSplitThreads) {
          const int idx = i < nA ? i : (ABLK * (kDwBlk / 16) + (i - nA));
          const float4 v = base[idx];
          // hi = the fp32 image itself: the tensor core reads tf32 operands by ignoring the low 13 mantissa bits
          // (truncation), so only lo = rn_tf32(a - trunc_tf32(a)) has to be produced: a = hi + lo to 2^-22 |a|
          float4 l;
          l.x = tf32_rn(v.x - tf32_trunc(v.x)); l.y = tf32_rn(v.y - tf32_trunc(v.y));
          l.z = tf32_rn(v.z - tf32_trunc(v.z)); l.w = tf32_rn(v.w - tf32_trunc(v.w));
          lob[idx] = l;
        }
        fence_proxy_async();
        mbar_arrive(&bar_split[st]);
      }
    }
    // ================= epilogue: D[k][c'] -> this group's record =================
    mbar_wait(&bar_done, 0);
    tc_fence_after();
    const bool add = phase == 1 && TA > 0;                 // phase B adds onto the phase-A record (fixed order: A + B)
    if (add && tid == 0) flag_wait(item_flag);
    asm volatile("bar.sync 2, %0;" ::"n"(kDwSplitThreads) : "memory");
    if (wid < 4) {
      const int m = wid * 32 + lane;                       // D lane = input channel k (Hp = ones channel)
      const uint32_t trow = tbase + ((uint32_t)(wid * 32) << 16);
      float* rec = a.big_rec + (int64_t)grp * a.gl.g0;
      const bool have = it > 0;
      for (int c0 = 0; c0 < N2; c0 += 8) {
        float v[8], vs[8];
        tmem_ld8(trow + c0, v);                            // warp-collective: every lane takes part
        tmem_ld8(trow + TM_DS + c0, vs);
        tmem_wait_ld();
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] += vs[u];
        if (!have) {
#pragma unroll
          for (int u = 0; u < 8; ++u) v[u] = 0.f;
        }
        float4* dst = nullptr;
        if (m < Hp) dst = (float4*)(rec + a.gl.gw + ((int64_t)(s - 1) * Hp + m) * N2 + c0);
        else if (m == Hp && c0 >= Hp) dst = (float4*)(rec + a.gl.gb + (int64_t)s * Hp + (c0 - Hp));
        if (dst) {
          if (add) {
            const float4 p0 = __ldcg(dst), p1 = __ldcg(dst + 1);
            v[0] += p0.x; v[1] += p0.y; v[2] += p0.z; v[3] += p0.w;
            v[4] += p1.x; v[5] += p1.y; v[6] += p1.z; v[7] += p1.w;
          }
          dst[0] = make_float4(v[0], v[1], v[2], v[3]);
          dst[1] = make_float4(v[4], v[5], v[6], v[7]);
        }
      }
    }
    if (phase == 0) {                                      // record complete: release it to the phase-B CTA of this item
      asm volatile("bar.sync 2, %0;" ::"n"(kDwSplitThreads) : "memory");
      if (tid == 0) { __threadfence(); flag_release(item_flag, 1); }
    }
  }
  if (ticket_s == (int)gridDim.x - 1 && tid == 0) {        // see the end of sage_bwd_tc_kernel
    const long long t0 = clock64();
    while (flag_acquire(a.sync + 1) < a.chain_grid) {
      __nanosleep(500);
      if (clock64() - t0 > 8000000000ll) __trap();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (wid == kDwSplitWarps + 1) tmem_dealloc(tbase, TM_COLS);
}

int tc_dw_groups(const Dims& d) {
  int g = bwd_grid_ctas() / (d.Lt - 1);
  return g < 1 ? 1 : g;
}

template <int NJ>
static int launch_dw_t(const DwArgs& a, cudaStream_t st) {
  auto kern = sage_dw_tc_kernel<NJ>;
  const size_t smem = (size_t)kDwStages * 2 * (4 * kDwBlk + NJ * kDwBlk) + 1024;
  HEX_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * (a.d.Lt - 1) * a.G);              // phase A + phase B items, ordered by ticket
  cfg.blockDim = dim3(kDwThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;   // may start while the chain kernel still runs
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  HEX_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, a));
  return HEXGNN_OK;
}

int launch_backward_tc_dw(const Dims& d, int64_t N, int64_t E, int64_t B, const void* csr, const float* stash,
                          const float* dz, float* big_rec, int* sync, cudaStream_t st) {
  DwArgs a;
  a.d = d; grad_layout(d, &a.gl); a.B = B;
  a.csr = csr_view(const_cast<void*>(csr), N, E, B);
  a.stash = stash; a.dz = dz; a.big_rec = big_rec; a.G = tc_dw_groups(d);
  a.chain_grid = bwd_grid_ctas(); a.sync = sync;
  switch (d.NJ) {
    case 1: return launch_dw_t<1>(a, st);
    case 2: return launch_dw_t<2>(a, st);
    case 3: return launch_dw_t<3>(a, st);
    case 4: return launch_dw_t<4>(a, st);
    case 5: return launch_dw_t<5>(a, st);
    case 6: return launch_dw_t<6>(a, st);
    case 7: return launch_dw_t<7>(a, st);
    default: break;
  }
  set_error("backward(tc dW): unsupported hidden=%d", d.H);
  return HEXGNN_E_UNSUPPORTED;
}

}  // namespace hexgnn
