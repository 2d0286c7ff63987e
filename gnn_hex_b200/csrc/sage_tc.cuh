// Shared definitions of the tcgen05 (tensor-core) fused kernels.
//
// Per SAGE layer the node-linear contraction is ONE tensor-core GEMM per 128-row tile,
//     Z[128, 2Hp] = h[128, Hp] * Wcat[Hp, 2Hp],   Wcat = [lin_l.W^T | lin_r.W^T]
// (transform first, aggregate after: mean_j(h_j) W_l^T == mean_j(h_j W_l^T)), computed as 3xTF32
// (hi*hi + lo*hi + hi*lo, fp32 accumulate) so the result keeps fp32 parity with the reference
// (GN0/torch_script_models.py:52-73 restates the SAGEConv math).  A = h lives in TENSOR MEMORY
// (lane = node row), B = weight images stream through a shared-memory ring from L2, D = Z accumulates
// in tensor memory; the epilogue threads (one per node row) dump Zl to shared memory, gather it over the
// CSR, add Zr + bias, ReLU, and write the next layer's h straight back into tensor memory.
#pragma once
#include "common.cuh"
#include "tc_common.cuh"

// Phase profile of the fused kernels (hexgnn_debug_phases, tools/phase_profile.py): compiled in only with
// -DHEXGNN_PHASES (HEXGNN_PHASES=1 python -m gnn_hex_b200.build --force); the counters cost ~24 registers.
#ifdef HEXGNN_PHASES
#define PH_NOW() clock64()
#define PH_ADD(acc, t0) do { (acc) += clock64() - (t0); } while (0)
#else
#define PH_NOW() 0ll
#define PH_ADD(acc, t0) do { (void)(t0); } while (0)
#endif

namespace hexgnn {

constexpr int kTcRows = 128;          // MMA M: node rows per tile
constexpr int kTcKC = 16;             // k values per streamed weight chunk (2 MMA k-steps)
constexpr int kTcStages = 4;          // ring depth
constexpr int kTcEpiWarps = 16;       // epilogue warps: 4 TMEM lane quarters x 4 column slices
constexpr int kTcSlices = kTcEpiWarps / 4;
constexpr int kTcEpiThreads = kTcEpiWarps * 32;
constexpr int kTcThreads = kTcEpiThreads + 64;  // + producer warp + MMA warp

// Weight images for the forward, one per SAGE layer s = 1..Lt-1 of the taken head:
//   img[s-1][which][p][part][...]  which 0 = lin_l, 1 = lin_r; p = 16-wide k chunk (Hp/16 of them); one (which, p) = one ring
//   stage = one contiguous bulk copy = the B operands of one MMA triple of the all-fp16 3-product scheme below
//   part 0 (Hp*32 bytes) = hi halves, part 1 (Hp*32 bytes) = 2^11-scaled lo halves of W[c][16 p + kk], both fp16, each a
//   K-major SWIZZLE_NONE canonical B operand B[c][kk], c in [0,Hp) the out channel, kk in [0,16):
//          byte = (kk/8) * (Hp*16) + (c/8)*128 + (c%8)*16 + (kk%8)*2
// A layer is two N = Hp GEMMs on the same A operand (Zl first; Zr runs under the CSR gather of Zl), see sage_fwd_tc.cu.
//
// All-fp16 3-product scheme of the forward.  a*b ~= hi_a*hi_b + (lo_a*hi_b + hi_a*lo_b) with 11-bit-mantissa factors, as in
// 3xTF32 -- but fp16 carries the same 11 significant bits as tf32 at half the bytes and twice the k per MMA (K = 16):
//     hi = fp16(x),  lo = fp16(2^11 (x - hi))        (x = hi + 2^-11 lo to 2^-22 |x|; the 2^11 keeps lo out of the subnormals)
//     D  += hi_a * hi_b                               one kind::f16 MMA per 16 k-values (every product exact in fp32)
//     Ds += lo_a * hi_b;  Ds += hi_a * lo_b           two more, into their OWN accumulator
//     Z = D + 2^-11 Ds                                 in the epilogue, fp32 round-to-nearest
// Against 3xTF32: half the weight bytes through the ring (the L2 -> shared stream bounds the GEMM phases), half the
// shared-memory operand reads, and only the Hp/16 large products truncate the large sum (the tensor core truncates its fp32
// accumulator after every MMA).  Exact while |activation|, |weight| < 65504 (fp16 range; below 6e-5 the error is ~1e-11
// absolute instead of relative); activations beyond it set HEXGNN_ST_FP16_RANGE.  The backward keeps 3xTF32: its operands
// (loss gradients, ~1e-6) sit outside fp16's range.
constexpr int kTcKF = 16;             // k values per forward ring stage (one fp16 MMA k-step)
struct TcPackView {
  float* wimg;    // forward images   (Lt-1) * Hp * 2Hp * 2 floats
  float* wimgT;   // backward images  (same size), layout above tc_imgB_floats
  float* bias;    // [Lt][Hp]
};
inline int64_t tc_img_floats(const Dims& d) { return (int64_t)(d.Lt - 1) * d.Hp * 2 * d.Hp * 2; }
inline int64_t tc_pack_floats(const Dims& d);
inline int64_t tc_pack_floats(const Dims& d) { return 2 * tc_img_floats(d) + (int64_t)d.Lt * d.Hp; }
inline TcPackView tc_pack_view(void* ws, const Dims& d) {
  TcPackView p;
  p.wimg = (float*)ws;
  p.wimgT = p.wimg + tc_img_floats(d);   // backward images (tc_imgB_floats == tc_img_floats)
  p.bias = p.wimgT + tc_img_floats(d);
  return p;
}

// ---- training stashes of the tcgen05 path ------------------------------------------------------------
// Both stashes are stored per (layer, tile) as "RM32" images: the row-major [row][channel] tile cut into
// 32-channel (128-byte) blocks, rows grouped by 4 with the 32-byte chunks XOR-swizzled by row%4 -- exactly
// the MN-major SWIZZLE_128B_BASE32B operand form, so the weight-gradient GEMM reads them from shared
// memory with no transposition, and 32 rows of a block are one contiguous 4 KB bulk copy.
//   byte(ch,row) = (ch/32)*16384 + (row/4)*512 + (row%4)*128 + ((((ch%32)/8) ^ (row%4))*32) + (ch%8)*4
// Activations: channels [0,Hp) = h, channel Hp = 1 for live rows (its row of h^T dZ is the bias gradient);
// rows [n, ceil32(n)) are written as zeros so whole 32-row k-chunks can be contracted.
constexpr int kImgBlockBytes = 16384;                       // one 32-channel block of a 128-row tile
__host__ __device__ inline int tc_act_blocks(int Hp) { return (Hp + 1 + 31) / 32; }
__host__ __device__ inline int tc_dz_blocks(int Hp) { return (2 * Hp) / 32; }
__device__ __forceinline__ uint32_t rm32_granule_off(int g, int row) {   // 8-channel granule g, 32 bytes
  return (uint32_t)(g >> 2) * kImgBlockBytes + (uint32_t)(row >> 2) * 512u + (uint32_t)(row & 3) * 128u +
         ((uint32_t)((g & 3) ^ (row & 3)) << 5);
}
// ReLU sign masks of every layer output: one u32 per (layer, tile, column slice, row); bit gi*8+u is set iff
// h[row][(slice + gi*kTcSlices)*8 + u] > 0 -- the (row, slice) thread mapping of both epilogues, so the backward
// chain reads ONE coalesced word per thread and layer instead of the activation image.
inline int64_t tc_mask_words(const Dims& d, int64_t B) { return (int64_t)d.Lt * B * kTcSlices * kTcRows; }
// forward stash: [Lt][B tiles][act_blocks*16KB] images + row-major embedding copy [N][Hp] (final_conv_acts)
// + sign masks [Lt][B][kTcSlices][128] u32
__host__ __device__ inline int64_t tc_stash_mask_off_floats(const Dims& d, int64_t N, int64_t B) {
  return (int64_t)d.Lt * B * tc_act_blocks(d.Hp) * (kImgBlockBytes / 4) + N * d.Hp;
}
// + the value MLP's hidden activations [B graphs][H/2 padded to 4] (the backward reads them instead of recomputing
// relu(W0 pooled + b0): 97 KB of weights per graph)
__host__ __device__ inline int64_t tc_stash_hid_off_floats(const Dims& d, int64_t N, int64_t B) {
  return tc_stash_mask_off_floats(d, N, B) + (int64_t)d.Lt * B * kTcSlices * kTcRows;
}
// + tanh(w_a . g_i + b_a) of every row [B tiles][128] (half the advantage: the backward needs 1 - th^2)
__host__ __device__ inline int64_t tc_stash_th_off_floats(const Dims& d, int64_t N, int64_t B) {
  return tc_stash_hid_off_floats(d, N, B) + B * (int64_t)((d.H2 + 3) & ~3);
}
inline int64_t tc_stash_bytes(const Dims& d, int64_t N, int64_t B) {
  return (tc_stash_th_off_floats(d, N, B) + B * (int64_t)kTcRows) * 4;
}
inline int64_t tc_dz_bytes(const Dims& d, int64_t B) { return (int64_t)(d.Lt - 1) * B * tc_dz_blocks(d.Hp) * kImgBlockBytes; }
inline bool tc_bwd_supported(const Dims& d) { return d.Hp + 1 <= 128; }   // the ones-channel must fit M = 128

// Backward weight images, per layer s>=1, per which (0: lin_r, 1: lin_l), per 16-wide chunk q of the OUT channel c
// (the contraction index of dH = dZ * W), hi half then lo half; inside a half the K-major SWIZZLE_NONE form
//   B[n = k (in channel)][kk = c]:  float = ((c%16)/4)*(Hp*4) + (k/8)*32 + (k%8)*4 + (c%4)
inline int64_t tc_imgB_floats(const Dims& d) { return (int64_t)(d.Lt - 1) * 2 * d.Hp * d.Hp * 2; }

// ---- the CSR gather of both fused kernels -----------------------------------------------------------------------
// agg[r] = sum over the CSR row r of src[col], r < n, for a [128 + 1][P] row-major shared tile whose row 128 is zeros.
// The kernels are issue / shared-bandwidth bound, so the gather runs off a per-tile TABLE built once and replayed by
// every layer.  A group = 4 accumulator slots walked in lock step; step k of a group is ONE int4 of byte offsets
// (the k-th neighbour of each slot; slots that have run out point at the zero row).  Normal groups hold 4 rows; a LONG
// row (degree > kGatherLong: the two terminal nodes of every Hex graph reach 20-30) gets a group of its own whose 4
// slots are 4 contiguous segments of its neighbour list, summed ((s0+s1)+s2)+s3 at the end -- otherwise one warp
// would walk 25 steps while fifteen wait at the barrier.  Groups go round-robin over the 16 epilogue warps.
// Per step: one broadcast 16-byte table load, four contiguous-row float4 loads (lane = one float4: no bank conflicts
// whatever the ids), eight packed FADD2.  Summation order is fixed (deterministic).
constexpr int kGatherMaxGroups = 48;                  // <= 16 long rows + 32 normal groups
constexpr int kGatherLong = 8;
constexpr int kGatherMaxLong = 16;
struct GatherTab {
  int4* tab;        // [<= kEdgeCapPerRow * kTcRows] steps
  int* gstart;      // [kGatherMaxGroups]
  int* glen;        // [kGatherMaxGroups]
  int4* grow;       // [kGatherMaxGroups] destination row per slot (-1 none; y == -2: long row x in 4 segments)
  int* ngroups;     // [1]
};
__host__ __device__ constexpr int gather_tab_bytes() {
  return kEdgeCapPerRow * kTcRows * 16 + kGatherMaxGroups * (4 + 4 + 16) + 16;
}
__device__ __forceinline__ GatherTab gather_tab_view(uint8_t* base) {
  GatherTab g;
  g.tab = (int4*)base;
  g.grow = (int4*)(base + kEdgeCapPerRow * kTcRows * 16);
  g.gstart = (int*)(g.grow + kGatherMaxGroups);
  g.glen = g.gstart + kGatherMaxGroups;
  g.ngroups = g.glen + kGatherMaxGroups;
  return g;
}
__device__ __forceinline__ int warp_scan_incl(int v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  return v;
}
// Build the table of one tile (all epilogue threads; rowptr_s / col_s complete; ends with the epilogue barrier).
template <int NJ>
__device__ __forceinline__ void gather_build(const GatherTab& g, const int* __restrict__ rowptr_s,
                                             const unsigned short* __restrict__ col_s, int n, int tid) {
  constexpr int PB = (16 * NJ + 4) * 4;   // row pitch in bytes
  const int wid = tid >> 5, lane = tid & 31;
  for (int i = tid; i < kGatherMaxGroups; i += kTcEpiThreads) { g.glen[i] = 0; g.grow[i] = make_int4(-1, -1, -1, -1); }
  asm volatile("bar.sync 1, %0;" ::"n"(kTcEpiThreads) : "memory");
  if (wid == 0) {   // lane = rows 4*lane .. 4*lane+3: classify, then place every row into (group, slot)
    int deg[4], cl = 0, cn = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = 4 * lane + j;
      deg[j] = r < n ? rowptr_s[r + 1] - rowptr_s[r] : -1;
      cl += deg[j] > kGatherLong;
      cn += deg[j] >= 0;
    }
    const int totl = warp_scan_incl(cl, lane);
    const int L_all = __shfl_sync(0xffffffffu, totl, 31);
    const int thr = L_all <= kGatherMaxLong ? kGatherLong : 0x7fffffff;   // too many long rows: split none
    cl = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) cl += deg[j] > thr;
    cn -= cl;
    const int v = (cl << 16) | cn;
    const int inc = warp_scan_incl(v, lane);
    const int tot = __shfl_sync(0xffffffffu, inc, 31);
    const int L = tot >> 16, NN = tot & 0xffff;
    int li = (inc - v) >> 16, ni = (inc - v) & 0xffff;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = 4 * lane + j, d = deg[j];
      if (d < 0) continue;
      if (d > thr) {
        g.grow[li] = make_int4(r, -2, -2, -2);
        g.glen[li] = (d + 3) >> 2;
        ++li;
      } else {
        const int grp = L + (ni >> 2);
        ((int*)&g.grow[grp])[ni & 3] = r;
        atomicMax(&g.glen[grp], d);
        ++ni;
      }
    }
    if (lane == 0) *g.ngroups = L + ((NN + 3) >> 2);
    __syncwarp();
    // exclusive scan of the group lengths (48 groups: lanes take group l and group 32 + l)
    const int l0 = g.glen[lane], l1 = lane < kGatherMaxGroups - 32 ? g.glen[32 + lane] : 0;
    const int i0 = warp_scan_incl(l0, lane), i1 = warp_scan_incl(l1, lane);
    const int t0 = __shfl_sync(0xffffffffu, i0, 31);
    g.gstart[lane] = i0 - l0;
    if (lane < kGatherMaxGroups - 32) g.gstart[32 + lane] = t0 + i1 - l1;
  }
  asm volatile("bar.sync 1, %0;" ::"n"(kTcEpiThreads) : "memory");
  const int ng = *g.ngroups;
  for (int gi = wid; gi < ng; gi += kTcEpiWarps) {
    const int len = g.glen[gi], st = g.gstart[gi];
    const int4 rw = g.grow[gi];
    int b[4], d[4];
    if (rw.y == -2) {   // long row: slot i = neighbours [i*len, (i+1)*len)
      const int b0 = rowptr_s[rw.x], d0 = rowptr_s[rw.x + 1] - b0;
#pragma unroll
      for (int i = 0; i < 4; ++i) { b[i] = b0 + i * len; d[i] = d0 - i * len; d[i] = d[i] > len ? len : d[i]; }
    } else {
      const int rr[4] = {rw.x, rw.y, rw.z, rw.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        b[i] = rr[i] >= 0 ? rowptr_s[rr[i]] : 0;
        d[i] = rr[i] >= 0 ? rowptr_s[rr[i] + 1] - b[i] : 0;
      }
    }
    for (int k = lane; k < len; k += 32) {
      int4 o;
      o.x = (k < d[0] ? (int)col_s[b[0] + k] : kTcRows) * PB;
      o.y = (k < d[1] ? (int)col_s[b[1] + k] : kTcRows) * PB;
      o.z = (k < d[2] ? (int)col_s[b[2] + k] : kTcRows) * PB;
      o.w = (k < d[3] ? (int)col_s[b[3] + k] : kTcRows) * PB;
      g.tab[st + k] = o;
    }
  }
  asm volatile("bar.sync 1, %0;" ::"n"(kTcEpiThreads) : "memory");
}
__device__ __forceinline__ void fadd2x2(float4& a, const float4& b) {   // a += b as two packed f32x2 adds
  unsigned long long a0, a1, b0, b1;
  asm("mov.b64 %0, {%1, %2};" : "=l"(a0) : "f"(a.x), "f"(a.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(a1) : "f"(a.z), "f"(a.w));
  asm("mov.b64 %0, {%1, %2};" : "=l"(b0) : "f"(b.x), "f"(b.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(b1) : "f"(b.z), "f"(b.w));
  asm("add.rn.f32x2 %0, %0, %1;" : "+l"(a0) : "l"(b0));
  asm("add.rn.f32x2 %0, %0, %1;" : "+l"(a1) : "l"(b1));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a.x), "=f"(a.y) : "l"(a0));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a.z), "=f"(a.w) : "l"(a1));
}
template <int NJ>
__device__ __forceinline__ void tile_gather(const GatherTab& g, const float* __restrict__ src, float* __restrict__ agg,
                                            int n, int wid, int lane) {
  constexpr int Hp = 16 * NJ, PV = (Hp + 4) / 4, NV = Hp / 4;
  if (lane >= NV) return;                       // warp-uniform loops below: no collective inside
  const char* sb = (const char*)src + lane * 16;
  float4* a4 = (float4*)agg + lane;
  const int ng = *g.ngroups;
  for (int gi = wid; gi < ng; gi += kTcEpiWarps) {
    const int len = g.glen[gi];
    const int4* tp = g.tab + g.gstart[gi];
    const int4 rw = g.grow[gi];
    float4 acc[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    int k = 0;
    for (; k + 2 <= len; k += 2) {
      const int4 o0 = tp[k], o1 = tp[k + 1];
      const float4 v0 = *(const float4*)(sb + o0.x), v1 = *(const float4*)(sb + o0.y);
      const float4 v2 = *(const float4*)(sb + o0.z), v3 = *(const float4*)(sb + o0.w);
      const float4 u0 = *(const float4*)(sb + o1.x), u1 = *(const float4*)(sb + o1.y);
      const float4 u2 = *(const float4*)(sb + o1.z), u3 = *(const float4*)(sb + o1.w);
      fadd2x2(acc[0], v0); fadd2x2(acc[1], v1); fadd2x2(acc[2], v2); fadd2x2(acc[3], v3);
      fadd2x2(acc[0], u0); fadd2x2(acc[1], u1); fadd2x2(acc[2], u2); fadd2x2(acc[3], u3);
    }
    if (k < len) {
      const int4 o0 = tp[k];
      const float4 v0 = *(const float4*)(sb + o0.x), v1 = *(const float4*)(sb + o0.y);
      const float4 v2 = *(const float4*)(sb + o0.z), v3 = *(const float4*)(sb + o0.w);
      fadd2x2(acc[0], v0); fadd2x2(acc[1], v1); fadd2x2(acc[2], v2); fadd2x2(acc[3], v3);
    }
    if (rw.y == -2) {
      fadd2x2(acc[0], acc[1]); fadd2x2(acc[0], acc[2]); fadd2x2(acc[0], acc[3]);
      a4[rw.x * PV] = acc[0];
    } else {
      if (rw.x >= 0) a4[rw.x * PV] = acc[0];
      if (rw.y >= 0) a4[rw.y * PV] = acc[1];
      if (rw.z >= 0) a4[rw.z * PV] = acc[2];
      if (rw.w >= 0) a4[rw.w * PV] = acc[3];
    }
  }
}

int launch_pack_tc(const Dims& d, const ParamMap& pm, const float* params, const float* x, TcPackView pv,
                   bool with_transpose, cudaStream_t st, int* zero_ints = nullptr, int n_zero = 0);
int launch_forward_tc(const Dims& d, const ParamMap& pm, const float* params, const float* x, int64_t N, int64_t E,
                      int64_t B, const void* csr, int flags, float* out_nodes, float* out_value, float* stash,
                      TcPackView pv, cudaStream_t st, int* tile_flag = nullptr);

int debug_read_fwd_phases(long long* host, int n);
int debug_read_bwd_phases(long long* host, int n);
int tc_dw_groups(const Dims& d);
struct TdFuse;
// Persistent-CTA schedule of the tcgen05 forward / chain kernels: tiles are visited in csr.tile_order (heaviest first).
// Stand-alone launch (zig = true): CTA c takes position rnd * G + c in even rounds and rnd * G + (G - 1 - c) in odd rounds
// (boustrophedon: the CTA with the heaviest tile of one round gets the lightest of the next -- measured forward 0.292 ->
// 0.285 ms).  Inside hexgnn_train_step (zig = false) the kernels are chained by programmatic dependent launch and the NEXT
// kernel's CTAs become resident in blockIdx order as this one's retire, so the low CTAs, which start first, keep the
// second round (boustrophedon there: trio 0.597 -> 0.642 ms).
__host__ __device__ inline int tile_position(int rnd, int cta, int G, bool zig) {
  return rnd * G + ((zig && (rnd & 1)) ? (G - 1 - cta) : cta);
}
int launch_backward_tc_chain(const Dims& d, const ParamMap& pm, const float* params, const float* x, int64_t N, int64_t E,
                             int64_t B, const void* csr, int flags, const float* d_nodes, const float* d_value,
                             const float* stash, float* dz, float* d_embeds, TcPackView pv, float* small_rec,
                             float* head_rec, int* sync, int grid, cudaStream_t st, const TdFuse* fuse = nullptr,
                             float* d_x = nullptr);
int launch_backward_tc_dw(const Dims& d, int64_t N, int64_t E, int64_t B, const void* csr, const float* stash,
                          const float* dz, float* big_rec, int* sync, cudaStream_t st);
int launch_reduce_tc(const Dims& d, const float* small_rec, const int* ntiles, const float* big_rec, int G,
                     const float* head_rec, int64_t B, float* packed_sum, cudaStream_t st, const TdLossArgs* td = nullptr);

// Per-graph record of the value head's backward: [dt (H/2 padded to 4) | pooled (4H padded to 4)]; the gradient of
// value_head.layers.0.weight is sum_g dt_g (x) pooled_g, summed in graph order by reduce_tc_kernel.
__host__ __device__ inline int tc_head_rec_floats(const Dims& d) { return ((d.H2 + 3) & ~3) + ((4 * d.H + 3) & ~3); }
// Workspace of the tcgen05 backward (floats): weight images | dZ stash | small records | group records | head
// records | packed sum
// `sync` (ints): [0] ticket counter of the dW kernel, [4 .. 4+B) one "dZ of this tile is complete" flag per tile,
// then one "phase-A record written" flag per (layer, group) item of the dW kernel, then [B] "forward of this tile is
// complete" flags (hexgnn_train_step only: the chain starts under the forward's last round); zeroed per call.
constexpr int kSyncTileOff = 4;
inline int64_t tc_sync_fwd_off(const Dims& d, int64_t B) { return kSyncTileOff + B + (int64_t)(d.Lt - 1) * tc_dw_groups(d); }
inline int64_t tc_sync_ints(const Dims& d, int64_t B) { return align_up(tc_sync_fwd_off(d, B) + B, 4); }

// Fused TD loss of hexgnn_train_step: the chain kernel derives dL/dQ itself (td_loss.cu's formula, same fp32 operations).
struct TdFuse {
  const int64_t* action;   // [B] local action per graph; nullptr = dL/dQ comes in as d_nodes
  const float* target;     // [B]
  const float* weight;     // [B] or nullptr
  const float* q;          // [N] forward output
  const int* fwd_flag;     // [tiles] released by the forward kernel, or nullptr
  int kind;
  float scale;
};
struct TcBwdWs { int64_t pack, dz, small_rec, big_rec, head_rec, sum, sync, total; };
inline TcBwdWs tc_bwd_ws(const Dims& d, int64_t B) {
  GradLayout gl;
  grad_layout(d, &gl);
  TcBwdWs w;
  int64_t o = 0;
  w.pack = o; o += align_up(tc_pack_floats(d), 256);
  w.dz = o; o += tc_dz_bytes(d, B) / 4;
  w.small_rec = o; o += align_up(B * (gl.total - gl.gb), 4);          // one record per tile (tiles <= B)
  w.big_rec = o; o += (int64_t)tc_dw_groups(d) * gl.g0;
  w.head_rec = o; o += align_up(B * tc_head_rec_floats(d), 4);
  w.sum = o; o += align_up(gl.total, 4);
  w.sync = o; o += tc_sync_ints(d, B);
  w.total = o;
  return w;
}

}  // namespace hexgnn
