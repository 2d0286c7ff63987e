// Device-side CSR builder: edge_index [2,E] int64 (+ batch / ptr) -> destination-sorted CSR + tiles.
//
// Replaces, for the kernels' needs, the host chain graph -> edge_index -> Batch of the reference
// (GN0/util/convert_graph.py:99-117, Batch.from_data_list at Rainbow/common/replay_buffer.py:215-216,
// cpp_hex/hex_graph_game/util.cpp:22-41).  Contract (bit-exact, tests/test_csr.py):
//   rowptr[i]  = #edges with destination < i
//   col[...]   = sources of the edges stably sorted by destination (ties keep input order)
// Integer work, HBM/L2-latency bound: 16 B per edge in, 4 B per edge + 4 B per node out.
#include <cooperative_groups.h>
#include <stdlib.h>
#include "common.cuh"

namespace hexgnn {
namespace cg = cooperative_groups;

// ---- 1. init: status, gptr, zero the degree histogram ------------------------------------------
__device__ __forceinline__ void csr_init_dev(CsrView v, const int64_t* __restrict__ batch, const int64_t* __restrict__ ptr,
                                             int N, int B, int tid, int nth) {
  if (tid == 0) { *v.status = 0; *v.ntiles = 0; v.scan_ticket[0] = 0; }
  for (int i = tid; i < (int)csr_scan_blocks(N); i += nth) v.scan_state[i] = 0ull;
  for (int i = tid; i <= N; i += nth) v.cursor[i] = 0;  // cursor doubles as the degree histogram
  if (ptr) {
    for (int g = tid; g <= B; g += nth) v.gptr[g] = (int)ptr[g];
  } else if (batch) {
    // batch is sorted: node i opens every graph in (batch[i-1], batch[i]]
    for (int i = tid; i < N; i += nth) {
      const long long b = batch[i];
      const long long pb = i == 0 ? -1 : batch[i - 1];
      if (b < pb || b < 0 || b >= B) { atomicOr(v.status, HEXGNN_ST_BATCH_UNSORTED); continue; }
      for (long long g = pb + 1; g <= b; ++g) v.gptr[g] = i;
      if (i == N - 1) for (long long g = b + 1; g <= B; ++g) v.gptr[g] = N;
    }
    if (N == 0) for (int g = tid; g <= B; g += nth) v.gptr[g] = 0;
  } else {
    for (int g = tid; g <= B; g += nth) v.gptr[g] = g == 0 ? 0 : N;
  }
}

__global__ void csr_init_kernel(CsrView v, const int64_t* __restrict__ batch, const int64_t* __restrict__ ptr,
                                int N, int B) {
  csr_init_dev(v, batch, ptr, N, B, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x);
}

// ---- 2. in-degree histogram ---------------------------------------------------------------------
__device__ __forceinline__ void csr_hist_dev(CsrView v, const int64_t* __restrict__ ei, int E, int N, int tid, int nth) {
  for (int e = tid; e < E; e += nth) {
    const long long s = ei[e], d = ei[(int64_t)E + e];
    if (s < 0 || s >= N || d < 0 || d >= N) { atomicOr(v.status, HEXGNN_ST_EDGE_CROSSES); continue; }
    atomicAdd(&v.cursor[d], 1);
  }
}
__global__ void csr_hist_kernel(CsrView v, const int64_t* __restrict__ ei, int E, int N) {
  csr_hist_dev(v, ei, E, N, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x);
}

// ---- 3. exclusive scan of the histogram: single pass, decoupled look-back ---------------------------------------
// Block (by ticket, so every predecessor is resident or done) = 1024 entries, one int4 per thread.  It publishes its
// aggregate, looks back over the predecessors' states (flag 1 = aggregate, 2 = inclusive prefix) one warp-wide window
// at a time, publishes its inclusive prefix, and writes rowptr.  All blocks run concurrently: two or three L2 round
// trips in total instead of one CTA walking the whole array (16 us for 20 k entries).
__device__ __forceinline__ void csr_scan_dev(CsrView v, int N) {
  __shared__ int warp_sums[8];
  __shared__ int bid_s, prefix_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) bid_s = atomicAdd(v.scan_ticket, 1);
  __syncthreads();
  const int bid = bid_s;
  const int M = N + 1;
  const int i0 = bid * kScanBlockElems + tid * 4;
  int a[4] = {0, 0, 0, 0};
  if (i0 + 3 < M) { const int4 t = *(const int4*)(v.cursor + i0); a[0] = t.x; a[1] = t.y; a[2] = t.z; a[3] = t.w; }
  else { for (int k = 0; k < 4; ++k) if (i0 + k < M) a[k] = v.cursor[i0 + k]; }
  const int ls = a[0] + a[1] + a[2] + a[3];
  int inc = ls;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) warp_sums[wid] = inc;
  __syncthreads();
  int wbase = 0, agg = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) { const int t = warp_sums[w]; if (w < wid) wbase += t; agg += t; }
  if (wid == 0) {
    volatile unsigned long long* st = v.scan_state;
    int prefix = 0;
    if (bid > 0) {
      if (lane == 0) { st[bid] = (1ull << 32) | (unsigned)agg; }
      int hi = bid - 1;                                   // look-back window [hi-31, hi]
      while (true) {
        const int j = hi - lane;
        unsigned long long sv = 2ull << 32;               // out-of-range lanes: a zero prefix
        if (j >= 0) { do { sv = st[j]; } while ((sv >> 32) == 0); }
        const unsigned full = __ballot_sync(0xffffffffu, (sv >> 32) == 2);
        const int stop = full ? (__ffs(full) - 1) : 32;   // nearest predecessor that already has its prefix
        int val = (lane <= stop) ? (int)(unsigned)sv : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
        prefix += val;
        if (full) break;
        hi -= 32;
      }
    }
    if (lane == 0) {
      __threadfence();
      st[bid] = (2ull << 32) | (unsigned)(prefix + agg);
      prefix_s = prefix;
    }
  }
  __syncthreads();
  int ex = prefix_s + wbase + inc - ls;
  if (i0 + 3 < M) {
    int4 r; r.x = ex; r.y = ex + a[0]; r.z = r.y + a[1]; r.w = r.z + a[2];
    *(int4*)(v.rowptr + i0) = r;
    *(int4*)(v.cursor + i0) = make_int4(0, 0, 0, 0);
  } else {
    for (int k = 0; k < 4; ++k) {
      if (i0 + k < M) { v.rowptr[i0 + k] = ex; v.cursor[i0 + k] = 0; }
      ex += a[k];
    }
  }
}

__global__ void __launch_bounds__(256) csr_scan_kernel(CsrView v, int N) { csr_scan_dev(v, N); }

// ---- 4. unstable fill with edge ids + greedy tile packing (last block) ----------------------------
__device__ __forceinline__ void csr_tiles_dev(CsrView v, int B, int tile_rows_packed, int tile_edges) {
  // low 16 bits: capacity of a tile (MMA M = 128); high 16 bits: rows up to which further graphs are PACKED into an open tile
  const int tile_rows = tile_rows_packed & 0xffff, pack_rows = (tile_rows_packed >> 16) ? (tile_rows_packed >> 16) : tile_rows;
  {
    // Greedy packing of whole graphs into tiles of <= tile_rows nodes, <= tile_edges edges, <= kMaxTileGraphs graphs:
    // tile starts s_0 = 0, s_{k+1} = next[s_k], where next[i] = first graph that no longer fits a tile opened at graph i
    // (graph i itself is always taken, also when it is too large on its own).  next[] is computed by all threads in
    // parallel from staged per-graph counts (a tile holds at most kMaxTileGraphs graphs, so the look-ahead is bounded);
    // one thread then only chases the pointers (one shared-memory load per TILE instead of a dependent
    // compare-and-add chain per GRAPH).
    constexpr int kChunk = 1024, kAhead = kMaxTileGraphs;
    __shared__ int n_s[kChunk + kAhead], ne_s[kChunk + kAhead], next_s[kChunk];
    __shared__ int hop2_s[kChunk], hop3_s[kChunk], hop4_s[kChunk];   // tile start 2 / 3 / 4 tiles on (-1: beyond this chunk)
    __shared__ int st_s[2];   // next tile start, tiles so far
    if (threadIdx.x == 0) { st_s[0] = 0; st_s[1] = 0; v.tile_g0[0] = 0; }
    int status = 0;
    for (int gb = 0; gb < B; gb += kChunk) {
      const int gn = (B - gb) < kChunk ? (B - gb) : kChunk;
      const int gl = (B - gb) < kChunk + kAhead ? (B - gb) : kChunk + kAhead;   // staged incl. look-ahead
      __syncthreads();
      for (int i = threadIdx.x; i < gl; i += blockDim.x) {
        const int r0 = v.gptr[gb + i], r1 = v.gptr[gb + i + 1];
        n_s[i] = r1 - r0;
        ne_s[i] = v.rowptr[r1] - v.rowptr[r0];
      }
      __syncthreads();
      for (int i = threadIdx.x; i < gn; i += blockDim.x) {
        int rows = n_s[i], edges = ne_s[i], cnt = 1, j = i + 1;
        if (rows > tile_rows || edges > tile_edges) status |= HEXGNN_ST_GRAPH_TOO_LARGE;
        while (j < gl && cnt < kMaxTileGraphs && rows + n_s[j] <= pack_rows && edges + ne_s[j] <= tile_edges) {
          rows += n_s[j]; edges += ne_s[j]; ++cnt; ++j;
        }
        next_s[i] = gb + j;
      }
      __syncthreads();
      // 2-, 3- and 4-tile look-ahead of every graph, in parallel: the serial chase below then advances four tiles per
      // dependent shared-memory round trip instead of one
      for (int i = threadIdx.x; i < gn; i += blockDim.x) {
        const int end = gb + gn;
        const int a = next_s[i];
        const int b = a < end ? next_s[a - gb] : -1;
        const int c = (b >= 0 && b < end) ? next_s[b - gb] : -1;
        const int d = (c >= 0 && c < end) ? next_s[c - gb] : -1;
        hop2_s[i] = b; hop3_s[i] = c; hop4_s[i] = d;
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        int g = st_s[0], t = st_s[1];
        while (g < gb + gn) {
          const int l = g - gb;
          const int a = next_s[l], b = hop2_s[l], c = hop3_s[l], d = hop4_s[l];
          v.tile_g0[++t] = a; g = a;
          if (b >= 0) { v.tile_g0[++t] = b; g = b; }
          if (c >= 0) { v.tile_g0[++t] = c; g = c; }
          if (d >= 0) { v.tile_g0[++t] = d; g = d; }
        }
        st_s[0] = g; st_s[1] = t;
      }
    }
    if (status) atomicOr(v.status, status);
    __syncthreads();
    const int T = st_s[1];
    if (threadIdx.x == 0) *v.ntiles = T;
    // Processing order of the tiles: heaviest first (cost ~ 8 * rows + edges), so that the LAST round of the persistent
    // kernels -- the one that ends the forward -> chain -> dW pipeline -- gets the lightest tiles.  Deterministic rank
    // sort (ties by tile index): the order decides which tiles a CTA accumulates, i.e. the gradient's summation order.
    constexpr int kMaxSorted = 2048;
    __shared__ int key_s[kMaxSorted];
    if (T <= kMaxSorted) {
      for (int t = threadIdx.x; t < T; t += blockDim.x) {
        const int r0 = v.gptr[v.tile_g0[t]], r1 = v.gptr[v.tile_g0[t + 1]];
        key_s[t] = 8 * (r1 - r0) + (v.rowptr[r1] - v.rowptr[r0]);
      }
      __syncthreads();
      for (int t = threadIdx.x; t < T; t += blockDim.x) {
        const int k = key_s[t];
        int rank = 0;
        for (int j = 0; j < T; ++j) { const int kj = key_s[j]; rank += (kj > k) || (kj == k && j < t); }
        v.tile_order[rank] = t;
      }
    } else {
      for (int t = threadIdx.x; t < T; t += blockDim.x) v.tile_order[t] = t;
    }
  }
}

__device__ __forceinline__ void csr_fill_dev(CsrView v, const int64_t* __restrict__ ei, int E, int N, int tid, int nth) {
  for (int e = tid; e < E; e += nth) {
    const long long s = ei[e], d = ei[(int64_t)E + e];
    if (s < 0 || s >= N || d < 0 || d >= N) continue;
    const int pos = atomicAdd(&v.cursor[d], 1);
    v.eid[v.rowptr[d] + pos] = e;
  }
}

__global__ void csr_fill_kernel(CsrView v, const int64_t* __restrict__ ei, int E, int N, int B, int tile_rows,
                                int tile_edges) {
  if (blockIdx.x == gridDim.x - 1) { csr_tiles_dev(v, B, tile_rows, tile_edges); return; }
  csr_fill_dev(v, ei, E, N, blockIdx.x * blockDim.x + threadIdx.x, (gridDim.x - 1) * blockDim.x);
}

// ---- 5. stable order inside every row: one thread per CSR slot ------------------------------------
// Slot p holds edge id eid[p] of row d = dst(eid[p]); its final position is the number of edge ids of the same row
// that are smaller (input order == stable sort by destination).  All threads independent: no dependent
// read-modify-write chains in global memory (the per-row insertion sort this replaces was latency bound).
__device__ __forceinline__ void csr_rank_dev(CsrView v, const int64_t* __restrict__ ei, int E, int N, int tid, int nth) {
  const int filled = v.rowptr[N];   // < E only when edges were rejected (status flagged): those slots hold no id
  for (int p = tid; p < filled; p += nth) {
    const int me = v.eid[p];
    const int d = (int)ei[(int64_t)E + me];
    const int b = v.rowptr[d], e = v.rowptr[d + 1];
    int rank = 0;
    for (int q = b; q < e; ++q) rank += v.eid[q] < me;
    v.col[b + rank] = (int)ei[me];
  }
}
__global__ void csr_rank_kernel(CsrView v, const int64_t* __restrict__ ei, int E, int N) {
  csr_rank_dev(v, ei, E, N, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x);
}

// ---- all five phases in ONE cooperative launch (grid-wide barriers instead of kernel boundaries) ----------------------
// The builder is 2.4 MB of integer work: five dependent launches cost more in launch / drain latency than in memory time
// (36 us for 103 k edges).  Here every phase is a grid-stride loop of the same resident grid; the tile packing (one block,
// the longest dependent chain) runs in the LAST phase beside the row ranking, which does not need it.
__global__ void __launch_bounds__(256) csr_build_fused_kernel(CsrView v, const int64_t* __restrict__ ei,
                                                              const int64_t* __restrict__ batch, const int64_t* __restrict__ ptr,
                                                              int N, int E, int B, int tile_rows, int tile_edges, int nscan) {
  cg::grid_group grid = cg::this_grid();
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
  csr_init_dev(v, batch, ptr, N, B, tid, nth);
  grid.sync();
  csr_hist_dev(v, ei, E, N, tid, nth);
  grid.sync();
  if ((int)blockIdx.x < nscan) csr_scan_dev(v, N);     // ticket-ordered look-back: every scan block is resident
  grid.sync();
  // rowptr is final: the last block packs and ranks the tiles (one block, the longest dependent chain of the build) while
  // the others fill and rank the rows.  Those G-1 blocks meet at their own barrier -- the scan's ticket counter, which
  // stands at nscan now, counts them in -- so the tile block is not waited for before the kernel ends.
  if (blockIdx.x == gridDim.x - 1) { csr_tiles_dev(v, B, tile_rows, tile_edges); return; }
  const int nth1 = (gridDim.x - 1) * blockDim.x;
  csr_fill_dev(v, ei, E, N, tid, nth1);
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(v.scan_ticket, 1);
    const int target = nscan + (int)gridDim.x - 1;
    while (*(volatile int*)v.scan_ticket < target) {}
    __threadfence();
  }
  __syncthreads();
  csr_rank_dev(v, ei, E, N, tid, nth1);
}

int launch_build_csr(const int64_t* edge_index, int64_t E, const int64_t* batch, const int64_t* ptr, int64_t N,
                     int64_t B, int tile_rows, void* csr, cudaStream_t st) {
  CsrView v = csr_view(csr, N, E, B);
  const int threads = 256;
  auto blocks = [&](int64_t n) { int64_t b = (n + threads - 1) / threads; return (int)(b < 1 ? 1 : (b > 1184 ? 1184 : b)); };
  const int64_t init_n = (N + 1 > B + 1) ? N + 1 : B + 1;
  // Rows up to which further graphs are packed into an open tile (<= tile_rows).  A batch with fewer than one full round of
  // 128-row tiles is packed LESS tightly, so that its tiles spread over all SMs (a tile's cost grows with its rows, and an
  // SM without a tile does nothing): about 1.25 N / #SMs rows, in steps of 16, at least 64.  Measured on GNN-S, 256 7x7
  // positions (8,139 rows): 0.229 ms (128: 66 tiles) -> 0.202 ms (80).  Batches of a full round or more keep 128 (fewer,
  // fuller tiles are cheaper there: 0.614 ms at 128, 0.649 at 96, 0.677 at 64 on the headline batch).  HEXGNN_PACK_ROWS pins it.
  static int pack_env = -1;
  if (pack_env < 0) {
    pack_env = 0;
    if (const char* e = getenv("HEXGNN_PACK_ROWS")) { const int c = atoi(e); if (c >= 1 && c <= tile_rows) pack_env = c; }
  }
  int pack_rows = pack_env;
  if (pack_rows == 0) {
    const int64_t sms = bwd_grid_ctas();
    pack_rows = tile_rows;
    if (N < (int64_t)tile_rows * sms) {
      int64_t p = (5 * N + 4 * sms - 1) / (4 * sms);
      p = (p + 15) / 16 * 16;
      pack_rows = (int)(p < 64 ? 64 : (p > tile_rows ? tile_rows : p));
    }
  }
  int trp = tile_rows | (pack_rows << 16);
  {
    // one cooperative launch when the whole grid can be resident (always, short of ~300 k nodes per call)
    static int coop_max = -1;     // resident 256-thread blocks of the fused kernel on this device, 0 = unsupported
    if (coop_max < 0) {
      int dev = 0, coop = 0, sms = 0, per_sm = 0;
      if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev) == cudaSuccess &&
          coop && cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess &&
          cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, csr_build_fused_kernel, threads, 0) == cudaSuccess)
      {
        int cap = 2;                                       // two blocks per SM: more only lengthens the barriers
        if (const char* e = getenv("HEXGNN_CSR_BLOCKS_PER_SM")) { const int c = atoi(e); if (c >= 0 && c <= 8) cap = c; }   // 0: five-launch path only
        coop_max = sms * (per_sm > cap ? cap : per_sm);
      }
      else
        coop_max = 0;
      (void)cudaGetLastError();
    }
    const int nscan = (int)csr_scan_blocks(N);
    int64_t want = (E > init_n ? E : init_n);
    want = (want + threads - 1) / threads + 1;
    if (want < nscan + 1) want = nscan + 1;
    int grid = (int)(want < coop_max ? want : coop_max);
    if (grid < 2) grid = 2;
    if (coop_max >= 2 && nscan < grid && grid <= coop_max) {
      int Ni = (int)N, Ei = (int)E, Bi = (int)B, te = tile_rows * kEdgeCapPerRow, ns = nscan;
      void* args[] = {&v, (void*)&edge_index, (void*)&batch, (void*)&ptr, &Ni, &Ei, &Bi, &trp, &te, &ns};
      if (cudaLaunchCooperativeKernel((const void*)csr_build_fused_kernel, dim3(grid), dim3(threads), args, 0, st) == cudaSuccess)
        return HEXGNN_OK;
      (void)cudaGetLastError();   // refused (e.g. a context that cannot guarantee co-residency): use the five-launch path
      coop_max = 0;               // from now on
    }
  }
  csr_init_kernel<<<blocks(init_n), threads, 0, st>>>(v, batch, ptr, (int)N, (int)B);
  if (E > 0) csr_hist_kernel<<<blocks(E), threads, 0, st>>>(v, edge_index, (int)E, (int)N);
  csr_scan_kernel<<<(int)csr_scan_blocks(N), 256, 0, st>>>(v, (int)N);
  csr_fill_kernel<<<blocks(E) + 1, threads, 0, st>>>(v, edge_index, (int)E, (int)N, (int)B, trp,
                                                    tile_rows * kEdgeCapPerRow);
  if (E > 0) csr_rank_kernel<<<blocks(E), threads, 0, st>>>(v, edge_index, (int)E, (int)N);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

}  // namespace hexgnn
