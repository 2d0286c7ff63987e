// Device-side CSR builder: edge_index [2,E] int64 (+ batch / ptr) -> destination-sorted CSR + tiles.
//
// Replaces, for the kernels' needs, the host chain graph -> edge_index -> Batch of the reference
// (GN0/util/convert_graph.py:99-117, Batch.from_data_list at Rainbow/common/replay_buffer.py:215-216,
// cpp_hex/hex_graph_game/util.cpp:22-41).  Contract (bit-exact, tests/test_csr.py):
//   rowptr[i]  = #edges with destination < i
//   col[...]   = sources of the edges stably sorted by destination (ties keep input order)
// Integer work, HBM/L2-latency bound: 16 B per edge in, 4 B per edge + 4 B per node out.
#include "common.cuh"

namespace hexgnn {

// ---- 1. init: status, gptr, zero the degree histogram ------------------------------------------
__global__ void csr_init_kernel(CsrView v, const int64_t* __restrict__ batch, const int64_t* __restrict__ ptr,
                                int N, int B) {
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const int nth = gridDim.x * blockDim.x;
  if (tid == 0) { *v.status = 0; *v.ntiles = 0; }
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

// ---- 2. in-degree histogram ---------------------------------------------------------------------
__global__ void csr_hist_kernel(CsrView v, const int64_t* __restrict__ ei, int E, int N) {
  const int nth = gridDim.x * blockDim.x;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < E; e += nth) {
    const long long s = ei[e], d = ei[(int64_t)E + e];
    if (s < 0 || s >= N || d < 0 || d >= N) { atomicOr(v.status, HEXGNN_ST_EDGE_CROSSES); continue; }
    atomicAdd(&v.cursor[d], 1);
  }
}

// ---- 3. exclusive scan of the histogram (single CTA; N is a few 10^4..10^5) ----------------------
// Warp w owns a contiguous range (a multiple of 128 entries); inside it a step is one coalesced int4 per lane.
// Pass 1: every lane sums its int4s of all steps (independent loads, all in flight), warp reduce, the 32 warp totals
// are scanned; pass 2 re-reads (cache hot), does the in-step scan with shuffles and writes rowptr.
__global__ void __launch_bounds__(1024) csr_scan_kernel(CsrView v, int N) {
  __shared__ int warp_sums[32];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int M = N + 1;
  const int per_warp = ((M + 31) / 32 + 127) / 128 * 128;
  const int w0 = wid * per_warp, w1 = (w0 + per_warp < M) ? w0 + per_warp : M;
  const int4* c4 = (const int4*)v.cursor;
  int tsum = 0;
  for (int base = w0; base < w1; base += 128) {
    const int i = base + lane * 4;
    if (i + 3 < M) { const int4 a = c4[i >> 2]; tsum += a.x + a.y + a.z + a.w; }
    else { for (int k = 0; k < 4; ++k) if (i + k < M) tsum += v.cursor[i + k]; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) tsum += __shfl_xor_sync(0xffffffffu, tsum, o);
  if (lane == 0) warp_sums[wid] = tsum;
  __syncthreads();
  if (wid == 0) {
    const int ws = warp_sums[lane];
    int winc = ws;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += t;
    }
    warp_sums[lane] = winc - ws;  // exclusive
  }
  __syncthreads();
  int carry = warp_sums[wid];
  for (int base = w0; base < w1; base += 128) {
    const int i = base + lane * 4;
    int a[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) a[k] = (i + k < M) ? v.cursor[i + k] : 0;
    const int ls = a[0] + a[1] + a[2] + a[3];
    int inc = ls;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    int ex = carry + inc - ls;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (i + k < M) { v.rowptr[i + k] = ex; v.cursor[i + k] = 0; }
      ex += a[k];
    }
    carry += __shfl_sync(0xffffffffu, inc, 31);
  }
}

// ---- 4. unstable fill with edge ids + greedy tile packing (last block) ----------------------------
__global__ void csr_fill_kernel(CsrView v, const int64_t* __restrict__ ei, int E, int N, int B, int tile_rows,
                                int tile_edges) {
  if (blockIdx.x == gridDim.x - 1) {
    // Greedy packing of whole graphs into tiles of <= tile_rows nodes, <= tile_edges edges,
    // <= kMaxTileGraphs graphs.  The scan is sequential by nature, so the block first stages the per-graph
    // node / edge counts in shared memory (parallel, coalesced: the dependent global loads were the whole cost),
    // then one thread walks them.
    constexpr int kChunk = 1024;
    __shared__ int n_s[kChunk], ne_s[kChunk];
    __shared__ int st_s[5];   // t, rows, edges, cnt, status
    if (threadIdx.x == 0) { st_s[0] = 0; st_s[1] = 0; st_s[2] = 0; st_s[3] = 0; st_s[4] = 0; v.tile_g0[0] = 0; }
    for (int gb = 0; gb < B; gb += kChunk) {
      const int gn = (B - gb) < kChunk ? (B - gb) : kChunk;
      __syncthreads();
      for (int i = threadIdx.x; i < gn; i += blockDim.x) {
        const int r0 = v.gptr[gb + i], r1 = v.gptr[gb + i + 1];
        n_s[i] = r1 - r0;
        ne_s[i] = v.rowptr[r1] - v.rowptr[r0];
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        int t = st_s[0], rows = st_s[1], edges = st_s[2], cnt = st_s[3], status = st_s[4];
        for (int i = 0; i < gn; ++i) {
          const int n = n_s[i], ne = ne_s[i];
          if (n > tile_rows || ne > tile_edges) status |= HEXGNN_ST_GRAPH_TOO_LARGE;
          if (cnt > 0 && (rows + n > tile_rows || edges + ne > tile_edges || cnt >= kMaxTileGraphs)) {
            v.tile_g0[++t] = gb + i;
            rows = edges = cnt = 0;
          }
          rows += n; edges += ne; ++cnt;
        }
        st_s[0] = t; st_s[1] = rows; st_s[2] = edges; st_s[3] = cnt; st_s[4] = status;
      }
    }
    if (threadIdx.x == 0) {
      int t = st_s[0];
      if (B > 0) v.tile_g0[++t] = B;
      *v.ntiles = t;
      if (st_s[4]) atomicOr(v.status, st_s[4]);
    }
    return;
  }
  const int nth = (gridDim.x - 1) * blockDim.x;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < E; e += nth) {
    const long long s = ei[e], d = ei[(int64_t)E + e];
    if (s < 0 || s >= N || d < 0 || d >= N) continue;
    const int pos = atomicAdd(&v.cursor[d], 1);
    v.eid[v.rowptr[d] + pos] = e;
  }
}

// ---- 5. stable order inside every row: one thread per CSR slot ------------------------------------
// Slot p holds edge id eid[p] of row d = dst(eid[p]); its final position is the number of edge ids of the same row
// that are smaller (input order == stable sort by destination).  All threads independent: no dependent
// read-modify-write chains in global memory (the per-row insertion sort this replaces was latency bound).
__global__ void csr_rank_kernel(CsrView v, const int64_t* __restrict__ ei, int E, int N) {
  const int nth = gridDim.x * blockDim.x;
  const int filled = v.rowptr[N];   // < E only when edges were rejected (status flagged): those slots hold no id
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < filled; p += nth) {
    const int me = v.eid[p];
    const int d = (int)ei[(int64_t)E + me];
    const int b = v.rowptr[d], e = v.rowptr[d + 1];
    int rank = 0;
    for (int q = b; q < e; ++q) rank += v.eid[q] < me;
    v.col[b + rank] = (int)ei[me];
  }
}

int launch_build_csr(const int64_t* edge_index, int64_t E, const int64_t* batch, const int64_t* ptr, int64_t N,
                     int64_t B, int tile_rows, void* csr, cudaStream_t st) {
  CsrView v = csr_view(csr, N, E, B);
  const int threads = 256;
  auto blocks = [&](int64_t n) { int64_t b = (n + threads - 1) / threads; return (int)(b < 1 ? 1 : (b > 1184 ? 1184 : b)); };
  const int64_t init_n = (N + 1 > B + 1) ? N + 1 : B + 1;
  csr_init_kernel<<<blocks(init_n), threads, 0, st>>>(v, batch, ptr, (int)N, (int)B);
  if (E > 0) csr_hist_kernel<<<blocks(E), threads, 0, st>>>(v, edge_index, (int)E, (int)N);
  csr_scan_kernel<<<1, 1024, 0, st>>>(v, (int)N);
  csr_fill_kernel<<<blocks(E) + 1, threads, 0, st>>>(v, edge_index, (int)E, (int)N, (int)B, tile_rows,
                                                    tile_rows * kEdgeCapPerRow);
  if (E > 0) csr_rank_kernel<<<blocks(E), threads, 0, st>>>(v, edge_index, (int)E, (int)N);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

}  // namespace hexgnn
