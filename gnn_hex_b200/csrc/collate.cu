// Device-side collate of B device-resident position graphs into one batch: the learner's
// `Batch.from_data_list(state)` over replay-buffer entries that already live on the GPU
// (GN0/RainbowDQN/Rainbow/common/replay_buffer.py:116-117 stores them there, :215-216 collates them) and the C++ twin
// `collate_batch` (cpp_hex/hex_graph_game/util.cpp:22-41): x = cat(x_g), edge_index = cat(edge_index_g + node_offset_g, dim 1),
// batch[i] = graph id of node i, ptr = node prefix sums.  The reference issues 2 B tiny cat / index_put kernels from a
// Python (or C++) loop; here the host only writes one table of B (pointer, size, offset) rows, and ONE launch
// (a CTA per graph, grid-stride) copies everything.  Byte/integer work: 12 n + 16 e bytes in, 20 n + 16 e + 8 out per graph.
#include "common.cuh"

namespace hexgnn {

struct CollateRow {          // one graph; filled on the host, copied to the device workspace
  const float* x;            // [n,3]
  const int64_t* ei;         // [2,e] row-major
  int64_t n, e, node_off, edge_off;
};

__global__ void __launch_bounds__(128) collate_kernel(const CollateRow* __restrict__ rows, int B, int64_t N, int64_t E,
                                                      float* __restrict__ x_out, int64_t* __restrict__ ei_out,
                                                      int64_t* __restrict__ batch_out, int64_t* __restrict__ ptr_out) {
  for (int g = blockIdx.x; g < B; g += gridDim.x) {
    const CollateRow r = rows[g];
    const int tid = threadIdx.x;
    float* xo = x_out + 3 * r.node_off;
    for (int64_t i = tid; i < 3 * r.n; i += blockDim.x) xo[i] = r.x[i];
    for (int64_t i = tid; i < r.n; i += blockDim.x) batch_out[r.node_off + i] = g;
    int64_t* s_out = ei_out + r.edge_off;
    int64_t* d_out = ei_out + E + r.edge_off;
    for (int64_t k = tid; k < r.e; k += blockDim.x) {
      s_out[k] = r.ei[k] + r.node_off;
      d_out[k] = r.ei[r.e + k] + r.node_off;
    }
    if (tid == 0) {
      ptr_out[g] = r.node_off;
      if (g == B - 1) ptr_out[B] = N;
    }
  }
}

int64_t collate_ws_bytes(int64_t B) { return B * (int64_t)sizeof(CollateRow); }

int launch_collate(const float* const* x_ptrs, const int64_t* const* ei_ptrs, const int64_t* n_nodes, const int64_t* n_edges,
                   int64_t B, float* x_out, int64_t* ei_out, int64_t* batch_out, int64_t* ptr_out, void* ws, void* host_ws,
                   cudaStream_t st) {
  CollateRow* h = (CollateRow*)host_ws;
  int64_t N = 0, E = 0;
  for (int64_t g = 0; g < B; ++g) {
    h[g].x = x_ptrs[g]; h[g].ei = ei_ptrs[g]; h[g].n = n_nodes[g]; h[g].e = n_edges[g];
    h[g].node_off = N; h[g].edge_off = E;
    N += n_nodes[g]; E += n_edges[g];
  }
  HEX_CUDA_CHECK(cudaMemcpyAsync(ws, h, (size_t)B * sizeof(CollateRow), cudaMemcpyHostToDevice, st));
  const int grid = (int)(B < 1184 ? B : 1184);
  collate_kernel<<<grid, 128, 0, st>>>((const CollateRow*)ws, (int)B, N, E, x_out, ei_out, batch_out, ptr_out);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

}  // namespace hexgnn
