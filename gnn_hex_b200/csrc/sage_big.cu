// Oversize graphs (more than one 128-row tile holds: 12x12 / 13x13 boards of the board-size-transfer evaluation,
// GN0/RainbowDQN/evaluate_elo.py:229-251, reach 146 / 171 nodes): inference-only CUDA-core path, one launch per SAGE
// layer, activations ping-pong in a caller-provided workspace.  It runs AFTER the fused forward on the same stream and
// overwrites the outputs of exactly the graphs the fused kernel refused (n > 128 rows or > 1024 edges); every other
// graph exits at once.  Same math as sage_fwd_*.cu (GN0/models.py:261-294, 368-384, 537-584), fp32, deterministic.
#include "common.cuh"

namespace hexgnn {

constexpr int kBigRows = 32;      // node rows per CTA
constexpr int kBigThreads = 256;

struct BigArgs {
  Dims d;
  ParamMap pm;
  const float* params;
  const float* x;
  CsrView csr;
  int B, max_rows, tile_rows, s, flags;
  const float* hin;
  float* hout;
  float* out_nodes;
  float* out_value;
  const float* wcat;   // pack.cu: [Lt-1][Hp][2Hp], wcat[k][c] = lin_l.W[c][k], wcat[k][Hp + c] = lin_r.W[c][k]
  const float* bias;   // [Lt][Hp]
};

__device__ __forceinline__ bool big_graph(const BigArgs& a, int g, int* r0, int* n) {
  *r0 = a.csr.gptr[g];
  *n = a.csr.gptr[g + 1] - *r0;
  const int ne = a.csr.rowptr[*r0 + *n] - a.csr.rowptr[*r0];
  return *n > a.tile_rows || ne > a.tile_rows * kEdgeCapPerRow;
}

__global__ void big_begin_kernel(CsrView v) { atomicAnd(v.status, ~HEXGNN_ST_GRAPH_TOO_LARGE); }

// One SAGE layer for rows [32*blockIdx.y, +32) of graph blockIdx.x: out = relu(W_l mean_j(h_j) + b + W_r h_i).
__global__ void __launch_bounds__(kBigThreads) big_layer_kernel(const __grid_constant__ BigArgs a) {
  extern __shared__ __align__(16) float A[];   // [kBigRows][2*Cp]: mean | self, Cp = Cin rounded up to 4 (zero padded)
  int r0, n;
  if (!big_graph(a, blockIdx.x, &r0, &n)) return;
  if (n > a.max_rows) {
    if (blockIdx.y == 0 && threadIdx.x == 0) atomicOr(a.csr.status, HEXGNN_ST_GRAPH_TOO_LARGE);
    return;
  }
  const int i0 = blockIdx.y * kBigRows;
  if (i0 >= n) return;
  const int rows = (n - i0) < kBigRows ? (n - i0) : kBigRows;
  const int H = a.d.H, Hp = a.d.Hp, s = a.s;
  const int Cin = s == 0 ? 2 : H;
  const int Cp = (Cin + 3) & ~3;
  const int head = (a.x[2] == 1.0f) ? 0 : 1;
  for (int idx = threadIdx.x; idx < kBigRows * Cp; idx += kBigThreads) {
    const int i = idx / Cp, k = idx - i * Cp;
    float mean = 0.f, self = 0.f;
    if (i < rows && k < Cin) {
      const int row = r0 + i0 + i;
      const int b = a.csr.rowptr[row], e = a.csr.rowptr[row + 1];
      float sum = 0.f;
      for (int p = b; p < e; ++p) {
        const int j = a.csr.col[p];
        sum += s == 0 ? a.x[(int64_t)j * 3 + k] : a.hin[(int64_t)j * Hp + k];
      }
      mean = sum / (float)(e - b > 0 ? e - b : 1);
      self = s == 0 ? a.x[(int64_t)row * 3 + k] : a.hin[(int64_t)row * Hp + k];
    }
    A[i * 2 * Cp + k] = mean;
    A[i * 2 * Cp + Cp + k] = self;
  }
  __syncthreads();
  if (s == 0) {
    // 2 input channels: read the parameters directly
    const float* wl = a.params + sage_wl(a.pm, head, s);
    const float* bl = a.params + sage_bl(a.pm, head, s, H);
    const float* wr = a.params + sage_wr(a.pm, head, s, H);
    for (int idx = threadIdx.x; idx < rows * Hp; idx += kBigThreads) {
      const int i = idx / Hp, c = idx - i * Hp;
      float acc = 0.f;
      if (c < H) {
        const float* ar = A + i * 2 * Cp;
        acc = bl[c];
        for (int k = 0; k < Cin; ++k) acc = fmaf(wl[c * Cin + k], ar[k], acc);
        for (int k = 0; k < Cin; ++k) acc = fmaf(wr[c * Cin + k], ar[Cp + k], acc);
        acc = fmaxf(acc, 0.f);
      }
      a.hout[(int64_t)(r0 + i0 + i) * Hp + c] = acc;
    }
    return;
  }
  // H -> H layers: thread = (output channel c, half of the rows); the k-major packed weights wcat[k][c | Hp + c]
  // (pack.cu) make the weight loads coalesced, the activations come as float4 broadcasts from shared memory.
  // Summation order per output: bias, then k ascending with the lin_l and lin_r terms interleaved.
  constexpr int R = kBigRows / 2;
  const int c = threadIdx.x & 127, rg = threadIdx.x >> 7;
  if (c >= Hp) return;
  const float* wcat = a.wcat + (int64_t)(s - 1) * Hp * 2 * Hp;
  float acc[R];
  const float b0 = a.bias[s * Hp + c];
#pragma unroll
  for (int r = 0; r < R; ++r) acc[r] = b0;
  const float* Ar = A + (rg * R) * 2 * Cp;
  for (int k = 0; k < Cp; k += 4) {
    float wl[4], wr[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) { wl[u] = wcat[(int64_t)(k + u) * 2 * Hp + c]; wr[u] = wcat[(int64_t)(k + u) * 2 * Hp + Hp + c]; }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float4 m = *(const float4*)(Ar + r * 2 * Cp + k);
      const float4 h = *(const float4*)(Ar + r * 2 * Cp + Cp + k);
      float v = acc[r];
      v = fmaf(wl[0], m.x, v); v = fmaf(wr[0], h.x, v);
      v = fmaf(wl[1], m.y, v); v = fmaf(wr[1], h.y, v);
      v = fmaf(wl[2], m.z, v); v = fmaf(wr[2], h.z, v);
      v = fmaf(wl[3], m.w, v); v = fmaf(wr[3], h.w, v);
      acc[r] = v;
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int i = rg * R + r;
    if (i < rows) a.hout[(int64_t)(r0 + i0 + i) * Hp + c] = c < H ? fmaxf(acc[r], 0.f) : 0.f;
  }
}

__device__ __forceinline__ float big_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Dueling head of one oversize graph (one CTA): hin = final activations g.
__global__ void __launch_bounds__(kBigThreads) big_head_kernel(const __grid_constant__ BigArgs a) {
  extern __shared__ float sm[];   // pool [4H] | hid [H2] | misc [4]
  int r0, n;
  if (!big_graph(a, blockIdx.x, &r0, &n) || n > a.max_rows) return;
  const int H = a.d.H, Hp = a.d.Hp, H2 = a.d.H2;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nw = kBigThreads / 32;
  const int head = (a.x[2] == 1.0f) ? 0 : 1;
  const float* wa = a.params + a.pm.lin_w[head];
  const float ba = a.params[a.pm.lin_b[head]];
  float* pool = sm;
  float* hid = sm + 4 * H;
  float* misc = hid + H2;
  for (int i = wid; i < n; i += nw) {   // a_i = 2 tanh(w_a . g_i + b_a)
    const float* gr = a.hin + (int64_t)(r0 + i) * Hp;
    float u = 0.f;
    for (int c = lane; c < H; c += 32) u = fmaf(wa[c], gr[c], u);
    u = big_warp_sum(u);
    if (lane == 0) a.out_nodes[r0 + i] = 2.0f * tanhf(u + ba);
  }
  if (a.flags & HEXGNN_ADV_ONLY) return;
  for (int c = tid; c < H; c += kBigThreads) {
    float sum = 0.f, mx = -INFINITY, mn = INFINITY;
    for (int i = 0; i < n; ++i) {
      const float v = a.hin[(int64_t)(r0 + i) * Hp + c];
      sum += v; mx = fmaxf(mx, v); mn = fminf(mn, v);
    }
    pool[c] = sum; pool[H + c] = mx; pool[2 * H + c] = mn; pool[3 * H + c] = sum / (float)n;
  }
  __syncthreads();   // also orders the advantage stores of this CTA before the reads below
  const float* w0 = a.params + a.pm.v0_w[head];
  const float* b0 = a.params + a.pm.v0_b[head];
  const float* w1 = a.params + a.pm.v1_w[head];
  const float b1 = a.params[a.pm.v1_b[head]];
  for (int k = wid; k < H2; k += nw) {
    float u = 0.f;
    const float* wrow = w0 + (int64_t)k * 4 * H;
    for (int q = lane; q < 4 * H; q += 32) u = fmaf(wrow[q], pool[q], u);
    u = big_warp_sum(u);
    if (lane == 0) hid[k] = fmaxf(u + b0[k], 0.f);
  }
  __syncthreads();
  if (wid == 0) {
    float u = 0.f;
    for (int k = lane; k < H2; k += 32) u = fmaf(w1[k], hid[k], u);
    u = big_warp_sum(u);
    if (lane == 0) misc[0] = tanhf(u + b1);
  } else if (wid == 1) {
    float u = 0.f;
    for (int i = lane; i < n; i += 32) u += a.out_nodes[r0 + i];
    u = big_warp_sum(u);
    if (lane == 0) misc[1] = u / (float)n;
  }
  __syncthreads();
  const float val = misc[0], amean = misc[1];
  for (int i = tid; i < n; i += kBigThreads) {
    const float adv = a.out_nodes[r0 + i];
    a.out_nodes[r0 + i] = (a.flags & HEXGNN_SEPARATE) ? adv - amean : val + (adv - amean);
  }
  if (tid == 0 && a.out_value) a.out_value[blockIdx.x] = val;
}

int launch_forward_big(const Dims& d, const ParamMap& pm, const float* params, const float* x, int64_t N, int64_t E,
                       int64_t B, const void* csr, int tile_rows, int flags, int max_rows, float* out_nodes,
                       float* out_value, float* ws, cudaStream_t st) {
  BigArgs a;
  a.d = d; a.pm = pm; a.params = params; a.x = x;
  a.csr = csr_view(const_cast<void*>(csr), N, E, B);
  a.B = (int)B; a.max_rows = max_rows; a.tile_rows = tile_rows; a.flags = flags;
  a.out_nodes = out_nodes; a.out_value = out_value;
  float* buf[2] = {ws, ws + N * d.Hp};
  PackView pv = pack_view(ws + 2 * N * d.Hp, d);
  int rc = launch_pack(d, pm, params, x, pv, false, st);
  if (rc) return rc;
  a.wcat = pv.wcat; a.bias = pv.bias;
  big_begin_kernel<<<1, 1, 0, st>>>(a.csr);
  const dim3 grid((unsigned)B, (unsigned)((max_rows + kBigRows - 1) / kBigRows));
  for (int s = 0; s < d.Lt; ++s) {
    a.s = s;
    a.hin = s == 0 ? nullptr : buf[(s - 1) & 1];
    a.hout = buf[s & 1];
    const size_t smem = (size_t)kBigRows * 2 * (s == 0 ? 4 : ((d.H + 3) & ~3)) * sizeof(float);
    big_layer_kernel<<<grid, kBigThreads, smem, st>>>(a);
  }
  a.hin = buf[(d.Lt - 1) & 1];
  a.hout = nullptr;
  big_head_kernel<<<(unsigned)B, kBigThreads, (size_t)(4 * d.H + d.H2 + 4) * sizeof(float), st>>>(a);
  HEX_CUDA_CHECK(cudaGetLastError());
  return HEXGNN_OK;
}

}  // namespace hexgnn
