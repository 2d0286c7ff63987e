// HexAra policy / value network on the layer-at-a-time tensor-core path.
// Reference: SAGE_torch_script, GN0/torch_script_models.py:286-379 (stacks: ModifiedBaseNet :96-189 -- ReLU after every
// conv except a stack's last, :178-183; ModifiedSAGEConv :52-73; MLP :263-284), built by get_current_model :495-500, called
// from C++ through NN_api::predict (cpp_hex/CrazyAra/nn_api.cpp:14-19) on collate_batch output (hex_graph_game/util.cpp:22-41)
// with the C++ feature convention (shannon_node_switching_game.cpp:690-726: x[:,0] = IS_TERMINAL, x[:,1:3] = swap flags).
//
//   embeds = body(x)                         in_channels (3) -> H, then H -> H; NO ReLU after the last layer
//   value  = tanh(MLP_v([sum|max|min|mean]_g value_stack(embeds)))
//   pi     = log_softmax_g( policy_stack(embeds)[non-terminal rows of g]  [+ MLP_s(pooled_g) if swapping is allowed in g] )
// Launches: weight pack, degrees, layer 0 (CUDA cores), one lw_sage_kernel per further conv (the policy stack's last conv has
// one output channel: the same kernel with N = 16), one per-graph head kernel.  Backward: hexara_bwd.cu.
#include "hexara.cuh"

namespace hexgnn {

namespace {

struct AraWs { int64_t degf, zero, bias, img, act, pi, total; int64_t act_stride; };
AraWs ara_ws(const AraDims& d, const AraMap& m, int64_t N) {
  AraWs w;
  int64_t o = 0;
  w.degf = o; o += align_up(2 * N * 4, 256);   // max(deg, 1) | its reciprocal
  w.zero = o; o += 256;
  w.bias = o; o += align_up((int64_t)d.nconv * d.Hp * 4, 256);
  w.img = o;
  for (int c = 1; c < d.nconv; ++c) o += align_up(lw_img_bytes(d.Hp, m.cout[c] == 1 ? 16 : d.Hp), 256);
  w.act_stride = align_up(N * d.Hp * 4, 256);
  w.act = o; o += 5 * w.act_stride;      // embeds | value ping, pong | policy ping, pong
  w.pi = o; o += align_up(N * 16 * 4, 256);
  w.total = o;
  return w;
}

constexpr int kAhThreads = 256;
struct AraHeadArgs {
  AraDims d; AraMap pm;
  const float* params; const float* x; const int* gptr;
  const float* vemb;        // [N][Hp] output of the value stack
  const float* pi_raw;      // [N][16] output of the policy stack (column 0)
  int B;
  float* pi; float* value; int64_t* out_gi; int64_t* out_ptr;
};

__device__ __forceinline__ float ah_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float ah_warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// One graph per CTA: pooling -> value MLP (+ swap MLP), terminal-row removal, swap entry, log-softmax (torch_scatter
// composite.scatter_log_softmax: x - max - log(sum exp(x - max) + 1e-12)), output graph ids and output ptr
// (torch_script_models.py:320-379).
__global__ void __launch_bounds__(kAhThreads) hexara_head_kernel(const __grid_constant__ AraHeadArgs a) {
  extern __shared__ float sm[];   // pool [4H] | hid [2][H2] | red [8] | misc [8]
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int H = a.d.H, Hp = a.d.Hp, H2 = a.d.H2;
  const int g = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nw = kAhThreads / 32;
  float* pool = sm;
  float* hid = pool + 4 * H;
  float* red = hid + 2 * H2;
  float* misc = red + 8;
  __shared__ int prefix_s;
  const int r0 = a.gptr[g], n = a.gptr[g + 1] - r0;
  // swap entries of the graphs in front of this one -> output offset (:351 cumsum)
  if (tid == 0) prefix_s = 0;
  __syncthreads();
  int my_swap = 0;
  if (a.d.swap) {
    int cnt = 0;
    for (int h = tid; h < g; h += kAhThreads) {
      const int e = a.gptr[h + 1];
      if (e > a.gptr[h] && a.x[(int64_t)(e - 1) * 3 + 2] != 0.f) ++cnt;
    }
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if (lane == 0 && cnt) atomicAdd(&prefix_s, cnt);          // integer: order-independent
    if (n > 0) my_swap = a.x[(int64_t)(r0 + n - 1) * 3 + 2] != 0.f ? 1 : 0;
  }
  // pooling over ALL rows of the graph, terminals included (:320-324)
  for (int c = tid; c < H; c += kAhThreads) {
    float sum = 0.f, mx = -INFINITY, mn = INFINITY;
#pragma unroll 4
    for (int i = 0; i < n; ++i) {
      const float v = a.vemb[(int64_t)(r0 + i) * Hp + c];
      sum += v; mx = fmaxf(mx, v); mn = fminf(mn, v);
    }
    if (n <= 0) { mx = 0.f; mn = 0.f; }
    pool[c] = sum; pool[H + c] = mx; pool[2 * H + c] = mn; pool[3 * H + c] = sum / (float)(n > 0 ? n : 1);
  }
  __syncthreads();
  const int out0 = r0 - 2 * g + prefix_s;
  const int nmlp = my_swap ? 2 : 1;
  for (int j = 0; j < nmlp; ++j) {
    const float* w0 = a.params + a.pm.mlp[j][0];
    const float* b0 = a.params + a.pm.mlp[j][1];
    for (int k = wid; k < H2; k += nw) {
      float u = 0.f;
      const float* wrow = w0 + (int64_t)k * 4 * H;
      for (int q = lane; q < 4 * H; q += 32) u = fmaf(wrow[q], pool[q], u);
      u = ah_warp_sum(u);
      if (lane == 0) hid[j * H2 + k] = fmaxf(u + b0[k], 0.f);
    }
  }
  __syncthreads();
  if (wid < nmlp) {
    const float* w1 = a.params + a.pm.mlp[wid][2];
    float u = 0.f;
    for (int k = lane; k < H2; k += 32) u = fmaf(w1[k], hid[wid * H2 + k], u);
    u = ah_warp_sum(u) + a.params[a.pm.mlp[wid][3]];
    if (lane == 0) misc[wid] = wid == 0 ? tanhf(u) : u;
  }
  __syncthreads();
  if (tid == 0) {
    a.value[g] = misc[0];
    a.out_ptr[g] = out0;
  }
  const int m = n > 2 ? n - 2 : 0;                 // policy entries of the nodes (rows r0+2 ..)
  const int mt = m + my_swap;
  if (g == a.B - 1 && tid == 0) a.out_ptr[a.B] = out0 + mt;
  if (mt == 0) return;
  const float slogit = misc[1];
  // max
  float mx = -INFINITY;
  for (int i = tid; i < m; i += kAhThreads) mx = fmaxf(mx, a.pi_raw[(int64_t)(r0 + 2 + i) * 16]);
  mx = ah_warp_max(mx);
  if (lane == 0) red[wid] = mx;
  __syncthreads();
  mx = red[0];
  for (int w = 1; w < nw; ++w) mx = fmaxf(mx, red[w]);
  if (my_swap) mx = fmaxf(mx, slogit);
  __syncthreads();
  // sum of exp, fixed order
  float s = 0.f;
  for (int i = tid; i < m; i += kAhThreads) s += expf(a.pi_raw[(int64_t)(r0 + 2 + i) * 16] - mx);
  s = ah_warp_sum(s);
  if (lane == 0) red[wid] = s;
  __syncthreads();
  s = 0.f;
  for (int w = 0; w < nw; ++w) s += red[w];
  if (my_swap) s += expf(slogit - mx);
  const float lse = logf(s + 1e-12f);
  for (int i = tid; i < mt; i += kAhThreads) {
    const float v = i < m ? a.pi_raw[(int64_t)(r0 + 2 + i) * 16] : slogit;
    a.pi[out0 + i] = (v - mx) - lse;
    a.out_gi[out0 + i] = g;
  }
}

__global__ void hexara_clear_status_kernel(int* status, int bits) { atomicAnd(status, ~bits); }

}  // namespace

}  // namespace hexgnn

using namespace hexgnn;

extern "C" {

int64_t hexgnn_hexara_param_count(const HexAraDesc* desc) {
  AraDims d;
  if (!make_ara_dims(desc, &d)) return -1;
  AraMap m;
  make_ara_map(d, &m);
  return m.total;
}

int32_t hexgnn_hexara_param_offsets(const HexAraDesc* desc, int64_t* offsets, int32_t max_entries) {
  AraDims d;
  if (!make_ara_dims(desc, &d)) return -1;
  AraMap m;
  make_ara_map(d, &m);
  int32_t n = 0;
  auto put = [&](int64_t v) { if (offsets && n < max_entries) offsets[n] = v; ++n; };
  for (int c = 0; c < d.nconv; ++c) { put(m.wl[c]); put(m.bl[c]); put(m.wr[c]); }
  for (int j = 0; j < 2; ++j) for (int k = 0; k < 4; ++k) put(m.mlp[j][k]);
  return n;
}

int hexgnn_hexara_sizes(const HexAraDesc* desc, int64_t N, int64_t E, int64_t B, HexAraSizes* out) {
  AraDims d;
  if (!make_ara_dims(desc, &d) || !out || N < 0 || E < 0 || B < 0) { set_error("hexara_sizes: bad argument"); return HEXGNN_E_BADARG; }
  if (d.Hp > kLwMaxK) { set_error("hexara_sizes: hidden=%d not supported (<= %d)", d.H, kLwMaxK); return HEXGNN_E_UNSUPPORTED; }
  AraMap m;
  make_ara_map(d, &m);
  out->ws_bytes = ara_ws(d, m, N).total;
  out->stash_bytes = ara_stash_floats(d, N) * 4;
  out->bwd_ws_bytes = (d.Hp <= kLwMaxK && d.nconv - 2 <= kMaxSage) ? ara_bwd_ws_bytes(d, m, N, B) : 0;
  out->max_policy_entries = (N - 2 * B > 0 ? N - 2 * B : 0) + (d.swap ? B : 0);
  out->hidden_padded = d.Hp;
  return HEXGNN_OK;
}

int hexgnn_hexara_forward(const HexAraDesc* desc, const float* params, const float* x, int64_t N, int64_t E, int64_t B,
                          const void* csr, float* pi, float* value, int64_t* out_graph_indices, int64_t* out_batch_ptr,
                          float* embeds, float* stash, void* ws, size_t ws_bytes, void* stream) {
  AraDims d;
  if (!make_ara_dims(desc, &d)) { set_error("hexara_forward: bad model descriptor"); return HEXGNN_E_BADARG; }
  if (!params || !x || !csr || !pi || !value || !out_graph_indices || !out_batch_ptr || !ws || N < 1 || B < 1 || E < 0 ||
      B > 0x7fffffff) { set_error("hexara_forward: bad argument"); return HEXGNN_E_BADARG; }
  if (d.Hp > kLwMaxK) { set_error("hexara_forward: hidden=%d not supported (<= %d)", d.H, kLwMaxK); return HEXGNN_E_UNSUPPORTED; }
  AraMap pm;
  make_ara_map(d, &pm);
  const AraWs W = ara_ws(d, pm, N);
  if ((int64_t)ws_bytes < W.total) { set_error("hexara_forward: workspace too small (%zu < %lld)", ws_bytes, (long long)W.total); return HEXGNN_E_WORKSPACE; }
  cudaStream_t st = (cudaStream_t)stream;
  const CsrView cv = csr_view(const_cast<void*>(csr), N, E, B);
  uint8_t* wb = (uint8_t*)ws;
  float* degf = (float*)(wb + W.degf);
  unsigned* zero = (unsigned*)(wb + W.zero);
  float* bias = (float*)(wb + W.bias);
  uint8_t* img = wb + W.img;
  float* act[5];
  for (int i = 0; i < 5; ++i) act[i] = (float*)(wb + W.act + i * W.act_stride);
  const int H = d.H, Hp = d.Hp;
  // with a stash (training) every conv writes its output into its own slot; without, the buffers ping-pong
  auto slot = [&](int c) { return stash + (int64_t)c * N * Hp; };
  const int c_emb = conv_index(d, 0, d.nl[0] - 1), c_pi = conv_index(d, 2, d.nl[2] - 1);
  float* emb = stash ? slot(c_emb) : (embeds ? embeds : act[0]);
  float* pi_raw = stash ? slot(c_pi) : (float*)(wb + W.pi);

  // ---- weight images of every conv but the first ------------------------------------------------------------------
  int64_t img_off[kAraMaxConvs];
  {
    LwPackArgs pa;
    pa.n = 0; pa.params = params; pa.x = nullptr; pa.img = img; pa.bias_out = bias;
    int64_t o = 0;
    for (int c = 1; c < d.nconv; ++c) {
      if (pa.n == kLwMaxPack) { int rc = launch_lw_pack(pa, st); if (rc) return rc; pa.n = 0; }
      LwPackEntry& e = pa.e[pa.n++];
      for (int v = 0; v < 2; ++v) { e.w_self[v] = pm.wr[c]; e.w_agg[v] = pm.wl[c]; e.bias[v] = pm.bl[c]; }
      e.ld = H; e.transposed = 0; e.K = H; e.Kp = Hp; e.Nn = pm.cout[c]; e.Np = pm.cout[c] == 1 ? 16 : Hp;
      e.img_off = img_off[c] = o;
      e.bias_off = (int64_t)c * Hp;
      o += align_up(lw_img_bytes(Hp, e.Np), 256);
    }
    int rc = launch_lw_pack(pa, st);
    if (rc) return rc;
  }
  int rc = launch_lw_prepare(cv.rowptr, N, degf, zero, 64, st);
  if (rc) return rc;
  hexara_clear_status_kernel<<<1, 1, 0, st>>>(cv.status, HEXGNN_ST_GRAPH_TOO_LARGE | HEXGNN_ST_FLAG_MIXED);   // no tile limit; x[:,2] is a feature here
  HEX_CUDA_CHECK(cudaGetLastError());

  auto sage = [&](int c, const float* in, float* out, int ldo, bool relu) {
    LwSageArgs a = {};
    a.x = in; a.ldx = Hp; a.N = N; a.Kp = Hp; a.Np = pm.cout[c] == 1 ? 16 : Hp; a.Nout = pm.cout[c];
    a.rowptr = cv.rowptr; a.col = cv.col; a.degf = degf; a.rdeg = degf + N; a.adjoint = 0;
    a.wimg = img + img_off[c];
    a.bias = bias + (int64_t)c * Hp;
    a.out = out; a.ldo = ldo; a.relu = relu ? 1 : 0;
    a.status = cv.status;
    return launch_lw_sage(a, st);
  };
  // ---- body -----------------------------------------------------------------------------------------------
  const int L = d.nl[0];
  {
    LwLayer0Args a0;
    a0.x = x; a0.ldx = 3; a0.xoff = 0; a0.Cin = d.Cin; a0.N = N;
    a0.rowptr = cv.rowptr; a0.col = cv.col; a0.degf = degf;
    a0.wl = params + pm.wl[0]; a0.bl = params + pm.bl[0]; a0.wr = params + pm.wr[0];
    a0.H = H; a0.Hp = Hp;
    a0.out = L == 1 ? emb : (stash ? slot(0) : act[1]); a0.ldo = Hp;
    a0.relu = L > 1 ? 1 : 0;
    a0.stats = nullptr;
    rc = launch_lw_layer0(a0, st);
    if (rc) return rc;
  }
  const float* cur = stash ? slot(0) : act[1];
  for (int s = 1; s < L; ++s) {
    float* out = s == L - 1 ? emb : (stash ? slot(s) : act[1 + (s & 1)]);
    rc = sage(conv_index(d, 0, s), cur, out, Hp, s < L - 1);
    if (rc) return rc;
    cur = out;
  }
  // ---- value stack (act[1], act[2]) and policy stack (act[3], act[4]) ------------------------------------------------------
  const float* vout = emb;
  for (int j = 0; j < d.nl[1]; ++j) {
    float* out = stash ? slot(conv_index(d, 1, j)) : act[1 + (j & 1)];
    rc = sage(conv_index(d, 1, j), vout, out, Hp, j < d.nl[1] - 1);
    if (rc) return rc;
    vout = out;
  }
  const float* pin = emb;
  for (int j = 0; j < d.nl[2]; ++j) {
    const bool last = j == d.nl[2] - 1;
    float* out = last ? pi_raw : (stash ? slot(conv_index(d, 2, j)) : act[3 + (j & 1)]);
    rc = sage(conv_index(d, 2, j), pin, out, last ? 16 : Hp, !last);
    if (rc) return rc;
    pin = out;
  }
  // ---- per-graph head -----------------------------------------------------------------------------------------------
  AraHeadArgs ha;
  ha.d = d; ha.pm = pm; ha.params = params; ha.x = x; ha.gptr = cv.gptr; ha.vemb = vout; ha.pi_raw = pi_raw; ha.B = (int)B;
  ha.pi = pi; ha.value = value; ha.out_gi = out_graph_indices; ha.out_ptr = out_batch_ptr;
  hexara_head_kernel<<<(unsigned)B, kAhThreads, (size_t)(4 * H + 2 * d.H2 + 16) * sizeof(float), st>>>(ha);
  HEX_CUDA_CHECK(cudaGetLastError());
  if (stash && embeds) HEX_CUDA_CHECK(cudaMemcpyAsync(embeds, emb, (size_t)N * Hp * 4, cudaMemcpyDeviceToDevice, st));
  return HEXGNN_OK;
}

int hexgnn_hexara_backward(const HexAraDesc* desc, const float* params, const float* x, int64_t N, int64_t E, int64_t B,
                           const void* csr, const float* pi, const int64_t* out_batch_ptr, const float* d_pi,
                           const float* d_value, const float* stash, float* grad_params, void* ws, size_t ws_bytes,
                           void* stream) {
  AraDims d;
  if (!make_ara_dims(desc, &d)) { set_error("hexara_backward: bad model descriptor"); return HEXGNN_E_BADARG; }
  if (!params || !x || !csr || !pi || !out_batch_ptr || !stash || !grad_params || !ws || N < 1 || B < 1 || E < 0 ||
      B > 0x7fffffff || (!d_pi && !d_value)) { set_error("hexara_backward: bad argument"); return HEXGNN_E_BADARG; }
  if (d.Hp > kLwMaxK || d.nconv - 2 > kMaxSage) { set_error("hexara_backward: unsupported shape (hidden %d, %d convs)", d.H, d.nconv); return HEXGNN_E_UNSUPPORTED; }
  AraMap pm;
  make_ara_map(d, &pm);
  const int64_t need = ara_bwd_ws_bytes(d, pm, N, B);
  if ((int64_t)ws_bytes < need) { set_error("hexara_backward: workspace too small (%zu < %lld)", ws_bytes, (long long)need); return HEXGNN_E_WORKSPACE; }
  return launch_hexara_backward(d, pm, params, x, N, E, B, csr, pi, out_batch_ptr, d_pi, d_value, stash, grad_params, ws,
                                (cudaStream_t)stream);
}

}  // extern "C"
