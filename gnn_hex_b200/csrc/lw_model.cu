// modern_two_headed on the layer-at-a-time tensor-core path (lw.cuh): forward orchestration + the dueling head per graph.
// Reference: DuellingTwoHeaded.forward GN0/models.py:537-584 (after_embed_norm :550-551), CachifiedGNN.forward :261-294
// (conv -> [LayerNorm] -> ReLU after EVERY layer, :158-162 appends the norm of the last layer), HeadNetwork.forward :368-384.
#include "lw_model.cuh"

namespace hexgnn {

namespace {

constexpr int kHeadThreads = 256;

__device__ __forceinline__ float hd_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

struct LwHeadArgs {
  Dims d;
  ParamMap pm;
  const float* params;
  const float* x;
  const int* gptr;
  const float* hin;      // [N][Hp] final activations of the taken head
  int flags;
  float* out_nodes;
  float* out_value;
};

// Dueling head of one graph per CTA (any node count): a_i = 2 tanh(w_a . g_i + b_a); pooled = [sum | max | min | mean] of g over
// the graph; v = tanh(W1 relu(W0 pooled + b0) + b1); Q_i = v + a_i - mean(a)   (models.py:368-384, 567-584).
__global__ void __launch_bounds__(kHeadThreads) lw_head_kernel(const __grid_constant__ LwHeadArgs a) {
  extern __shared__ float sm[];   // pool [4H] | hid [H2] | misc [4]
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int r0 = a.gptr[blockIdx.x], n = a.gptr[blockIdx.x + 1] - r0;
  if (n <= 0) {
    if (threadIdx.x == 0 && a.out_value) a.out_value[blockIdx.x] = 0.f;
    return;
  }
  const int H = a.d.H, Hp = a.d.Hp, H2 = a.d.H2;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nw = kHeadThreads / 32;
  const int head = (a.x[2] == 1.0f) ? 0 : 1;
  const float* wa = a.params + a.pm.lin_w[head];
  const float ba = a.params[a.pm.lin_b[head]];
  float* pool = sm;
  float* hid = sm + 4 * H;
  float* misc = hid + H2;
  for (int i = wid; i < n; i += nw) {
    const float* gr = a.hin + (int64_t)(r0 + i) * Hp;
    float u = 0.f;
    for (int c = lane; c < H; c += 32) u = fmaf(wa[c], gr[c], u);
    u = hd_warp_sum(u);
    if (lane == 0) a.out_nodes[r0 + i] = 2.0f * tanhf(u + ba);
  }
  if (a.flags & HEXGNN_ADV_ONLY) return;
  for (int c = tid; c < H; c += kHeadThreads) {
    float sum = 0.f, mx = -INFINITY, mn = INFINITY;
#pragma unroll 4
    for (int i = 0; i < n; ++i) {
      const float v = a.hin[(int64_t)(r0 + i) * Hp + c];
      sum += v; mx = fmaxf(mx, v); mn = fminf(mn, v);
    }
    pool[c] = sum; pool[H + c] = mx; pool[2 * H + c] = mn; pool[3 * H + c] = sum / (float)n;
  }
  __syncthreads();   // also orders this CTA's advantage stores before the reads below
  const float* w0 = a.params + a.pm.v0_w[head];
  const float* b0 = a.params + a.pm.v0_b[head];
  const float* w1 = a.params + a.pm.v1_w[head];
  const float b1 = a.params[a.pm.v1_b[head]];
  for (int k = wid; k < H2; k += nw) {
    float u = 0.f;
    const float* wrow = w0 + (int64_t)k * 4 * H;
    for (int q = lane; q < 4 * H; q += 32) u = fmaf(wrow[q], pool[q], u);
    u = hd_warp_sum(u);
    if (lane == 0) hid[k] = fmaxf(u + b0[k], 0.f);
  }
  __syncthreads();
  if (wid == 0) {
    float u = 0.f;
    for (int k = lane; k < H2; k += 32) u = fmaf(w1[k], hid[k], u);
    u = hd_warp_sum(u);
    if (lane == 0) misc[0] = tanhf(u + b1);
  } else if (wid == 1) {
    float u = 0.f;
    for (int i = lane; i < n; i += 32) u += a.out_nodes[r0 + i];
    u = hd_warp_sum(u);
    if (lane == 0) misc[1] = u / (float)n;
  }
  __syncthreads();
  const float val = misc[0], amean = misc[1];
  for (int i = tid; i < n; i += kHeadThreads) {
    const float adv = a.out_nodes[r0 + i];
    a.out_nodes[r0 + i] = (a.flags & HEXGNN_SEPARATE) ? adv - amean : val + (adv - amean);
  }
  if (tid == 0 && a.out_value) a.out_value[blockIdx.x] = val;
}

__global__ void lw_clear_status_kernel(int* status, int bits) { atomicAnd(status, ~bits); }

}  // namespace

LwModelWs lw_model_ws(const Dims& d, int64_t N, bool backward) {
  LwModelWs w;
  int64_t o = 0;
  w.degf = o; o += align_up(2 * N * 4, 256);   // max(deg, 1) | its reciprocal
  w.zero = o; o += 256;
  w.bias = o; o += align_up((int64_t)d.Lt * d.Hp * 4, 256);
  w.img = o; o += align_up((int64_t)(d.Lt - 1) * lw_img_bytes(d.Hp, d.Hp), 256);
  w.imgT = o; if (backward) o += align_up((int64_t)(d.Lt - 1) * lw_img_bytes(d.Hp, d.Hp), 256);
  w.parts = lw_stat_parts(N);
  w.stats = o; o += align_up((int64_t)(2 * d.Lt + 2) * w.parts * 16, 256);
  w.act = o; o += 3 * align_up(N * d.Hp * 4, 256);
  w.total = o;
  return w;
}

int64_t lw_stash_floats(const Dims& d, int64_t N, bool norm) {
  int64_t f = (int64_t)d.Lt * N * d.Hp;
  if (norm) f += (int64_t)(d.Lt + 1) * N * d.Hp + (int64_t)(2 * d.Lt + 2) * lw_stat_parts(N) * 4;
  return f;
}

int launch_forward_lw(const Dims& d, const ParamMap& pm, const float* params, const float* normp, const float* x, int64_t N,
                      int64_t E, int64_t B, const void* csr, int flags, float* out_nodes, float* out_value, float* stash,
                      void* ws, cudaStream_t st) {
  if (d.Hp > kLwMaxK) { set_error("forward_lw: hidden=%d not supported (<= %d)", d.H, kLwMaxK); return HEXGNN_E_UNSUPPORTED; }
  const CsrView cv = csr_view(const_cast<void*>(csr), N, E, B);
  const bool norm = normp != nullptr;
  const LwModelWs W = lw_model_ws(d, N, false);
  uint8_t* wb = (uint8_t*)ws;
  float* degf = (float*)(wb + W.degf);
  unsigned* zero = (unsigned*)(wb + W.zero);
  float* bias = (float*)(wb + W.bias);
  uint8_t* img = wb + W.img;
  const int H = d.H, Hp = d.Hp, L = d.L, Lt = d.Lt;
  const int64_t NH = N * Hp;
  const int64_t act_stride = align_up(NH * 4, 256) / 4;
  float* act = (float*)(wb + W.act);
  // statistics: with a stash they are kept for the backward (slot s < Lt: pre_s; slot Lt: the embedding before after_embed_norm)
  double* stats = norm && stash ? (double*)(stash + (int64_t)(2 * Lt + 1) * NH) : (double*)(wb + W.stats);
  auto stat = [&](int slot) { return stats + (int64_t)slot * W.parts * 2; };
  auto hbuf = [&](int s) { return stash ? stash + (int64_t)s * NH : act + (int64_t)(s & 1) * act_stride; };
  auto prebuf = [&](int s) { return stash ? stash + (int64_t)(Lt + s) * NH : act + 2 * act_stride; };
  float* emb2 = stash ? stash + (int64_t)2 * Lt * NH : act + (int64_t)(L & 1) * act_stride;   // after_embed_norm output
  auto gamma = [&](int idx, int v) { return normp + ((int64_t)idx + (idx > L ? (int64_t)v * d.Lh : 0)) * 2 * H; };
  // norm index: body layer s -> s; after_embed -> L; head layer s >= L -> L + 1 + (s - L) (+ Lh for the breaker variant)

  // ---- weight images of layers 1..Lt-1 (the taken head's variant is chosen on device) -------------------------
  {
    LwPackArgs pa;
    pa.n = 0; pa.params = params; pa.x = x; pa.img = img; pa.bias_out = bias;
    for (int s = 1; s < Lt; ++s) {
      if (pa.n == kLwMaxPack) { int rc = launch_lw_pack(pa, st); if (rc) return rc; pa.n = 0; }
      LwPackEntry& e = pa.e[pa.n++];
      for (int v = 0; v < 2; ++v) {
        e.w_self[v] = sage_wr(pm, v, s, H);
        e.w_agg[v] = sage_wl(pm, v, s);
        e.bias[v] = sage_bl(pm, v, s, H);
      }
      e.ld = H; e.transposed = 0; e.K = H; e.Kp = Hp; e.Nn = H; e.Np = Hp;
      e.img_off = (int64_t)(s - 1) * lw_img_bytes(Hp, Hp);
      e.bias_off = (int64_t)s * Hp;
    }
    int rc = launch_lw_pack(pa, st);
    if (rc) return rc;
  }
  int rc = launch_lw_prepare(cv.rowptr, N, degf, zero, 64, st);
  if (rc) return rc;
  lw_clear_status_kernel<<<1, 1, 0, st>>>(cv.status, HEXGNN_ST_GRAPH_TOO_LARGE);   // no tile limit on this path

  // ---- layer 0 (2 -> H) --------------------------------------------------------------------------------
  {
    LwLayer0Args a0;
    a0.x = x; a0.ldx = 3; a0.xoff = 0; a0.Cin = 2; a0.N = N;
    a0.rowptr = cv.rowptr; a0.col = cv.col; a0.degf = degf;
    a0.wl = params + sage_wl(pm, 0, 0); a0.bl = params + sage_bl(pm, 0, 0, H); a0.wr = params + sage_wr(pm, 0, 0, H);
    a0.H = H; a0.Hp = Hp;
    a0.out = norm ? prebuf(0) : hbuf(0); a0.ldo = Hp;
    a0.relu = norm ? 0 : 1;
    a0.stats = norm ? stat(0) : nullptr;
    rc = launch_lw_layer0(a0, st);
    if (rc) return rc;
  }
  auto run_norm = [&](const float* in, float* out, int idx, int slot_in, int nparts, int relu, int slot_out) {
    LwNormArgs na;
    na.in = in; na.ldi = Hp; na.out = out; na.ldo = Hp; na.N = N; na.H = H; na.Hp = Hp;
    na.stats_in = stat(slot_in); na.nparts = nparts;
    for (int v = 0; v < 2; ++v) { na.gamma[v] = gamma(idx, v); na.beta[v] = gamma(idx, v) + H; }
    na.x = x; na.relu = relu;
    na.stats_out = slot_out >= 0 ? stat(slot_out) : nullptr;
    na.eps = 1e-5f;
    return launch_lw_norm(na, st);
  };
  if (norm) {
    rc = run_norm(prebuf(0), hbuf(0), 0, 0, lw_layer0_blocks(N), 1, L == 1 ? Lt : -1);
    if (rc) return rc;
    if (L == 1) { rc = run_norm(hbuf(0), emb2, L, Lt, lw_norm_blocks(N), 0, -1); if (rc) return rc; }
  }
  // ---- layers 1 .. Lt-1 -----------------------------------------------------------------------------------
  for (int s = 1; s < Lt; ++s) {
    LwSageArgs a = {};
    a.x = (norm && s == L) ? emb2 : hbuf(s - 1); a.ldx = Hp; a.N = N;
    a.Kp = Hp; a.Np = Hp; a.Nout = H;
    a.rowptr = cv.rowptr; a.col = cv.col; a.degf = degf; a.rdeg = degf + N; a.adjoint = 0;
    a.wimg = img + (int64_t)(s - 1) * lw_img_bytes(Hp, Hp);
    a.bias = bias + (int64_t)s * Hp;
    a.out = norm ? prebuf(s) : hbuf(s); a.ldo = Hp;
    a.relu = norm ? 0 : 1;
    a.stats = norm ? stat(s) : nullptr;
    a.status = cv.status;
    rc = launch_lw_sage(a, st);
    if (rc) return rc;
    if (norm) {
      const int idx = s < L ? s : s + 1;
      rc = run_norm(prebuf(s), hbuf(s), idx, s, lw_sage_blocks(N), 1, s == L - 1 ? Lt : -1);
      if (rc) return rc;
      if (s == L - 1) { rc = run_norm(hbuf(s), emb2, L, Lt, lw_norm_blocks(N), 0, -1); if (rc) return rc; }
    }
  }
  // ---- dueling head ------------------------------------------------------------------------------------------
  {
    LwHeadArgs ha;
    ha.d = d; ha.pm = pm; ha.params = params; ha.x = x; ha.gptr = cv.gptr; ha.hin = hbuf(Lt - 1);
    ha.flags = flags; ha.out_nodes = out_nodes; ha.out_value = out_value;
    lw_head_kernel<<<(unsigned)B, kHeadThreads, (size_t)(4 * H + d.H2 + 4) * sizeof(float), st>>>(ha);
    HEX_CUDA_CHECK(cudaGetLastError());
  }
  return HEXGNN_OK;
}

}  // namespace hexgnn
