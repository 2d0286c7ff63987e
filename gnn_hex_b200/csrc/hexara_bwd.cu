// Backward of the HexAra policy / value network (hexara.cu): replaces `combined_loss.backward()` through SAGE_torch_script
// (rl_loop/trainer_agent_pytorch.py:245-275; the module: GN0/torch_script_models.py:286-379) on the layer-at-a-time kernels.
//   head      hexara_head_bwd_kernel  per graph: log-softmax backward (d logit = d pi - exp(pi) * sum_graph d pi), terminal rows
//                                     get zero, swap entry -> swap MLP; d value -> tanh' -> value MLP; 4-way pooling backward
//                                     (max / min route to the first row attaining the extremum) -> dG of the value stack
//   stacks    lw_sage_kernel (adjoint) value stack, policy stack (its last conv has ONE output channel: K = 16), their
//                                     gradients w.r.t. the body embedding are SUMMED (second writer accumulates), body
//   weights   lw_dw_kernel over a table of (input rows, [dpre | dZl] rows) for every H -> H conv; the 1-channel policy conv and
//             the 3-channel first conv on the CUDA cores; fixed-order reductions into the flat gradient (deterministic)
#include "hexara.cuh"

namespace hexgnn {

namespace {

__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ float wsum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float wmaxf(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// per-graph record (floats): dt_v [H2p] | gv1 [H2p] | dt_s [H2p] | gs1 [H2p] | pooled [4H pad 4] | misc [4] = (dvp, dsl)
struct AraRec { int dtv, gv1, dts, gs1, pooled, misc, total; };
__host__ __device__ inline AraRec ara_rec(const AraDims& d) {
  AraRec r;
  const int H2p = (d.H2 + 3) & ~3;
  r.dtv = 0; r.gv1 = H2p; r.dts = 2 * H2p; r.gs1 = 3 * H2p; r.pooled = 4 * H2p;
  r.misc = r.pooled + ((4 * d.H + 3) & ~3);
  r.total = r.misc + 4;
  return r;
}

constexpr int kHbThreads = 256;
struct AraHeadBwdArgs {
  AraDims d; AraMap pm;
  const float* params; const int* gptr;
  const float* vemb;            // [N][Hp] output of the value stack
  const float* pi; const int64_t* out_ptr; const float* d_pi; const float* d_value;
  float* dpi_raw;               // [N][32]: column 0 = d logit of the row (dpre of the policy stack's last conv), rest zeros
  float* dG; int ldg;           // dpre of the value stack's last conv ([N][ldg], columns [0, Hp))
  unsigned* absmax_pi; unsigned* absmax_v;
  float* rec;
};

__global__ void __launch_bounds__(kHbThreads) hexara_head_bwd_kernel(const __grid_constant__ AraHeadBwdArgs a) {
  extern __shared__ float sm[];   // pool [4H] | hid [2][H2] | dt [2][H2] | dpool [4H] | amax [H] | amin [H] | red [8] | misc [8]
  pdl_wait();
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int H = a.d.H, Hp = a.d.Hp, H2 = a.d.H2;
  const AraRec R = ara_rec(a.d);
  const int g = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nw = kHbThreads / 32;
  float* rec = a.rec + (int64_t)g * R.total;
  float* pool = sm;
  float* hid = pool + 4 * H;
  float* dt = hid + 2 * H2;
  float* dpool = dt + 2 * H2;
  int* amax = (int*)(dpool + 4 * H);
  int* amin = amax + H;
  float* red = (float*)(amin + H);
  float* misc = red + 8;
  const int r0 = a.gptr[g], n = a.gptr[g + 1] - r0;
  const int64_t out0 = a.out_ptr[g];
  const int mt = (int)(a.out_ptr[g + 1] - out0);
  const int m = n > 2 ? n - 2 : 0;
  const int my_swap = mt > m ? 1 : 0;
  // ---- log-softmax backward over the graph's policy entries (torch_script_models.py:378) --------------------------------
  float S = 0.f;
  if (a.d_pi) {
    float s = 0.f;
    for (int i = tid; i < mt; i += kHbThreads) s += a.d_pi[out0 + i];
    s = wsum(s);
    if (lane == 0) red[wid] = s;
    __syncthreads();
    for (int w = 0; w < nw; ++w) S += red[w];
  }
  float mx = 0.f;
  for (int i = tid; i < n; i += kHbThreads) {
    float dl = 0.f;
    if (i >= 2 && a.d_pi) dl = a.d_pi[out0 + i - 2] - expf(a.pi[out0 + i - 2]) * S;
    float4* o = (float4*)(a.dpi_raw + (int64_t)(r0 + i) * 32);
    o[0] = make_float4(dl, 0.f, 0.f, 0.f);
#pragma unroll
    for (int q = 1; q < 8; ++q) o[q] = make_float4(0.f, 0.f, 0.f, 0.f);
    mx = fmaxf(mx, fabsf(dl));
  }
  mx = wmaxf(mx);
  if (lane == 0 && mx > 0.f) atomicMax(a.absmax_pi, __float_as_uint(mx));
  const float dsl = (my_swap && a.d_pi) ? a.d_pi[out0 + m] - expf(a.pi[out0 + m]) * S : 0.f;
  // ---- pooling of the value stack's output, with the first row attaining each extremum --------------------------------
  for (int c = tid; c < H; c += kHbThreads) {
    float sum = 0.f, vmax = -INFINITY, vmin = INFINITY;
    int imx = 0, imn = 0;
#pragma unroll 4
    for (int i = 0; i < n; ++i) {
      const float v = a.vemb[(int64_t)(r0 + i) * Hp + c];
      sum += v;
      if (v > vmax) { vmax = v; imx = i; }
      if (v < vmin) { vmin = v; imn = i; }
    }
    if (n <= 0) { vmax = 0.f; vmin = 0.f; imx = imn = -1; }
    pool[c] = sum; pool[H + c] = vmax; pool[2 * H + c] = vmin; pool[3 * H + c] = sum / (float)(n > 0 ? n : 1);
    amax[c] = imx; amin[c] = imn;
  }
  __syncthreads();
  // ---- the two MLPs forward again (Linear -> ReLU -> Linear; value: tanh), then backward -------------------------------
  const int nmlp = my_swap ? 2 : 1;
  for (int j = 0; j < nmlp; ++j) {
    const float* w0 = a.params + a.pm.mlp[j][0];
    const float* b0 = a.params + a.pm.mlp[j][1];
    for (int k = wid; k < H2; k += nw) {
      float u = 0.f;
      const float* wrow = w0 + (int64_t)k * 4 * H;
      for (int q = lane; q < 4 * H; q += 32) u = fmaf(wrow[q], pool[q], u);
      u = wsum(u);
      if (lane == 0) hid[j * H2 + k] = fmaxf(u + b0[k], 0.f);
    }
  }
  __syncthreads();
  if (wid == 0) {
    const float* w1 = a.params + a.pm.mlp[0][2];
    float u = 0.f;
    for (int k = lane; k < H2; k += 32) u = fmaf(w1[k], hid[k], u);
    u = wsum(u) + a.params[a.pm.mlp[0][3]];
    if (lane == 0) {
      const float v = tanhf(u);
      misc[0] = a.d_value ? a.d_value[g] * (1.0f - v * v) : 0.f;
    }
  }
  __syncthreads();
  const float dvp = misc[0];
  for (int k = tid; k < 2 * H2; k += kHbThreads) {
    const int j = k / H2, kk = k - j * H2;
    const float up = j == 0 ? dvp : (my_swap ? dsl : 0.f);
    const bool on = j < nmlp && hid[k] > 0.f;
    dt[k] = on ? up * a.params[a.pm.mlp[j][2] + kk] : 0.f;
    if (j >= nmlp) hid[k] = 0.f;
  }
  __syncthreads();
  for (int q = tid; q < 4 * H; q += kHbThreads) {
    float u = 0.f;
    for (int j = 0; j < nmlp; ++j) {
      const float* w0 = a.params + a.pm.mlp[j][0];
      for (int k = 0; k < H2; ++k) u = fmaf(w0[(int64_t)k * 4 * H + q], dt[j * H2 + k], u);
    }
    dpool[q] = u;
  }
  // ---- record ------------------------------------------------------------------------------------------------------
  for (int k = tid; k < ((H2 + 3) & ~3); k += kHbThreads) {
    const bool ok = k < H2;
    rec[R.dtv + k] = ok ? dt[k] : 0.f;
    rec[R.gv1 + k] = ok ? dvp * hid[k] : 0.f;
    rec[R.dts + k] = ok ? dt[H2 + k] : 0.f;
    rec[R.gs1 + k] = ok ? dsl * hid[H2 + k] : 0.f;
  }
  for (int q = tid; q < ((4 * H + 3) & ~3); q += kHbThreads) rec[R.pooled + q] = q < 4 * H ? pool[q] : 0.f;
  if (tid == 0) { rec[R.misc] = dvp; rec[R.misc + 1] = dsl; rec[R.misc + 2] = 0.f; rec[R.misc + 3] = 0.f; }
  __syncthreads();
  // ---- dG: gradient of the pooled vector back onto the graph's rows (no activation after the stack's last conv) ----------
  float gx = 0.f;
  const float invn = 1.0f / (float)(n > 0 ? n : 1);
  for (int idx = tid; idx < n * Hp; idx += kHbThreads) {
    const int i = idx / Hp, c = idx - i * Hp;
    float v = 0.f;
    if (c < H) {
      v = dpool[c] + dpool[3 * H + c] * invn;
      if (i == amax[c]) v += dpool[H + c];
      if (i == amin[c]) v += dpool[2 * H + c];
    }
    a.dG[(int64_t)(r0 + i) * a.ldg + c] = v;
    gx = fmaxf(gx, fabsf(v));
  }
  gx = wmaxf(gx);
  if (lane == 0 && gx > 0.f) atomicMax(a.absmax_v, __float_as_uint(gx));
}

// ---- the policy stack's last conv (H -> 1): d lin_r.W[0][k] = sum_i dpre_i h_i[k], d lin_l.W[0][k] = sum_j dZl_j h_j[k],
//      d bias = sum_i dpre_i; per-block partials [blocks][3][Hp] ---------------------------------------------------------
constexpr int kDpThreads = 128, kDpRows = 64;
struct AraDwpArgs {
  const float* h; int64_t N; int Hp;
  const float* y;            // [N][32]: column 0 = dpre, column 16 = dZl
  float* part;
};
__global__ void __launch_bounds__(kDpThreads) ara_dwp_kernel(const __grid_constant__ AraDwpArgs a) {
  __shared__ float sy[kDpRows][2];
  pdl_wait();
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int64_t row0 = (int64_t)blockIdx.x * kDpRows;
  const int nrows = (int)((a.N - row0) < kDpRows ? (a.N - row0) : kDpRows);
  if ((int)threadIdx.x < kDpRows) {
    const bool live = (int)threadIdx.x < nrows;
    sy[threadIdx.x][0] = live ? a.y[(row0 + threadIdx.x) * 32] : 0.f;
    sy[threadIdx.x][1] = live ? a.y[(row0 + threadIdx.x) * 32 + 16] : 0.f;
  }
  __syncthreads();
  float* o = a.part + (int64_t)blockIdx.x * 3 * a.Hp;
  for (int k = threadIdx.x; k < a.Hp; k += kDpThreads) {
    float gr_ = 0.f, gl = 0.f;
#pragma unroll 8
    for (int r = 0; r < nrows; ++r) {
      const float hv = a.h[(row0 + r) * a.Hp + k];
      gr_ = fmaf(sy[r][0], hv, gr_);
      gl = fmaf(sy[r][1], hv, gl);
    }
    o[k] = gr_;
    o[a.Hp + k] = gl;
  }
  if (threadIdx.x == 0) {
    float b = 0.f;
    for (int r = 0; r < nrows; ++r) b += sy[r][0];
    o[2 * a.Hp] = b;
  }
}

// ---- reductions into the flat gradient ---------------------------------------------------------------------------------
struct AraReduceArgs {
  AraDims d; AraMap pm;
  float* grad;
  int items[kMaxSage]; int nitems;           // conv index of every H -> H conv that went through lw_dw_kernel
  const float* rec; int G;                   // [nitems][G][(Hp + 1)][2Hp]
  const float* part0; int nblk0;             // first conv: [nblk0][2 Cin + 1][Hp]
  const float* partp; int nblkp; int c_pi;   // 1-channel policy conv: [nblkp][3][Hp]
  const float* hrec; int B;
};
__global__ void ara_reduce_sage_kernel(const __grid_constant__ AraReduceArgs a) {
  pdl_wait();
  const int H = a.d.H, Hp = a.d.Hp, N2 = 2 * Hp;
  const int t = blockIdx.y, c = a.items[t];
  const float* rec = a.rec + (int64_t)t * a.G * (Hp + 1) * N2;
  float* wl = a.grad + a.pm.wl[c];
  float* bl = a.grad + a.pm.bl[c];
  float* wr = a.grad + a.pm.wr[c];
  const int total = (H + 1) * 2 * H;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int k = i / (2 * H), cc = i - k * 2 * H;
    const int which = cc / H, co = cc - which * H;          // 0: dpre columns (lin_r, bias), 1: dZl columns (lin_l)
    const int mrow = k < H ? k : Hp;                        // the ones channel sits at row Hp
    float s = 0.f;
    for (int g = 0; g < a.G; ++g) s += rec[((int64_t)g * (Hp + 1) + mrow) * N2 + which * Hp + co];
    if (k < H) (which ? wl : wr)[(int64_t)co * H + k] = s;
    else if (!which) bl[co] = s;
  }
}
// d W0[k][q] = sum_g dt_g[k] * pooled_g[q], graphs in order, for both MLPs (blockIdx.y: 0 value_linear, 1 swap_linear)
constexpr int kV0Threads = 256, kV0Q = 32, kV0G = 128, kV0K = 64;
__global__ void __launch_bounds__(kV0Threads) ara_reduce_w0_kernel(const __grid_constant__ AraReduceArgs a) {
  __shared__ float sd[kV0G][kV0K];
  __shared__ float sp[kV0G][kV0Q];
  pdl_wait();
  const int H = a.d.H, H2 = a.d.H2, j = blockIdx.y;
  const AraRec R = ara_rec(a.d);
  const int off_dt = j == 0 ? R.dtv : R.dts;
  const int tid = threadIdx.x, ql = tid & 31, kb = tid >> 5, q0 = blockIdx.x * kV0Q;
  float acc[kV0K / 8];
#pragma unroll
  for (int u = 0; u < kV0K / 8; ++u) acc[u] = 0.f;
  for (int g0 = 0; g0 < a.B; g0 += kV0G) {
    __syncthreads();
    for (int i = tid; i < kV0G * kV0K; i += kV0Threads) {
      const int g = i / kV0K, k = i - g * kV0K;
      sd[g][k] = (g0 + g < a.B && k < H2) ? a.hrec[(int64_t)(g0 + g) * R.total + off_dt + k] : 0.f;
    }
    for (int i = tid; i < kV0G * kV0Q; i += kV0Threads) {
      const int g = i / kV0Q, q = q0 + (i - g * kV0Q);
      sp[g][i - g * kV0Q] = (g0 + g < a.B && q < 4 * H) ? a.hrec[(int64_t)(g0 + g) * R.total + R.pooled + q] : 0.f;
    }
    __syncthreads();
    for (int g = 0; g < kV0G; ++g) {
      const float pq = sp[g][ql];
#pragma unroll
      for (int u = 0; u < kV0K / 8; ++u) acc[u] = fmaf(sd[g][kb + 8 * u], pq, acc[u]);
    }
  }
  float* w0 = a.grad + a.pm.mlp[j][0];
  if (q0 + ql < 4 * H) {
#pragma unroll
    for (int u = 0; u < kV0K / 8; ++u)
      if (kb + 8 * u < H2) w0[(int64_t)(kb + 8 * u) * 4 * H + q0 + ql] = acc[u];
  }
}
// one warp per output: lane l adds elements l, l + 32, ... in order, then a butterfly over the lanes (fixed order)
__device__ __forceinline__ float warp_strided_sum(const float* p, int64_t stride, int n, int lane) {
  float t = 0.f;
#pragma unroll 4
  for (int i = lane; i < n; i += 32) t += p[(int64_t)i * stride];
  return wsum(t);
}
__global__ void ara_reduce_small_kernel(const __grid_constant__ AraReduceArgs a) {
  pdl_wait();
  const int H = a.d.H, Hp = a.d.Hp, H2 = a.d.H2, Cin = a.d.Cin;
  const AraRec R = ara_rec(a.d);
  const int lane = threadIdx.x & 31;
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
  // MLP biases and output layers: sums of record entries over the graphs
  for (int i = gw; i < 2 * (2 * H2 + 1); i += nw) {
    const int j = i / (2 * H2 + 1), r = i - j * (2 * H2 + 1);
    int off; float* dst;
    if (r < H2) { off = (j == 0 ? R.dtv : R.dts) + r; dst = a.grad + a.pm.mlp[j][1] + r; }
    else if (r < 2 * H2) { off = (j == 0 ? R.gv1 : R.gs1) + (r - H2); dst = a.grad + a.pm.mlp[j][2] + (r - H2); }
    else { off = R.misc + j; dst = a.grad + a.pm.mlp[j][3]; }
    const float s = warp_strided_sum(a.hrec + off, R.total, a.B, lane);
    if (lane == 0) *dst = s;
  }
  // first conv (Cin -> H)
  for (int i = gw; i < (2 * Cin + 1) * H; i += nw) {
    const int row = i / H, c = i - row * H;
    const float s = warp_strided_sum(a.part0 + (int64_t)row * Hp + c, (int64_t)(2 * Cin + 1) * Hp, a.nblk0, lane);
    if (lane != 0) continue;
    if (row < Cin) a.grad[a.pm.wl[0] + (int64_t)c * Cin + row] = s;
    else if (row == Cin) a.grad[a.pm.bl[0] + c] = s;
    else a.grad[a.pm.wr[0] + (int64_t)c * Cin + (row - Cin - 1)] = s;
  }
  // 1-channel policy conv
  for (int i = gw; i < 2 * H + 1; i += nw) {
    const int row = i / H, k = i - row * H;     // row 0: lin_r.W[0][k], 1: lin_l.W[0][k], 2 (k == 0): bias
    const float s = warp_strided_sum(a.partp + (int64_t)row * Hp + k, (int64_t)3 * Hp, a.nblkp, lane);
    if (lane != 0) continue;
    if (row == 0) a.grad[a.pm.wr[a.c_pi] + k] = s;
    else if (row == 1) a.grad[a.pm.wl[a.c_pi] + k] = s;
    else a.grad[a.pm.bl[a.c_pi]] = s;
  }
}

template <typename K, typename A>
int launch_pdl3(K kern, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const A& args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  HEX_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, args));
  return HEXGNN_OK;
}

inline int conv_k(const AraMap& m, int c, int Hp) { return m.cout[c] == 1 ? 16 : Hp; }
// the conv whose OUTPUT is conv c's input (-1: the network input x)
inline int conv_pred(const AraDims& d, int c) {
  const int c_emb = d.nl[0] - 1;
  if (c < d.nl[0]) return c - 1;
  if (c < d.nl[0] + d.nl[1]) return c == d.nl[0] ? c_emb : c - 1;
  return c == d.nl[0] + d.nl[1] ? c_emb : c - 1;
}
// does conv c's output pass through a ReLU (every conv but the last of its stack)
inline bool conv_relu(const AraDims& d, int c) {
  return !(c == d.nl[0] - 1 || c == d.nl[0] + d.nl[1] - 1 || c == d.nconv - 1);
}

struct AraBwdWs {
  int64_t degf, zero, imgT, y, hrec, rec, part0, partp, total;
  int64_t img_off[kAraMaxConvs], y_off[kAraMaxConvs];
  int nitems, items[kAraMaxConvs];
};
AraBwdWs ara_bwd_ws(const AraDims& d, const AraMap& m, int64_t N, int64_t B) {
  AraBwdWs w;
  int64_t o = 0;
  w.degf = o; o += align_up(2 * N * 4, 256);
  w.zero = o; o += 512;
  w.imgT = o;
  for (int c = 1; c < d.nconv; ++c) { w.img_off[c] = o - w.imgT; o += align_up(lw_img_bytes(conv_k(m, c, d.Hp), d.Hp), 256); }
  w.y = o;
  for (int c = 0; c < d.nconv; ++c) { w.y_off[c] = o - w.y; o += align_up(N * 2 * conv_k(m, c, d.Hp) * 4, 256); }
  w.hrec = o; o += align_up(B * (int64_t)ara_rec(d).total * 4, 256);
  w.nitems = 0;
  for (int c = 1; c < d.nconv; ++c) if (m.cout[c] != 1) w.items[w.nitems++] = c;
  w.rec = o; o += align_up(lw_dw_rec_floats(d.Hp, w.nitems) * 4, 256);
  w.part0 = o; o += align_up((int64_t)lw_dw0_blocks(N) * (2 * d.Cin + 1) * d.Hp * 4, 256);
  w.partp = o; o += align_up((int64_t)((N + kDpRows - 1) / kDpRows) * 3 * d.Hp * 4, 256);
  w.total = o;
  return w;
}

}  // namespace

int64_t ara_bwd_ws_bytes(const AraDims& d, const AraMap& m, int64_t N, int64_t B) { return ara_bwd_ws(d, m, N, B).total; }

int launch_hexara_backward(const AraDims& d, const AraMap& pm, const float* params, const float* x, int64_t N, int64_t E,
                           int64_t B, const void* csr, const float* pi, const int64_t* out_ptr, const float* d_pi,
                           const float* d_value, const float* stash, float* grad_params, void* ws, cudaStream_t st) {
  const CsrView cv = csr_view(const_cast<void*>(csr), N, E, B);
  const AraBwdWs W = ara_bwd_ws(d, pm, N, B);
  if (W.nitems > kMaxSage) { set_error("hexara_backward: %d H -> H convs (max %d)", W.nitems, kMaxSage); return HEXGNN_E_UNSUPPORTED; }
  uint8_t* wb = (uint8_t*)ws;
  float* degf = (float*)(wb + W.degf);
  unsigned* absmax = (unsigned*)(wb + W.zero);           // slot c: max |dpre of conv c|
  uint8_t* imgT = wb + W.imgT;
  auto Y = [&](int c) { return (float*)(wb + W.y + W.y_off[c]); };
  float* hrec = (float*)(wb + W.hrec);
  float* rec = (float*)(wb + W.rec);
  float* part0 = (float*)(wb + W.part0);
  float* partp = (float*)(wb + W.partp);
  const int H = d.H, Hp = d.Hp;
  auto slot = [&](int c) { return stash + (int64_t)c * N * Hp; };
  const int c_emb = d.nl[0] - 1, c_v0 = d.nl[0], c_vl = d.nl[0] + d.nl[1] - 1, c_p0 = d.nl[0] + d.nl[1], c_pi = d.nconv - 1;

  HEX_CUDA_CHECK(cudaMemsetAsync(grad_params, 0, (size_t)pm.total * 4, st));
  // ---- transposed weight images (d input = [dpre | dZl] * [W_r ; W_l]) ---------------------------------------------------
  {
    LwPackArgs pa;
    pa.n = 0; pa.params = params; pa.x = nullptr; pa.img = imgT; pa.bias_out = nullptr;
    for (int c = 1; c < d.nconv; ++c) {
      if (pa.n == kLwMaxPack) { int rc = launch_lw_pack(pa, st); if (rc) return rc; pa.n = 0; }
      LwPackEntry& e = pa.e[pa.n++];
      for (int v = 0; v < 2; ++v) { e.w_self[v] = pm.wr[c]; e.w_agg[v] = pm.wl[c]; e.bias[v] = -1; }
      e.ld = H; e.transposed = 1; e.K = pm.cout[c]; e.Kp = conv_k(pm, c, Hp); e.Nn = H; e.Np = Hp;
      e.img_off = W.img_off[c];
      e.bias_off = 0;
    }
    int rc = launch_lw_pack(pa, st);
    if (rc) return rc;
  }
  int rc = launch_lw_prepare(cv.rowptr, N, degf, absmax, 128, st);
  if (rc) return rc;
  // ---- head ----------------------------------------------------------------------------------------------------------
  {
    AraHeadBwdArgs ha = {};
    ha.d = d; ha.pm = pm; ha.params = params; ha.gptr = cv.gptr; ha.vemb = slot(c_vl);
    ha.pi = pi; ha.out_ptr = out_ptr; ha.d_pi = d_pi; ha.d_value = d_value;
    ha.dpi_raw = Y(c_pi); ha.dG = Y(c_vl); ha.ldg = 2 * Hp;
    ha.absmax_pi = absmax + c_pi; ha.absmax_v = absmax + c_vl;
    ha.rec = hrec;
    const size_t smem = (size_t)(10 * H + 4 * d.H2 + 16) * 4;
    rc = launch_pdl3(hexara_head_bwd_kernel, dim3((unsigned)B), dim3(kHbThreads), smem, st, ha);
    if (rc) return rc;
  }
  // ---- adjoint of conv c: [dpre_c | sum_i dpre_i / deg_i] * [W_r ; W_l] -> d (input of c), masked by the ReLU of the conv that
  //      produced that input, written (or added) into that conv's dpre columns -------------------------------------------
  auto adjoint = [&](int c, bool accum, bool track_max) {
    const int p = conv_pred(d, c), Kc = conv_k(pm, c, Hp);
    LwSageArgs a = {};
    a.x = Y(c); a.ldx = 2 * Kc; a.N = N; a.Kp = Kc; a.Np = Hp; a.Nout = H;
    a.rowptr = cv.rowptr; a.col = cv.col; a.degf = degf; a.rdeg = degf + N; a.adjoint = 1;
    a.wimg = imgT + W.img_off[c];
    a.in_absmax = absmax + c;
    a.agg_out = Y(c) + Kc; a.lda = 2 * Kc;
    a.out = Y(p); a.ldo = 2 * Hp;
    if (conv_relu(d, p)) { a.mask_src = slot(p); a.ldm = Hp; }
    a.out_absmax = track_max ? absmax + p : nullptr;
    a.accum = accum ? 1 : 0;
    a.status = cv.status;
    return launch_lw_sage(a, st);
  };
  // the aggregated columns of convs that are never an adjoint source stay unused; conv 0 needs only its dpre columns
  for (int c = c_vl; c >= c_v0; --c) { rc = adjoint(c, false, c != c_v0); if (rc) return rc; }      // value stack -> d embeds
  for (int c = c_pi; c >= c_p0; --c) { rc = adjoint(c, c == c_p0, true); if (rc) return rc; }       // policy stack: += d embeds
  for (int c = c_emb; c >= 1; --c) { rc = adjoint(c, false, true); if (rc) return rc; }              // body
  // ---- weight gradients ------------------------------------------------------------------------------------------------
  {
    const float* xs[kMaxSage]; const float* ys[kMaxSage];
    for (int t = 0; t < W.nitems; ++t) { xs[t] = slot(conv_pred(d, W.items[t])); ys[t] = Y(W.items[t]); }
    rc = launch_lw_dw_tab(xs, ys, W.nitems, N, Hp, rec, st);
    if (rc) return rc;
  }
  const int nblkp = (int)((N + kDpRows - 1) / kDpRows);
  {
    AraDwpArgs pa;
    pa.h = slot(conv_pred(d, c_pi)); pa.N = N; pa.Hp = Hp; pa.y = Y(c_pi); pa.part = partp;
    rc = launch_pdl3(ara_dwp_kernel, dim3((unsigned)nblkp), dim3(kDpThreads), 0, st, pa);
    if (rc) return rc;
  }
  rc = launch_lw_dw0(x, 3, d.Cin, N, cv.rowptr, cv.col, degf, Y(0), 2 * Hp, Hp, part0, st);
  if (rc) return rc;
  AraReduceArgs ra = {};
  ra.d = d; ra.pm = pm; ra.grad = grad_params; ra.nitems = W.nitems;
  for (int t = 0; t < W.nitems; ++t) ra.items[t] = W.items[t];
  ra.rec = rec; ra.G = lw_dw_groups(W.nitems);
  ra.part0 = part0; ra.nblk0 = lw_dw0_blocks(N); ra.partp = partp; ra.nblkp = nblkp; ra.c_pi = c_pi;
  ra.hrec = hrec; ra.B = (int)B;
  if (W.nitems > 0) {
    rc = launch_pdl3(ara_reduce_sage_kernel, dim3(32, (unsigned)W.nitems), dim3(256), 0, st, ra);
    if (rc) return rc;
  }
  if (d.H2 > kV0K) { set_error("hexara_backward: hidden=%d not supported by the MLP reduction", H); return HEXGNN_E_UNSUPPORTED; }
  rc = launch_pdl3(ara_reduce_w0_kernel, dim3((unsigned)((4 * H + kV0Q - 1) / kV0Q), 2), dim3(kV0Threads), 0, st, ra);
  if (rc) return rc;
  return launch_pdl3(ara_reduce_small_kernel, dim3((unsigned)bwd_grid_ctas()), dim3(256), 0, st, ra);
}

}  // namespace hexgnn
