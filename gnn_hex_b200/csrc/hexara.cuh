// HexAra network (hexara.cu forward, hexara_bwd.cu backward): shared shapes, parameter map and stash layout.
#pragma once
#include "lw_model.cuh"

namespace hexgnn {

constexpr int kAraMaxConvs = 96;

struct AraDims {
  int H, Hp, H2, Cin, swap;
  int nl[3];          // layers of the stacks: 0 body, 1 value, 2 policy
  int nconv;
};
struct AraMap {
  int64_t wl[kAraMaxConvs], bl[kAraMaxConvs], wr[kAraMaxConvs];   // conv index = stack-major: body, value, policy
  int cin[kAraMaxConvs], cout[kAraMaxConvs];
  int64_t mlp[2][4];  // [value_linear | swap_linear][layers.0.weight, layers.0.bias, layers.1.weight, layers.1.bias]
  int64_t total;
};

inline bool make_ara_dims(const HexAraDesc* s, AraDims* d) {
  if (!s || s->hidden < 2 || s->hidden_layers < 1 || s->policy_layers < 1 || s->value_layers < 1 || s->in_channels != 3)
    return false;
  d->H = s->hidden;
  d->Hp = lw_pad16(s->hidden);
  d->H2 = s->hidden / 2;
  d->Cin = s->in_channels;
  d->swap = s->swap_allowed ? 1 : 0;
  d->nl[0] = s->hidden_layers; d->nl[1] = s->value_layers; d->nl[2] = s->policy_layers;
  d->nconv = d->nl[0] + d->nl[1] + d->nl[2];
  return d->nconv <= kAraMaxConvs && d->H2 >= 1;
}

inline int conv_index(const AraDims& d, int stack, int i) { return (stack > 0 ? d.nl[0] : 0) + (stack > 1 ? d.nl[1] : 0) + i; }

inline void make_ara_map(const AraDims& d, AraMap* m) {
  int64_t o = 0;
  for (int k = 0; k < 3; ++k)
    for (int i = 0; i < d.nl[k]; ++i) {
      const int c = conv_index(d, k, i);
      m->cin[c] = (k == 0 && i == 0) ? d.Cin : d.H;
      m->cout[c] = (k == 2 && i == d.nl[2] - 1) ? 1 : d.H;
      m->wl[c] = o; o += (int64_t)m->cout[c] * m->cin[c];
      m->bl[c] = o; o += m->cout[c];
      m->wr[c] = o; o += (int64_t)m->cout[c] * m->cin[c];
    }
  for (int j = 0; j < 2; ++j) {
    m->mlp[j][0] = o; o += (int64_t)d.H2 * 4 * d.H;
    m->mlp[j][1] = o; o += d.H2;
    m->mlp[j][2] = o; o += d.H2;
    m->mlp[j][3] = o; o += 1;
  }
  m->total = o;
}


// Training stash (floats): slot c = output of conv c AFTER its ReLU (the last conv of a stack has none), [N][Hp] per slot
// (the 1-channel policy conv uses the first [N][16] of its slot); conv index = stack-major (body, value, policy).
inline int64_t ara_stash_floats(const AraDims& d, int64_t N) { return (int64_t)d.nconv * N * d.Hp; }

int64_t ara_bwd_ws_bytes(const AraDims& d, const AraMap& m, int64_t N, int64_t B);
int launch_hexara_backward(const AraDims& d, const AraMap& pm, const float* params, const float* x, int64_t N, int64_t E,
                           int64_t B, const void* csr, const float* pi, const int64_t* out_ptr, const float* d_pi,
                           const float* d_value, const float* stash, float* grad_params, void* ws, cudaStream_t st);

}  // namespace hexgnn
