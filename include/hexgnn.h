/* hexgnn.h -- C-ABI of libhexgnn.so: B200 (sm_100a) fused forward/backward of GNN_Hex's
 * `modern_two_headed` Q-network over batched Shannon node-switching Hex graphs.
 *
 * The reference has no FFI for this path: its boundary is the Python nn.Module protocol
 *   get_pre_defined("modern_two_headed", args)            GN0/models.py:892,928-947
 *   DuellingTwoHeaded.forward(x, edge_index, graph_indices, ptr, set_cache,
 *                             advantages_only, seperate)  GN0/models.py:537-584
 *   loss.backward() through it                            Rainbow/common/rainbow.py:316
 * and, on the C++ side, collate_batch + NN_api::predict   cpp_hex/hex_graph_game/util.cpp:22-41,
 *                                                         cpp_hex/CrazyAra/nn_api.cpp:14-19.
 * These entry points are what a binding for that path calls (INTEGRATION.md shows the ctypes
 * stub used by gnn_hex_b200/ and the libtorch stub a C++ maintainer would add).
 *
 * Conventions
 *  - plain pointers and sizes; no torch types.  Every pointer except the descriptors and the
 *    host out-parameters is a DEVICE pointer valid on `stream` (a cudaStream_t passed as void*).
 *  - the caller owns every buffer (inputs, outputs, csr blob, stash, workspace); the library
 *    allocates nothing, keeps no mutable global state and is re-entrant.  Work is enqueued on
 *    `stream`; nothing synchronises the host.
 *  - return 0 on success, a negative HEXGNN_E_* otherwise; hexgnn_last_error() gives a
 *    thread-local message.
 *  - data-dependent faults (a graph larger than the tile capacity, an edge crossing graphs)
 *    cannot be seen by the host without a sync: kernels record them in the status word at the
 *    start of the csr blob (see hexgnn_csr_layout) and poison the outputs with NaN.
 */
#ifndef HEXGNN_H_
#define HEXGNN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HEXGNN_ABI_VERSION 11

enum {
  HEXGNN_OK = 0,
  HEXGNN_E_BADARG = -1,      /* null pointer, negative size, inconsistent descriptor            */
  HEXGNN_E_UNSUPPORTED = -2, /* hidden size / tile capacity outside what the kernels were built for */
  HEXGNN_E_WORKSPACE = -3,   /* a caller-provided buffer is smaller than hexgnn_sizes() says      */
  HEXGNN_E_CUDA = -4         /* a CUDA runtime call failed (message has the cudaError string)     */
};

/* flags for hexgnn_forward / hexgnn_backward */
enum {
  HEXGNN_ADV_ONLY = 1,   /* advantages_only=True: out_nodes = 2*tanh(linear(g)), no value branch (models.py:567-569) */
  HEXGNN_SEPARATE = 2,   /* seperate=True: out_value = tanh(value), out_nodes = adv - mean(adv)  (models.py:580-581) */
  HEXGNN_PATH_SIMT = 0x100, /* force the fp32 CUDA-core kernels (default: tcgen05 3xTF32 kernels when the shape allows) */
  HEXGNN_PATH_TC = 0x200,   /* force the tcgen05 kernels; HEXGNN_E_UNSUPPORTED if the shape does not allow      */
  HEXGNN_TF32X1 = 0x400     /* hexgnn_forward without a stash only (inference, opt-in): ONE tf32 tensor-core pass per
                               k-step instead of the three of the fp32-parity 3xTF32 scheme -- ~1e-3 relative error, a third
                               of the tensor-core time.  For self-play acting, where only an arg-max is taken
                               (rainbow.py:214-239; HexAra infers in fp16, rl_loop/rl_config.py:31).  Never for training. */
};

/* device-side status bits (csr blob word 0) */
enum {
  HEXGNN_ST_GRAPH_TOO_LARGE = 1, /* a graph has more nodes/edges than one tile holds   */
  HEXGNN_ST_EDGE_CROSSES = 2,    /* an edge joins two different graphs (or is out of range) */
  HEXGNN_ST_BATCH_UNSORTED = 4,  /* `batch` is not non-decreasing                     */
  HEXGNN_ST_FLAG_MIXED = 8,      /* x[:,2] is not constant (the reference asserts this, models.py:538) */
  HEXGNN_ST_FP16_RANGE = 16      /* an activation reached |h| >= 60000: outside the exact range of the forward kernel's
                                    mixed fp16 / tf32 product scheme (csrc/sage_tc.cuh); use the fp32 path (PATH_SIMT)   */
};

/* Hyper-parameters of modern_two_headed (GN0/models.py:928-947): in_channels is 2. */
typedef struct HexModelDesc {
  int32_t hidden;          /* args.hidden_channels (35 / 110)   */
  int32_t num_layers;      /* args.num_layers (10 / 15)         */
  int32_t num_head_layers; /* args.num_head_layers (default 2)  */
  int32_t in_channels;     /* must be 2                         */
} HexModelDesc;

/* Byte sizes of the caller-allocated buffers for one (N nodes, E directed edges, B graphs) batch. */
typedef struct HexSizes {
  int64_t csr_bytes;     /* csr blob for hexgnn_build_csr                                  */
  int64_t fwd_ws_bytes;  /* workspace of hexgnn_forward                                    */
  int64_t stash_bytes;   /* activation stash written by forward, read by backward           */
  int64_t bwd_ws_bytes;  /* workspace of hexgnn_backward (per-CTA weight-gradient partials) */
  int64_t hidden_padded; /* Hp: row pitch (floats) of stash rows and of d_embeds            */
} HexSizes;

/* Offsets (in int32 words from the blob start) of the arrays inside the csr blob. */
typedef struct HexCsrLayout {
  int64_t status;   /* [1]   status bits (HEXGNN_ST_*)                                   */
  int64_t ntiles;   /* [1]   number of tiles                                             */
  int64_t gptr;     /* [B+1] node prefix sums per graph (== PyG `ptr`, int32)            */
  int64_t rowptr;   /* [N+1] CSR row pointers, rows = destination nodes                  */
  int64_t col;      /* [E]   source node id of each edge, stably sorted by destination  */
  int64_t tile_g0;  /* [B+1] first graph of each tile; tile t = graphs [tile_g0[t], tile_g0[t+1]) */
  int64_t scratch;  /* internal                                                          */
  int64_t total_words;
} HexCsrLayout;

int hexgnn_abi_version(void);
const char* hexgnn_last_error(void);

/* Number of fp32 parameters of the flat parameter buffer (486,974 for GNN-L, 37,382 for GNN-S). */
int64_t hexgnn_param_count(const HexModelDesc* desc);

/* Flat parameter layout.  Writes up to `max_entries` float offsets; returns the tensor count.
 * Order (each tensor row-major as torch stores it, weight[out,in]):
 *   body layer l = 0..L-1 :  convs.l.lin_l.weight, convs.l.lin_l.bias, convs.l.lin_r.weight
 *   maker_head            :  gnn.convs.{0..Lh-1}.(lin_l.weight, lin_l.bias, lin_r.weight),
 *                            linear.weight[1,H], linear.bias[1],
 *                            value_head.layers.0.weight[H/2,4H], .bias[H/2],
 *                            value_head.layers.1.weight[1,H/2],  .bias[1]
 *   breaker_head          :  same as maker_head                                           */
int32_t hexgnn_param_offsets(const HexModelDesc* desc, int64_t* offsets, int32_t max_entries);

/* tile_rows: node capacity of one CTA tile; 128 is the only accepted value (anything else: HEXGNN_E_UNSUPPORTED).
 * Graphs are packed into tiles whole; a tile also holds at most 1024 directed edges and 64 graphs. */
int hexgnn_sizes(const HexModelDesc* desc, int64_t N, int64_t E, int64_t B, int32_t tile_rows,
                 HexSizes* out);
int hexgnn_csr_layout(int64_t N, int64_t E, int64_t B, HexCsrLayout* out);

/* Device CSR builder: replaces the host-side graph -> edge_index -> Batch chain of
 * GN0/util/convert_graph.py:99-117 + Batch.from_data_list (replay_buffer.py:215-216) for the
 * kernels' needs.  edge_index is [2,E] int64 row-major (messages flow row0 -> row1, as PyG);
 * `batch` [N] int64 sorted (PyG Batch.batch) and/or `ptr` [B+1] int64; both NULL => one graph.
 * Result (bit-exact contract): rowptr/col == stable sort of the edges by destination.        */
int hexgnn_build_csr(const int64_t* edge_index, int64_t E, const int64_t* batch, const int64_t* ptr,
                     int64_t N, int64_t B, int32_t tile_rows, void* csr, size_t csr_bytes, void* stream);

/* Fused forward (replaces DuellingTwoHeaded.forward, GN0/models.py:537-584).
 *   params     flat fp32 parameters (layout above)
 *   x          [N,3] fp32: degree, is_terminal, maker flag (convert_graph.py:78-85)
 *   out_nodes  [N]: Q (default) | 2*tanh(adv) (ADV_ONLY) | adv - mean(adv) (SEPARATE)
 *   out_value  [B] tanh(value) (SEPARATE; optional otherwise, may be NULL)
 *   stash      NULL for inference; else stash_bytes, kept for hexgnn_backward.  Layer l's output
 *              h_{l+1} is rows [l*N, (l+1)*N) of pitch Hp floats; the body embedding
 *              (`final_conv_acts`, models.py:553) is layer num_layers-1.                       */
int hexgnn_forward(const HexModelDesc* desc, const float* params, const float* x, int64_t N, int64_t E,
                   int64_t B, const void* csr, int32_t tile_rows, int32_t flags, float* out_nodes,
                   float* out_value, float* stash, void* ws, size_t ws_bytes, void* stream);

/* hexgnn_forward for inference WITHOUT a caller workspace and without any allocation.  No counterpart in the reference:
 * added for callers that cannot call cudaMalloc (code under CUDA-graph capture,
 * tight serving loops, harnesses that forbid allocation).  The only scratch hexgnn_forward needs -- the tensor-core weight
 * images and padded bias rows, rebuilt from `params` on every call -- lives in a file-scope static __device__ array of
 * HEXGNN_STATIC_WS_BYTES inside libhexgnn.so (held by every GPU that loads the library until the process exits).  Same
 * arguments, kernels and results as hexgnn_forward with stash == NULL (tensor-core path; HEXGNN_PATH_SIMT is refused); the
 * caller still passes inputs and outputs.  A model whose images need more than HEXGNN_STATIC_WS_BYTES (GNN-L, hidden 110 /
 * 15 + 2 layers, needs 1.62 MB) is refused with HEXGNN_E_UNSUPPORTED and a message that names both sizes.  Calls no
 * cudaMalloc / cudaFree / allocating library routine (keep hexgnn_profile_enable off: its event brackets are created on
 * demand); the first call on each device looks the array's address up (cudaGetSymbolAddress, that device current) -- make it
 * once outside a capture.  ALL calls on a device share the array: it keeps the last call's images, every byte a call reads
 * is rewritten by that call's own pack kernel (nothing relies on zeros), and two calls that overlap on the GPU (different
 * streams without a dependency) race on it -- serialise them. */
#define HEXGNN_STATIC_WS_BYTES (2 * 1024 * 1024)
int hexgnn_forward_static(const HexModelDesc* desc, const float* params, const float* x, int64_t N, int64_t E, int64_t B,
                          const void* csr, int32_t tile_rows, int32_t flags, float* out_nodes, float* out_value,
                          void* stream);

/* Fused backward (replaces loss.backward() through the model, rainbow.py:316).
 *   d_nodes      [N] gradient w.r.t. out_nodes
 *   d_value      [B] gradient w.r.t. out_value (SEPARATE) or NULL
 *   grad_params  flat fp32, same layout as params; OVERWRITTEN (head not taken gets zeros).
 *                Deterministic: per-CTA partial sums are reduced in a fixed order, no atomics.
 *   d_embeds     optional [N,Hp] gradient w.r.t. the body embedding (`final_conv_grads`,
 *                models.py:534-535), or NULL.                                                  */
int hexgnn_backward(const HexModelDesc* desc, const float* params, const float* x, int64_t N, int64_t E,
                    int64_t B, const void* csr, int32_t tile_rows, int32_t flags, const float* d_nodes,
                    const float* d_value, const float* stash, float* grad_params, float* d_embeds,
                    void* ws, size_t ws_bytes, void* stream);

/* hexgnn_backward that also writes the input gradient d_x [N,3] f32 = dL/dx (saliency maps, rl_loop/explain.py:26-36:
 * `input.grad` after `loss.backward()`); column 2 (the head selector x[:,2], never a network input, models.py:539-540) is
 * zero.  Every row of a graph the kernels processed is written; tensor-core path only (HEXGNN_E_UNSUPPORTED otherwise).
 * All other arguments as in hexgnn_backward. */
int hexgnn_backward_dx(const HexModelDesc* desc, const float* params, const float* x, int64_t N, int64_t E, int64_t B,
                       const void* csr, int32_t tile_rows, int32_t flags, const float* d_nodes, const float* d_value,
                       const float* stash, float* grad_params, float* d_embeds, float* d_x, void* ws, size_t ws_bytes,
                       void* stream);

/* Oversize graphs (board-size transfer, evaluate_elo.py:229-251: 12x12 / 13x13 boards have 146 / 171 nodes).  Call AFTER
 * hexgnn_forward, same stream, same arguments: every graph the fused kernel refused (more than tile_rows nodes or
 * tile_rows*8 edges; its outputs were poisoned with NaN) is evaluated by a CUDA-core path, one launch per SAGE layer,
 * and its rows of out_nodes / out_value are overwritten; HEXGNN_ST_GRAPH_TOO_LARGE is cleared unless a graph exceeds
 * max_rows.  Inference only (no stash).  ws: 2 * N * hidden_padded floats (activations) + fwd_ws_bytes of
 * hexgnn_sizes (k-major packed weights, rebuilt per call). */
int hexgnn_forward_big(const HexModelDesc* desc, const float* params, const float* x, int64_t N, int64_t E, int64_t B,
                       const void* csr, int32_t tile_rows, int32_t flags, int32_t max_rows, float* out_nodes,
                       float* out_value, void* ws, size_t ws_bytes, void* stream);

/* ---- Layer-at-a-time tensor-core path ("lw") ----------------------------------------------------------------------
 * The same network evaluated one SAGE layer per launch over ALL node rows (tcgen05, fp32-parity fp16 split), activations in
 * global memory.  It covers what the fused one-tile-per-CTA kernels cannot hold on chip:
 *   - graphs of any size (12x12 .. 19x19 boards of the transfer evaluation, evaluate_elo.py:228-251, plotting.py:61-80),
 *     forward AND training;
 *   - `--norm=True`: PyG LayerNorm in graph mode without a batch vector (statistics over all nodes and channels of the
 *     batch) after every SAGE layer and once more between body and heads (GN0/models.py:283-287,158-162,481-482,550-551).
 * norm_params == NULL: norm=False.  Otherwise (L + 1 + 2 Lh) blocks of gamma[H] | beta[H]: body layers 0..L-1
 * (`gnn.norms.i`), `after_embed_norm`, maker head layers (`maker_head.gnn.norms.j`), breaker head layers.
 * Everything else (params, x, csr, flags, outputs) as in hexgnn_forward / hexgnn_backward; HEXGNN_ST_GRAPH_TOO_LARGE is
 * cleared (no tile limit here).  stash: stash_bytes of hexgnn_lw_sizes (NULL = inference); the body embedding
 * (`final_conv_acts`, after after_embed_norm when norm is on) is the row-major [N, hidden_padded] block at
 * embed_offset_floats.  grad_norm: same layout as norm_params, OVERWRITTEN.  Deterministic. */
typedef struct HexLwSizes {
  int64_t fwd_ws_bytes, bwd_ws_bytes, stash_bytes;
  int64_t norm_floats;          /* length of norm_params / grad_norm */
  int64_t embed_offset_floats;  /* final_conv_acts inside the stash  */
  int64_t hidden_padded;
} HexLwSizes;
int hexgnn_lw_sizes(const HexModelDesc* desc, int64_t N, int64_t E, int64_t B, int32_t norm, HexLwSizes* out);
int hexgnn_forward_lw(const HexModelDesc* desc, const float* params, const float* norm_params, const float* x, int64_t N,
                      int64_t E, int64_t B, const void* csr, int32_t flags, float* out_nodes, float* out_value, float* stash,
                      void* ws, size_t ws_bytes, void* stream);
int hexgnn_backward_lw(const HexModelDesc* desc, const float* params, const float* norm_params, const float* x, int64_t N,
                       int64_t E, int64_t B, const void* csr, int32_t flags, const float* d_nodes, const float* d_value,
                       const float* stash, float* grad_params, float* grad_norm, float* d_embeds, void* ws, size_t ws_bytes,
                       void* stream);

/* ---- HexAra policy / value network ---------------------------------------------------------------------
 * SAGE_torch_script (GN0/torch_script_models.py:286-379; factory get_current_model :495-500, defaults hidden 60, 15 body
 * layers, 2 policy / 2 value layers), the TorchScript module the C++ engine evaluates through NN_api::predict
 * (cpp_hex/CrazyAra/nn_api.cpp:14-19) on collate_batch output (cpp_hex/hex_graph_game/util.cpp:22-41): inference forward on
 * the layer-at-a-time tensor-core kernels (graphs of any size).
 *   x [N,3] f32 in the C++ feature convention (shannon_node_switching_game.cpp:690-726): IS_TERMINAL, two swap flags -- all
 *   three are network inputs; rows 0, 1 of every graph are its terminal nodes.  csr: hexgnn_build_csr of the collated batch.
 *   params: flat fp32 in state_dict order -- gnn.convs.i.(lin_l.weight, lin_l.bias, lin_r.weight); my_modules.value_head.convs.j.*;
 *   my_modules.policy_head.convs.j.* (the last conv has ONE output channel); my_modules.value_linear.layers.{0,1}.(weight, bias);
 *   my_modules.swap_linear.layers.{0,1}.(weight, bias)   (hexgnn_hexara_param_offsets gives the float offsets).
 * Outputs = the module's 4-tuple (torch_script_models.py:379):
 *   pi [max_policy_entries] f32   per graph: log-softmax over its non-terminal nodes, in node order, followed -- when
 *                                 swap_allowed and x[last node of the graph, 2] != 0 -- by the swap entry (:337-366); the
 *                                 number of entries written is out_batch_ptr[B]  (= N - 2B without swap)
 *   value [B] f32                 tanh(value_linear([sum|max|min|mean] pooling of the value stack))
 *   out_graph_indices [max_policy_entries] i64, out_batch_ptr [B+1] i64
 *   embeds [N, hidden_padded] f32 or NULL: the body output (`final_conv_acts`, :314)
 *   stash: NULL for inference; stash_bytes to keep every conv's output for hexgnn_hexara_backward                    */
typedef struct HexAraDesc {
  int32_t hidden;         /* hidden_channels (60)                        */
  int32_t hidden_layers;  /* body convs (15)                             */
  int32_t policy_layers;  /* policy stack convs (2)                      */
  int32_t value_layers;   /* value stack convs (2)                       */
  int32_t in_channels;    /* must be 3                                   */
  int32_t swap_allowed;   /* SAGE_torch_script(swap_allowed=...)         */
} HexAraDesc;
typedef struct HexAraSizes {
  int64_t ws_bytes;            /* workspace of hexgnn_hexara_forward                     */
  int64_t max_policy_entries;  /* N - 2B (+ B with swap_allowed): size pi / out_graph_indices */
  int64_t hidden_padded;
  int64_t stash_bytes;         /* training: activation stash written by forward, read by backward */
  int64_t bwd_ws_bytes;        /* workspace of hexgnn_hexara_backward                    */
} HexAraSizes;
int64_t hexgnn_hexara_param_count(const HexAraDesc* desc);
int32_t hexgnn_hexara_param_offsets(const HexAraDesc* desc, int64_t* offsets, int32_t max_entries);
int hexgnn_hexara_sizes(const HexAraDesc* desc, int64_t N, int64_t E, int64_t B, HexAraSizes* out);
int hexgnn_hexara_forward(const HexAraDesc* desc, const float* params, const float* x, int64_t N, int64_t E, int64_t B,
                          const void* csr, float* pi, float* value, int64_t* out_graph_indices, int64_t* out_batch_ptr,
                          float* embeds, float* stash, void* ws, size_t ws_bytes, void* stream);
/* Backward of the same network (replaces `combined_loss.backward()` of rl_loop/trainer_agent_pytorch.py:245-275: value MSE
 * + policy cross-entropy on the module's outputs).  stash = the buffer hexgnn_hexara_forward filled (stash != NULL there;
 * stash_bytes), pi / out_batch_ptr = that call's outputs, d_pi [out_batch_ptr[B]] and d_value [B] = the loss gradients w.r.t.
 * pi and value (either may be NULL = zero).  grad_params: flat fp32 in the parameter layout, OVERWRITTEN (parameters the loss
 * does not reach -- swap_linear without a swap entry -- get zeros).  Deterministic.  ws = bwd_ws_bytes. */
int hexgnn_hexara_backward(const HexAraDesc* desc, const float* params, const float* x, int64_t N, int64_t E, int64_t B,
                           const void* csr, const float* pi, const int64_t* out_batch_ptr, const float* d_pi,
                           const float* d_value, const float* stash, float* grad_params, void* ws, size_t ws_bytes,
                           void* stream);

/* One learner step in one call (replaces rainbow.py:301-316: Q = model(states); td_est = Q[action + ptr[:-1]];
 * loss = mean(w * l(td_tgt, td_est)); loss.backward()): forward with stash, TD loss, backward and the deterministic
 * gradient reduction.  Same results as hexgnn_forward + hexgnn_td_loss + hexgnn_backward (bit-identical gradients),
 * but because dL/dQ is derived inside the backward kernel, the three tensor-core kernels are chained by programmatic
 * dependent launch and hand 128-row tiles to each other through release/acquire flags: the dH chain starts on the SMs
 * the forward's last round of tiles leaves idle, the dW GEMM on those the chain leaves idle.
 *   q [N] (out) Q-values; loss [1] (out); td_err [B] (out, or NULL); grad_params (out, OVERWRITTEN); stash = stash_bytes;
 *   ws = bwd_ws_bytes of hexgnn_sizes.  flags must be 0 (default Q output, tensor-core path).
 *   ptr [B+1] i64, action [B] i64, target [B] f32, weight [B] f32 or NULL, kind / scale as in hexgnn_td_loss.          */
int hexgnn_train_step(const HexModelDesc* desc, const float* params, const float* x, int64_t N, int64_t E, int64_t B,
                      const void* csr, int32_t tile_rows, int32_t flags, const int64_t* ptr, const int64_t* action,
                      const float* target, const float* weight, int32_t kind, float scale, float* q, float* loss,
                      float* td_err, float* stash, float* grad_params, void* ws, size_t ws_bytes, void* stream);

/* Collate B device-resident graphs into one batch in ONE launch (replaces Batch.from_data_list over GPU-resident
 * replay-buffer entries, replay_buffer.py:116-117,215-216, and C++ collate_batch, cpp_hex/hex_graph_game/util.cpp:22-41):
 *   x_out [N,3] = cat(x_g);  ei_out [2,E] = cat(edge_index_g + node_offset_g, dim 1);  batch_out [N] = graph id per node;
 *   ptr_out [B+1] = node prefix sums.   N = sum n_nodes, E = sum n_edges (the caller sizes the outputs).
 * x_ptrs / ei_ptrs / n_nodes / n_edges are HOST arrays of length B; x_ptrs[g] -> device [n_g,3] f32, ei_ptrs[g] -> device
 * [2,e_g] i64 row-major.  ws: device, host_ws: host (pinned preferred; it must stay untouched until the stream has
 * passed this call), both hexgnn_collate_ws_bytes(B).  Bit-exact. */
int64_t hexgnn_collate_ws_bytes(int64_t B);
int hexgnn_collate(const float* const* x_ptrs, const int64_t* const* ei_ptrs, const int64_t* n_nodes, const int64_t* n_edges,
                   int64_t B, float* x_out, int64_t* ei_out, int64_t* batch_out, int64_t* ptr_out, void* ws, size_t ws_bytes,
                   void* host_ws, size_t host_ws_bytes, void* stream);

/* Optimizer step over the flat buffers, two launches (replaces nn.utils.clip_grad_norm_ + torch.optim.Adam.step over
 * ~70 tensors, rainbow.py:79,320,325): total L2 norm of `grads` -> grad_norm[0] (may be NULL); when max_grad_norm > 0
 * the gradient is scaled IN PLACE by min(1, max_grad_norm / (norm + 1e-6)); then Adam (no weight decay, no amsgrad)
 * updates params / exp_avg / exp_avg_sq in place.  `state` = 4 int32 (zero them once): step counters of the body, the
 * maker head and the breaker head, and the head taken last.  torch keeps a step counter per parameter and skips
 * parameters without gradient; here a head's counter starts with the first step that takes it (x[0,2], models.py:556)
 * and then advances every step -- torch 1.13 `opt.zero_grad()` (rainbow.py:326) leaves zero gradients, so Adam keeps
 * applying that head's momentum -- unless HEXGNN_ADAM_SKIP_UNTAKEN is set (torch >= 2 set_to_none semantics).
 * ws = hexgnn_adam_ws_bytes().  Deterministic. */
#define HEXGNN_ADAM_SKIP_UNTAKEN 1
int64_t hexgnn_adam_ws_bytes(void);
int hexgnn_adam_step(const HexModelDesc* desc, float* params, float* grads, float* exp_avg, float* exp_avg_sq,
                     int32_t* state, const float* x, float lr, float beta1, float beta2, float eps, float max_grad_norm,
                     int32_t flags, float* grad_norm, void* ws, size_t ws_bytes, void* stream);

/* FactorizedNoisyLinear on the advantage Linear(H,1) of both heads, `--noisy_dqn=True` (GN0/models.py:84-141 the layer,
 * :331-334 HeadNetwork uses it for `self.linear`, :944-945 factory arguments; Rainbow resets / disables the noise at
 * rainbow.py:167-192).  The fused kernels read the advantage weights from the flat parameter buffer, so:
 *   hexgnn_noisy_compose: params[linear.W | linear.b of head h] = mu + sigma * epsilon            (models.py:134-141)
 *   hexgnn_noisy_grads:   d_mu = g, d_sigma = g * epsilon, g = the composed weights' slice of the flat gradient `grads`;
 *                         accumulate != 0 adds into d_noisy (autograd `+=`), else overwrites.
 * noisy [2][3][H+1] f32: per head (maker, breaker) the rows mu, sigma, epsilon, each = weight[H] then the bias;
 * d_noisy [2][2][H+1]: per head the rows d_mu, d_sigma.  One launch each, element-wise, deterministic. */
int hexgnn_noisy_compose(const HexModelDesc* desc, float* params, const float* noisy, void* stream);
int hexgnn_noisy_grads(const HexModelDesc* desc, const float* grads, const float* noisy, float* d_noisy, int32_t accumulate,
                       void* stream);

/* Learner loss glue on device (replaces the index / loss / mean kernels of rainbow.py:301-312 and their autograd):
 *   td_est[g] = q[ptr[g] + action[g]];  td_err[g] = td_est[g] - target[g]  (PER priorities = |td_err| + 1e-6, :307-308)
 *   loss[0]   = scale * mean_g( weight[g] * l(target[g], td_est[g]) ),  kind 0: l = MSE, 1: l = SmoothL1 (beta 1), :85
 *   d_q [N]   = d loss / d q (zero except the B selected rows), OVERWRITTEN; may be NULL (evaluation only)
 * q [N] f32, ptr [B+1] i64, action [B] i64, target [B] f32, weight [B] f32 or NULL (uniform), td_err [B] or NULL.
 * `scale` = 1/world_size for data-parallel shards (the mean is over the GLOBAL batch).  One launch, deterministic. */
#define HEXGNN_LOSS_MSE 0
#define HEXGNN_LOSS_SMOOTH_L1 1
int hexgnn_td_loss(const float* q, const int64_t* ptr, const int64_t* action, const float* target, const float* weight,
                   int64_t N, int64_t B, int32_t kind, float scale, float* loss, float* td_err, float* d_q, void* stream);

/* Acting / target-network glue (rainbow.py:222-225 `act`, :263-269 `td_target`): when mask_terminals != 0 the two
 * terminal rows of every graph (q[ptr[g]], q[ptr[g]+1]: not actions) are overwritten with mask_value (-5 in the
 * reference), then out_max[g] / out_arg[g] = maximum of graph g and the GLOBAL row index attaining it (ties: the lowest
 * index -- torch_scatter.scatter_max leaves ties unspecified; an empty graph gives 0 and N like torch_scatter).
 * q [N] f32 in/out, ptr [B+1] i64, out_max [B] f32 or NULL, out_arg [B] i64 or NULL.  One launch, one warp per graph. */
int hexgnn_mask_argmax(float* q, const int64_t* ptr, int64_t N, int64_t B, int32_t mask_terminals, float mask_value,
                       float* out_max, int64_t* out_arg, void* stream);
/* n-step double-DQN target (rainbow.py:268-269): target[g] = reward[g] + gamma_n * q_target[best[g]] * (1 - done[g]),
 * best [B] i64 global row ids (hexgnn_mask_argmax of the policy net), reward / done / target [B] f32. */
int hexgnn_td_target(const float* q_target, const int64_t* best, const float* reward, const float* done, float gamma_n,
                     int64_t N, int64_t B, float* target, void* stream);

/* Data-parallel gradient all-reduce over NVLink peer memory (no reference counterpart).  peer_bufs is a
 * DEVICE array of `world` pointers, entry r = rank r's flat gradient of n floats in peer-mapped (symmetric) memory,
 * 16-byte aligned; out [n] (private, NOT one of the peer buffers) = their sum in rank order, bit-identical on every
 * rank.  The caller brackets the call with cross-rank barriers: all ranks have written before it, none overwrites its
 * buffer until all have returned from it (torch.distributed._symmetric_memory.barrier in the Python host side). */
int hexgnn_allreduce_p2p(const float* const* peer_bufs, int32_t world, int64_t n, float* out, void* stream);
/* The same all-reduce through the NVSwitch, two-shot and IN PLACE (no reference counterpart).  mc_buf = the
 * MULTICAST (NVLS) address of a symmetric allocation that holds every rank's n floats (16-byte aligned, n a multiple of
 * 4 * world: pad with zeros).  Rank r reduces slice r with multimem.ld_reduce (the switch adds the W copies in fp32) and
 * writes the sums into all W buffers with multimem.st; 2n / W floats cross each GPU's links instead of the W * n of the
 * one-shot pull.  peer_flags: DEVICE array of `world` pointers, entry r = rank r's int32 flag array [world] in peer-mapped
 * memory (zeroed once); epoch: 1, 2, 3, ... (one more per call, the same on every rank); counter: one device int32, zeroed
 * once.  Both barriers are inside the kernel (flag words, values only grow): it starts reducing when every rank has reached the
 * call -- the gradient writes precede it in stream order -- and returns only when every peer's slice has landed in this
 * rank's buffer.  Every rank must make the same sequence of calls.  Every rank ends with identical bits (each slice is summed
 * once, by one rank). */
int hexgnn_allreduce_mc(float* mc_buf, int32_t* const* peer_flags, int32_t rank, int32_t world, int64_t n, int32_t epoch,
                        int32_t* counter, void* stream);

/* Where forward (with `flags`) leaves the body embedding h_L (`final_conv_acts`, models.py:553) inside the stash:
 * float offset of a row-major [N, row_pitch] block.  (The two kernel paths keep different stash formats.) */
int hexgnn_stash_embed(const HexModelDesc* desc, int64_t N, int64_t B, int32_t flags, int64_t* offset_floats,
                       int64_t* row_pitch);

/* Self-test of the tcgen05 building blocks the fused kernels are made of: one CTA computes
 * D[128,N] = A[128,K] * B[K,N] (row-major device fp32, 3xTF32).  `modes` = a_mode*16 + b_mode selects the
 * operand staging (a_mode 0 tensor memory | 1 smem K-major SWIZZLE_NONE | 3 smem MN-major
 * SWIZZLE_128B_BASE32B; b_mode 0 smem K-major SWIZZLE_NONE | 3 smem MN-major SWIZZLE_128B_BASE32B).
 * N%16==0, K%8==0, K<=128. */
int hexgnn_selftest_mma(const float* A, const float* B, float* D, int32_t N, int32_t K, int32_t modes, void* stream);

/* Phase profile of the last tcgen05 launch (which: 0 forward, 1 backward): per CTA (148 x 16 int64, HOST
 * buffer) cycle totals of the kernel's internal phases, see the table above g_fwd_phase / g_bwd_phase.
 * Synchronises the device.  Diagnostic only. */
int hexgnn_debug_phases(int32_t which, int64_t* host_out, int32_t n);

/* Measurement aid (bench.py `roofline`): while enabled (process-wide; NOT re-entrant, use from one
 * stream at a time), hexgnn_forward and
 * hexgnn_backward bracket their kernels (which: 0 = fused forward, 1 = backward dH chain + dW GEMM (overlapped, one
 * pair), 2 = the forward + chain + dW trio of hexgnn_train_step) with
 * CUDA events recorded on `stream`.  hexgnn_profile_read synchronises on the recorded events, returns
 * the summed device time (ms) and the number of bracketed launches since the last read, and resets.
 * This is the only mutable state the library keeps besides the thread-local error string. */
int hexgnn_profile_enable(int on);
int hexgnn_profile_read(int which, double* total_ms, int64_t* launches);

#ifdef __cplusplus
}
#endif
#endif /* HEXGNN_H_ */
