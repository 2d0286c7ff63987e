"""`torch.library` custom ops over the C-ABI: the functional, flat-parameter face of the fused path.

    torch.ops.hexgnn.build_csr(edge_index, batch, ptr, num_nodes, num_graphs, tile_rows)          -> csr blob (int32)
    torch.ops.hexgnn.fused_forward(params, x, csr, E, B, hidden, L, L_head, tile_rows, flags, want_stash)
                                                                                  -> (out_nodes [N], out_value [B], stash)
    torch.ops.hexgnn.fused_backward(params, x, csr, stash, d_nodes, d_value, E, B, hidden, L, L_head, tile_rows, flags)
                                                                                  -> flat parameter gradient [P]

`params` is the flat fp32 parameter buffer in the order of include/hexgnn.h (`ModernTwoHeaded.flat_parameters`); `flags` are
HEXGNN_ADV_ONLY / HEXGNN_SEPARATE / HEXGNN_PATH_*.  Each op has a fake (meta) implementation, so the ops trace under
FakeTensorMode / torch.export / CUDA-graph capture tooling, and `fused_forward` carries an autograd formula that calls
`fused_backward` (gradient w.r.t. `params` only, as `loss.backward()` in rainbow.py:316 needs).  The `nn.Module`
(`ModernTwoHeaded`, the drop-in for GN0/models.py:477-590) stays the primary interface; these ops are what a TorchScript /
C++ caller (cpp_hex/CrazyAra/nn_api.cpp:14-19) or a graph-capturing front end binds to.  CUDA only: there is no CPU kernel."""
from typing import Optional, Tuple

import torch
from torch import Tensor

from . import _lib, ops


def _desc(hidden, num_layers, num_head_layers):
    return _lib.make_desc(int(hidden), int(num_layers), int(num_head_layers))


@torch.library.custom_op("hexgnn::build_csr", mutates_args=(), device_types="cuda")
def build_csr(edge_index: Tensor, batch: Optional[Tensor], ptr: Optional[Tensor], num_nodes: int, num_graphs: int,
              tile_rows: int = 128) -> Tensor:
    return ops.build_csr(edge_index, batch, ptr, num_nodes=num_nodes, num_graphs=num_graphs, tile_rows=tile_rows).blob


@build_csr.register_fake
def _(edge_index, batch, ptr, num_nodes, num_graphs, tile_rows=128):
    lay = _lib.csr_layout(int(num_nodes), int(edge_index.shape[1]), int(num_graphs))
    return edge_index.new_empty(int(lay.total_words), dtype=torch.int32)


@torch.library.custom_op("hexgnn::fused_forward", mutates_args=(), device_types="cuda")
def fused_forward(params: Tensor, x: Tensor, csr: Tensor, num_edges: int, num_graphs: int, hidden: int, num_layers: int,
                  num_head_layers: int, tile_rows: int = 128, flags: int = 0,
                  want_stash: bool = True) -> Tuple[Tensor, Tensor, Tensor]:
    blob = ops.CsrBlob(csr, int(x.shape[0]), num_edges, num_graphs, tile_rows)
    out_nodes, out_value, stash, _ = ops.fused_forward(_desc(hidden, num_layers, num_head_layers), params.contiguous(), x, blob,
                                                       flags, want_stash=want_stash)
    if stash is None:
        stash = x.new_empty(0)
    return out_nodes, out_value, stash


@fused_forward.register_fake
def _(params, x, csr, num_edges, num_graphs, hidden, num_layers, num_head_layers, tile_rows=128, flags=0, want_stash=True):
    N = int(x.shape[0])
    sz = _lib.sizes(_desc(hidden, num_layers, num_head_layers), N, int(num_edges), int(num_graphs), int(tile_rows))
    return (x.new_empty(N), x.new_empty(int(num_graphs)), x.new_empty(int(sz.stash_bytes) // 4 if want_stash else 0))


@torch.library.custom_op("hexgnn::fused_backward", mutates_args=(), device_types="cuda")
def fused_backward(params: Tensor, x: Tensor, csr: Tensor, stash: Tensor, d_nodes: Tensor, d_value: Optional[Tensor],
                   num_edges: int, num_graphs: int, hidden: int, num_layers: int, num_head_layers: int,
                   tile_rows: int = 128, flags: int = 0) -> Tensor:
    blob = ops.CsrBlob(csr, int(x.shape[0]), num_edges, num_graphs, tile_rows)
    if stash.numel() == 0:
        raise _lib.HexGnnError("fused_backward: the forward ran with want_stash=False")
    grad, _ = ops.fused_backward(_desc(hidden, num_layers, num_head_layers), params.contiguous(), x, blob, flags, d_nodes,
                                 d_value, stash)
    return grad


@fused_backward.register_fake
def _(params, x, csr, stash, d_nodes, d_value, num_edges, num_graphs, hidden, num_layers, num_head_layers, tile_rows=128,
      flags=0):
    return params.new_empty(params.numel())


def _setup_context(ctx, inputs, output):
    params, x, csr, num_edges, num_graphs, hidden, num_layers, num_head_layers, tile_rows, flags, want_stash = inputs
    ctx.save_for_backward(params, x, csr, output[2])
    ctx.meta = (num_edges, num_graphs, hidden, num_layers, num_head_layers, tile_rows, flags)


def _backward(ctx, d_nodes, d_value, d_stash):
    params, x, csr, stash = ctx.saved_tensors
    num_edges, num_graphs, hidden, num_layers, num_head_layers, tile_rows, flags = ctx.meta
    if d_nodes is None:
        d_nodes = x.new_zeros(x.shape[0])
    use_dv = bool(flags & ops.HEXGNN_SEPARATE) and d_value is not None
    grad = torch.ops.hexgnn.fused_backward(params, x, csr, stash, d_nodes.contiguous(), d_value.contiguous() if use_dv else None,
                                           num_edges, num_graphs, hidden, num_layers, num_head_layers, tile_rows, flags)
    return (grad.view_as(params),) + (None,) * 10


fused_forward.register_autograd(_backward, setup_context=_setup_context)


# ---- HexAra network: the op a TorchScript / C++ front end binds instead of the traced SAGE_torch_script ----
@torch.library.custom_op("hexgnn::hexara_forward", mutates_args=(), device_types="cuda")
def hexara_forward(params: Tensor, x: Tensor, edge_index: Tensor, graph_indices: Tensor, batch_ptr: Tensor, hidden: int,
                   hidden_layers: int, policy_layers: int, value_layers: int,
                   swap_allowed: bool = False) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """`SAGE_torch_script.forward(x, edge_index, graph_indices, batch_ptr)` (GN0/torch_script_models.py:309-379) on the flat
    parameter buffer (`SAGETorchScript.flat_parameters`): (pi, value, output_graph_indices, output_batch_ptr).  pi and
    output_graph_indices have the CAPACITY length N - 2B (+ B with swap_allowed); output_batch_ptr[-1] entries are valid."""
    desc = _lib.HexAraDesc(int(hidden), int(hidden_layers), int(policy_layers), int(value_layers), 3, 1 if swap_allowed else 0)
    csr = ops.build_csr(edge_index, graph_indices, batch_ptr, num_nodes=int(x.shape[0]), num_graphs=int(batch_ptr.shape[0]) - 1)
    pi, value, gi, bp, _, _, _ = ops.hexara_forward(desc, params.contiguous(), x, csr)
    return pi, value, gi, bp


@hexara_forward.register_fake
def _(params, x, edge_index, graph_indices, batch_ptr, hidden, hidden_layers, policy_layers, value_layers, swap_allowed=False):
    N, B = int(x.shape[0]), int(batch_ptr.shape[0]) - 1
    cap = max(N - 2 * B, 0) + (B if swap_allowed else 0)
    return (x.new_empty(cap), x.new_empty(B), batch_ptr.new_empty(cap), batch_ptr.new_empty(B + 1))
