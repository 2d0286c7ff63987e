"""Host-side mirror of the HexAra policy / value network `SAGE_torch_script` (GN0/torch_script_models.py:286-379; factory
`get_current_model`, :495-500), the module the reference traces to TorchScript (rl_loop/trace_model.py) and the C++ engine
evaluates through `NN_api::predict` (cpp_hex/CrazyAra/nn_api.cpp:14-19).  Same constructor arguments, parameter NAMES
(`gnn.convs.i.lin_l/lin_r`, `my_modules.{value_head,policy_head}.convs.j...`, `my_modules.{value_linear,swap_linear}.layers.k`)
and the same 4-tuple from forward, so reference checkpoints load with `load_state_dict`.  All arithmetic runs in libhexgnn.so
(`hexgnn_hexara_forward` / `hexgnn_hexara_backward`: layer-at-a-time tcgen05 kernels): inference, and training through
autograd (`loss.backward()` fills every `p.grad`, rl_loop/trainer_agent_pytorch.py:245-275)."""
import ctypes as C

import torch
from torch import nn

from . import _lib, ops
from .models import _Conv, _MLPParams


class _AraStackParams(nn.Module):
    """Parameter holder named like `ModifiedBaseNet` (torch_script_models.py:96-155): `.convs`; in -> hidden, hidden -> hidden
    ..., hidden -> out_channels (a one-conv stack maps in -> out directly)."""

    def __init__(self, in_channels, hidden, num_layers, out_channels=None):
        super().__init__()
        self.in_channels, self.hidden_channels, self.num_layers = in_channels, hidden, num_layers
        self.out_channels = hidden if out_channels is None else out_channels
        self.convs = nn.ModuleList()
        cin = in_channels
        for i in range(num_layers):
            self.convs.append(_Conv(cin, self.out_channels if i == num_layers - 1 else hidden))
            cin = hidden


class _AraFn(torch.autograd.Function):
    """Autograd node around hexgnn_hexara_forward (with stash) / hexgnn_hexara_backward.  `anchor` (one parameter) makes the
    outputs require grad; the backward kernel writes the whole flat gradient and every `p.grad` becomes (or receives) its slice."""

    @staticmethod
    def forward(ctx, model, x, csr, anchor):
        pi, value, gi, bp, emb, model._ws, stash = ops.hexara_forward(model._desc, model._flat, x, csr,
                                                                      want_embeds=model.capture_embeds, ws=model._ws,
                                                                      want_stash=True)
        ctx.model, ctx.csr, ctx.stash, ctx.flat = model, csr, stash, model._flat
        ctx.save_for_backward(x, pi, bp)
        ctx.mark_non_differentiable(gi, bp)
        ctx.set_materialize_grads(False)
        model._emb = emb
        return pi, value, gi, bp

    @staticmethod
    def backward(ctx, d_pi, d_value, _gi, _bp):
        model = ctx.model
        x, pi, bp = ctx.saved_tensors
        if ctx.flat is not model._flat:
            raise _lib.HexGnnError("parameters were re-flattened between forward and backward")
        if d_pi is None and d_value is None:
            return None, None, None, None
        grad, model._bws = ops.hexara_backward(model._desc, model._flat, x, ctx.csr, pi, bp, d_pi, d_value, ctx.stash,
                                               ws=model._bws)
        for p, (o, n, shp) in zip(model._params_cache, model._slices):
            if not p.requires_grad:
                continue
            g = grad[o:o + n].view(shp)
            if p.grad is None:
                p.grad = g
            else:
                p.grad = p.grad + g
        model.flat_gradient = grad
        return None, None, None, None


class SAGETorchScript(nn.Module):
    """`SAGE_torch_script(hidden_channels, hidden_layers, policy_layers, value_layers, in_channels=3, swap_allowed=False)`.

    forward(x, edge_index, graph_indices, batch_ptr) -> (pi, value, output_graph_indices, output_batch_ptr)
      x [N,3] in the C++ feature convention (IS_TERMINAL, swap flags: shannon_node_switching_game.cpp:690-726); rows 0, 1 of
      every graph are terminals and are removed from the policy; with `swap_allowed`, a graph whose last node has
      x[:,2] != 0 gets one more policy entry (the swap logit) after its nodes.  `pi` is the per-graph log-softmax."""

    def __init__(self, hidden_channels, hidden_layers, policy_layers, value_layers, in_channels=3, swap_allowed=False,
                 norm=None):
        super().__init__()
        if norm is not None:
            raise NotImplementedError("SAGE_torch_script(norm=...) is not on the fused path (get_current_model's default is None)")
        if in_channels != 3:
            raise NotImplementedError("in_channels must be 3 (the C++ feature convention)")
        self.swap_allowed = bool(swap_allowed)
        self.hidden_channels = int(hidden_channels)
        self.final_conv_acts = None
        self.gnn = _AraStackParams(in_channels, hidden_channels, hidden_layers)
        self.my_modules = nn.ModuleDict()
        self.my_modules["value_head"] = _AraStackParams(hidden_channels, hidden_channels, value_layers)
        self.my_modules["policy_head"] = _AraStackParams(hidden_channels, hidden_channels, policy_layers, out_channels=1)
        self.my_modules["value_linear"] = _MLPParams(hidden_channels // 2, hidden_channels * 4, 1)
        self.my_modules["swap_linear"] = _MLPParams(hidden_channels // 2, hidden_channels * 4, 1)
        self._desc = _lib.HexAraDesc(int(hidden_channels), int(hidden_layers), int(policy_layers), int(value_layers),
                                     int(in_channels), 1 if swap_allowed else 0)
        self._flat = None
        self._ws = self._bws = None
        self._slices = self._params_cache = None
        self._emb = None
        self.flat_gradient = None         # the flat gradient of the last backward (every p.grad is a slice of one)
        self.capture_embeds = False       # True: fill final_conv_acts (torch_script_models.py:314)
        self.last_csr = None

    def _ordered_params(self):
        ps = []
        for stack in (self.gnn, self.my_modules["value_head"], self.my_modules["policy_head"]):
            for c in stack.convs:
                ps += [c.lin_l.weight, c.lin_l.bias, c.lin_r.weight]
        for name in ("value_linear", "swap_linear"):
            mlp = self.my_modules[name]
            ps += [mlp.layers[0].weight, mlp.layers[0].bias, mlp.layers[1].weight, mlp.layers[1].bias]
        return ps

    def _ensure_flat(self):
        """One flat fp32 buffer in the C-ABI's order (hexgnn_hexara_param_offsets); every named parameter is a view of it."""
        ps = self._ordered_params()
        if self._flat is not None and self._flat.device == ps[0].device and \
                ps[0].data_ptr() == self._flat.data_ptr() and \
                ps[-1].data_ptr() == self._flat.data_ptr() + 4 * (self._flat.numel() - 1):
            return
        l = _lib.lib()
        n = l.hexgnn_hexara_param_offsets(C.byref(self._desc), None, 0)
        if n < 0 or n != len(ps):
            raise _lib.HexGnnError("hexara: parameter table mismatch between hexara.py and libhexgnn.so")
        offs = (C.c_int64 * n)()
        l.hexgnn_hexara_param_offsets(C.byref(self._desc), offs, n)
        total = int(l.hexgnn_hexara_param_count(C.byref(self._desc)))
        flat = torch.empty(total, dtype=torch.float32, device=ps[0].device)
        with torch.no_grad():
            for p, o in zip(ps, offs):
                k = p.numel()
                flat[o:o + k].copy_(p.detach().reshape(-1).to(torch.float32))
                p.data = flat[o:o + k].view(p.shape)
        if int(offs[-1]) + ps[-1].numel() != total:
            raise _lib.HexGnnError("hexara: flat parameter size mismatch")
        self._flat = flat
        self._params_cache = ps
        self._slices = [(int(o), p.numel(), tuple(p.shape)) for p, o in zip(ps, offs)]

    def _apply(self, fn, *args, **kwargs):
        out = super()._apply(fn, *args, **kwargs)
        self._flat = None
        return out

    @property
    def flat_parameters(self):
        self._ensure_flat()
        return self._flat

    def forward(self, x, edge_index, graph_indices, batch_ptr):
        self._ensure_flat()
        csr = ops.build_csr(edge_index, graph_indices, batch_ptr, num_nodes=int(x.shape[0]))
        self.last_csr = csr
        anchor = self._params_cache[0]
        if torch.is_grad_enabled() and any(p.requires_grad for p in self._params_cache):
            # training (rl_loop/trainer_agent_pytorch.py:245-275): the activations are stashed for hexgnn_hexara_backward
            pi, value, gi, bp = _AraFn.apply(self, x, csr, anchor)
            emb = self._emb
        else:
            pi, value, gi, bp, emb, self._ws, _ = ops.hexara_forward(self._desc, self._flat, x, csr,
                                                                    want_embeds=self.capture_embeds, ws=self._ws)
        self.final_conv_acts = emb[:, :self.hidden_channels] if emb is not None else None
        count = int(x.shape[0]) - 2 * csr.B
        if self.swap_allowed:
            count = int(bp[-1])            # data dependent (the reference's boolean-mask indexing syncs as well)
        return pi[:count], value, gi[:count], bp


def get_current_model(net_type="SAGE", hidden_channels=60, hidden_layers=15, policy_layers=2, value_layers=2, in_channels=3,
                      swap_allowed=False, norm=None):
    """`get_current_model` (GN0/torch_script_models.py:495-500) for the one network type this package implements."""
    if net_type != "SAGE":
        raise NotImplementedError(f"gnn_hex_b200 implements only the 'SAGE' HexAra network, not {net_type!r}")
    return SAGETorchScript(hidden_channels=hidden_channels, hidden_layers=hidden_layers, policy_layers=policy_layers,
                           value_layers=value_layers, in_channels=in_channels, swap_allowed=swap_allowed, norm=norm)
