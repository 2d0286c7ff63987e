"""Host-side mirror of the reference model API for the `modern_two_headed` Q-network.

`get_pre_defined("modern_two_headed", args)` (GN0/models.py:892,928-947) returns a module whose
forward/simple_forward signatures, attributes and parameter NAMES match the reference's
`DuellingTwoHeaded` (GN0/models.py:477-590) over PyG GraphSAGE, so checkpoints and callers
(Rainbow/common/rainbow.py:221,263,268,302; GN0/RainbowDQN/evaluate_elo.py:251) carry over.
All arithmetic runs in libhexgnn.so; this file only owns parameters and plumbing.
"""
from typing import Optional

import torch
from torch import nn

from . import _lib, ops


class _Conv(nn.Module):
    """Parameter holder named like PyG 2.2.0 `SAGEConv`: lin_l (weight+bias), lin_r (weight)."""

    def __init__(self, cin, cout):
        super().__init__()
        self.lin_l = nn.Linear(cin, cout, bias=True)
        self.lin_r = nn.Linear(cin, cout, bias=False)


class _NormParams(nn.Module):
    """Affine parameters named like PyG `LayerNorm(hidden)` (`weight` ones, `bias` zeros), the `norm` the reference passes
    to GraphSAGE with `--norm=True` (GN0/models.py:935,945).  Evaluated inside libhexgnn.so (graph-mode statistics over
    the whole batch, because models.py:285 calls it without `batch`)."""

    def __init__(self, channels):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(channels))
        self.bias = nn.Parameter(torch.zeros(channels))


class _SageParams(nn.Module):
    """Holder named like `cachify_gnn(GraphSAGE)` (GN0/models.py:144-160): `.convs` (+ `.norms` with norm=True, one per
    layer incl. the last: models.py:158-162), plus the attributes callers read (`hidden_channels` train.py:301,
    `has_cache` rainbow.py:344)."""
    supports_cache = True

    def __init__(self, in_channels, hidden, num_layers, norm=False):
        super().__init__()
        self.in_channels = in_channels
        self.hidden_channels = hidden
        self.num_layers = num_layers
        self.has_cache = False
        self.convs = nn.ModuleList([_Conv(in_channels if i == 0 else hidden, hidden) for i in range(num_layers)])
        self.norms = nn.ModuleList([_NormParams(hidden) for _ in range(num_layers)]) if norm else None

    @torch.no_grad()
    def grow_depth(self, additional_layers):
        """GN0/models.py:166-185: append layers that start as the identity (lin_l = 0, lin_r = I; the input is a ReLU
        output, so relu(I x) = x and the network's function is unchanged)."""
        if self.norms is not None:
            raise NotImplementedError("grow_depth with norm=True (the reference's identity layer is not the identity "
                                      "through a LayerNorm either)")
        dev = self.convs[0].lin_l.weight.device
        H = self.hidden_channels
        for _ in range(int(additional_layers)):
            c = _Conv(H, H).to(dev)
            c.lin_l.weight.zero_()
            c.lin_l.bias.zero_()
            c.lin_r.weight.copy_(torch.eye(H, device=dev))
            self.convs.append(c)
        self.num_layers += int(additional_layers)
        self.has_cache = False

    @torch.no_grad()
    def grow_width(self, new_width, new_in_channels=None):
        """GN0/models.py:187-238: freshly initialised wider layers whose top-left block is the old layer; the old rows see
        nothing of the new input columns (zeros), the new rows keep their random init."""
        if self.num_layers < 2:
            raise NotImplementedError("grow_width of a one-layer stack")
        if self.norms is not None:
            raise NotImplementedError("grow_width with norm=True (models.py:187-238 rebuilds the convs only)")
        dev = self.convs[0].lin_l.weight.device
        old, H = self.convs, self.hidden_channels
        old_in = self.in_channels
        if new_in_channels is not None:
            self.in_channels = int(new_in_channels)
        new = nn.ModuleList([_Conv(self.in_channels if i == 0 else new_width, new_width).to(dev)
                             for i in range(self.num_layers)])
        for i, (c, o) in enumerate(zip(new, old)):
            cin_old = (old_in if i == 0 else H)
            for lin, lin_o in ((c.lin_l, o.lin_l), (c.lin_r, o.lin_r)):
                lin.weight[:H, :cin_old] = lin_o.weight
                lin.weight[:H, cin_old:] = 0
            c.lin_l.bias[:H] = o.lin_l.bias
        self.convs = new
        self.has_cache = False
        self.hidden_channels = int(new_width)


class _MLPParams(nn.Module):
    """Holder named like `MLP(hidden//2, 1, 4*hidden, 1)` (GN0/models.py:36-49): `.layers.{0,1}`."""

    def __init__(self, hidden, num_input, num_output):
        super().__init__()
        self.layers = nn.ModuleList([nn.Linear(num_input, hidden), nn.Linear(hidden, num_output)])
        self.hidden_channels, self.num_input, self.num_output = hidden, num_input, num_output

    @torch.no_grad()
    def grow_input_width(self, new_input_width, new_hidden_channels):
        """GN0/models.py:52-73 for the one-hidden-layer MLP: zero-filled wider layers with the old weights in the top-left
        block.  As in the reference the old input columns keep their POSITIONS (the pooled vector is [sum | max | min |
        mean] of the new width, so they no longer line up with the old pooling blocks)."""
        dev = self.layers[0].weight.device
        l0, l1 = self.layers[0], self.layers[1]
        n0 = nn.Linear(new_input_width, new_hidden_channels).to(dev)
        n0.weight.zero_()
        n0.bias.zero_()
        n0.weight[:self.hidden_channels, :self.num_input] = l0.weight
        n0.bias[:self.hidden_channels] = l0.bias
        n1 = nn.Linear(new_hidden_channels, 1).to(dev)
        n1.weight.zero_()
        n1.weight[:, :self.hidden_channels] = l1.weight
        n1.bias[:] = l1.bias
        self.layers[0], self.layers[1] = n0, n1
        self.num_input, self.hidden_channels = int(new_input_width), int(new_hidden_channels)


class FactorizedNoisyLinear(nn.Module):
    """Parameter / noise holder named like the reference's noisy-nets layer (GN0/models.py:84-141): `weight_mu`,
    `weight_sigma`, `bias_mu`, `bias_sigma` parameters and `weight_epsilon`, `bias_epsilon` buffers, with the reference's
    `reset_parameters` / `reset_noise` / `disable_noise` (called by rainbow.py:167-192).  The layer has no forward of its
    own: the owning model composes `mu + sigma * epsilon` into the fused kernels' weight slot (hexgnn_noisy_compose) and
    chains the gradient back (hexgnn_noisy_grads)."""

    def __init__(self, in_features: int, out_features: int, sigma_0: float):
        super().__init__()
        if out_features != 1:
            raise NotImplementedError("the fused path has the noisy layer on the advantage Linear(H, 1) only")
        self.in_features, self.out_features, self.sigma_0 = in_features, out_features, sigma_0
        self.weight_mu = nn.Parameter(torch.empty(out_features, in_features))
        self.weight_sigma = nn.Parameter(torch.empty(out_features, in_features))
        self.register_buffer("weight_epsilon", torch.empty(out_features, in_features))
        self.bias_mu = nn.Parameter(torch.empty(out_features))
        self.bias_sigma = nn.Parameter(torch.empty(out_features))
        self.register_buffer("bias_epsilon", torch.empty(out_features))
        # the composed weights' slots of the flat parameter buffer (derived values: neither parameters nor buffers)
        self._w_eff = torch.zeros(out_features, in_features)
        self._b_eff = torch.zeros(out_features)
        self.reset_parameters()
        self.reset_noise()

    @torch.no_grad()
    def reset_parameters(self) -> None:
        """Reference init (models.py:105-114): mu ~ U(+-1/sqrt(fan_in)), sigma = sigma_0 / sqrt(fan_in); weight first,
        then bias, so a seeded construction draws the same numbers as the reference class."""
        bound = self.in_features ** -0.5
        for mu, sigma in ((self.weight_mu, self.weight_sigma), (self.bias_mu, self.bias_sigma)):
            mu.uniform_(-bound, bound)
            sigma.fill_(self.sigma_0 * bound)

    @torch.no_grad()
    def reset_noise(self) -> None:
        """Fresh factorised Gaussian noise (models.py:116-128; NoisyNets eq. 10-11): with f(z) = sgn(z) sqrt|z|,
        eps_w = f(z_out) f(z_in)^T and eps_b = f(z_out); z_in is drawn before z_out, as in the reference."""
        dev = self.weight_mu.device
        z_in = torch.randn(self.in_features, device=dev)
        z_out = torch.randn(self.out_features, device=dev)
        f_in = torch.copysign(z_in.abs().sqrt(), z_in)
        f_out = torch.copysign(z_out.abs().sqrt(), z_out)
        torch.outer(f_out, f_in, out=self.weight_epsilon)
        self.bias_epsilon.copy_(f_out)

    @torch.no_grad()
    def disable_noise(self) -> None:
        """Evaluation with the mean weights (models.py:130-133)."""
        self.weight_epsilon.zero_()
        self.bias_epsilon.zero_()

    def forward(self, input):
        raise _lib.HexGnnError("FactorizedNoisyLinear is evaluated inside the fused kernels (call the owning model)")


class _HeadParams(nn.Module):
    """Holder named like `HeadNetwork` (GN0/models.py:318-334): gnn, linear, value_head."""
    supports_cache = True

    def __init__(self, hidden, num_layers, noisy_dqn=False, noise_sigma=0.0, norm=False):
        super().__init__()
        self.gnn = _SageParams(hidden, hidden, num_layers, norm=norm)
        self.linear = FactorizedNoisyLinear(hidden, 1, noise_sigma) if noisy_dqn else nn.Linear(hidden, 1)
        self.value_head = _MLPParams(hidden // 2, 4 * hidden, 1)
        self.hidden_channels = hidden

    @torch.no_grad()
    def grow_width(self, new_width, new_in_channels=None):
        """GN0/models.py:336-358."""
        if not isinstance(self.linear, nn.Linear):
            raise AssertionError("grow_width needs a plain advantage Linear (models.py:337)")
        H = self.hidden_channels
        self.gnn.grow_width(new_width, new_in_channels=new_in_channels)
        old = self.linear
        self.linear = nn.Linear(new_width, 1).to(old.weight.device)
        self.linear.weight.zero_()
        self.linear.weight[:, :H] = old.weight
        self.linear.bias[:] = old.bias
        self.value_head.grow_input_width(new_width * 4, new_width // 2)
        self.hidden_channels = int(new_width)

    def grow_depth(self, additional_layers):
        """GN0/models.py:360-361 (train.py:306-309 grows both heads)."""
        self.gnn.grow_depth(additional_layers)


class _FusedQ(torch.autograd.Function):
    """Autograd node around hexgnn_forward / hexgnn_backward.

    Only one parameter (`anchor`) is an autograd input: it makes the outputs require grad.  The
    backward kernel writes the whole flat gradient in one pass and the module points every
    `p.grad` at its slice of that buffer (accumulating when the slices are still in place, exactly
    like autograd's `+=`), so a step costs no per-parameter kernels."""

    @staticmethod
    def forward(ctx, model, x, csr, flags, anchor):
        out_nodes, out_value, stash, hp = ops.fused_forward(model._desc, model._flat, x, csr, flags,
                                                            want_stash=True, ws=model._fwd_ws(csr))
        ctx.model, ctx.csr, ctx.flags, ctx.stash = model, csr, flags, stash
        ctx.save_for_backward(x)
        ctx.flat_version = model._flat_version
        off, pitch = _lib.stash_embed(model._desc, csr.N, csr.B, flags)
        model.final_conv_acts = stash[off:off + csr.N * pitch].view(csr.N, pitch)[:, :model.hidden_channels]
        return out_nodes, out_value

    @staticmethod
    def backward(ctx, d_nodes, d_value):
        model = ctx.model
        (x,) = ctx.saved_tensors
        if ctx.flat_version != model._flat_version:
            raise _lib.HexGnnError("parameters were re-flattened between forward and backward")
        if d_nodes is None:
            d_nodes = torch.zeros(ctx.csr.N, dtype=torch.float32, device=x.device)
        use_dv = bool(ctx.flags & ops.HEXGNN_SEPARATE) and d_value is not None
        accumulate = model._grads_in_place()
        out = model._grad_scratch() if accumulate else model._flat_grad_buffer()
        # x.requires_grad: the caller wants dL/dx (`input.grad`, rl_loop/explain.py:26-36)
        d_x = torch.zeros_like(x, dtype=torch.float32) if ctx.needs_input_grad[1] else None
        grad, d_emb = ops.fused_backward(model._desc, model._flat, x, ctx.csr, ctx.flags, d_nodes,
                                         d_value if use_dv else None, ctx.stash,
                                         want_d_embeds=model.capture_embed_grads, ws=model._bwd_ws(ctx.csr),
                                         grad_out=out, d_x=d_x)
        if accumulate:
            model._flat_grad.add_(grad)
        else:
            model._point_grads()
        model._noisy_backward(grad)
        if d_emb is not None:
            model.final_conv_grads = d_emb[:, :model.hidden_channels]
        return None, d_x, None, None, None


class _LwQ(torch.autograd.Function):
    """Autograd node around hexgnn_forward_lw / hexgnn_backward_lw (layer-at-a-time tensor-core path: graphs of any size,
    norm=True).  Same contract as `_FusedQ`; the LayerNorm affine gradients go to a second small flat buffer."""

    @staticmethod
    def forward(ctx, model, x, csr, flags, anchor):
        out_nodes, out_value, stash, sz = ops.lw_forward(model._desc, model._flat, model._norm_flat, x, csr, flags,
                                                         want_stash=True, ws=model._lw_ws(csr, "fwd"))
        ctx.model, ctx.csr, ctx.flags, ctx.stash = model, csr, flags, stash
        ctx.save_for_backward(x)
        ctx.flat_version = model._flat_version
        off, pitch = int(sz.embed_offset_floats), int(sz.hidden_padded)
        model.final_conv_acts = stash[off:off + csr.N * pitch].view(csr.N, pitch)[:, :model.hidden_channels]
        return out_nodes, out_value

    @staticmethod
    def backward(ctx, d_nodes, d_value):
        model = ctx.model
        (x,) = ctx.saved_tensors
        if ctx.flat_version != model._flat_version:
            raise _lib.HexGnnError("parameters were re-flattened between forward and backward")
        if ctx.needs_input_grad[1]:
            raise _lib.HexGnnError("dL/dx is available on the fused path only (norm=False, graphs within one tile)")
        if d_nodes is None:
            d_nodes = torch.zeros(ctx.csr.N, dtype=torch.float32, device=x.device)
        use_dv = bool(ctx.flags & ops.HEXGNN_SEPARATE) and d_value is not None
        model._lw_backward(x, ctx.csr, ctx.flags, d_nodes, d_value if use_dv else None, ctx.stash)
        return None, None, None, None, None


class ModernTwoHeaded(nn.Module):
    """Drop-in for `get_pre_defined("modern_two_headed", args)` (norm, noisy_dqn included).

    forward(x, edge_index, graph_indices=None, ptr=None, set_cache=False, advantages_only=False,
            seperate=False)            -- GN0/models.py:537
    simple_forward(data)               -- GN0/models.py:586-590
    Extra, non-reference keyword: `num_graphs` (avoids the host sync of `int(graph_indices.max())+1`,
    models.py:576, when `ptr` is not passed) and `csr` (reuse a CSR built by `build_csr`).
    """
    supports_cache = True  # class attribute in the reference (models.py:148,486); read at train.py:98

    def __init__(self, hidden_channels: int, num_layers: int, num_head_layers: int = 2, tile_rows: int = 128,
                 strict: bool = False, path: str = "auto", max_graph_nodes: int = 0, noisy_dqn: bool = False,
                 noisy_sigma0: float = 0.5, inference_precision: str = "tf32x3", norm: bool = False):
        super().__init__()
        self.hidden_channels = int(hidden_channels)
        self.num_layers = int(num_layers)
        self.num_head_layers = int(num_head_layers)
        self.tile_rows = int(tile_rows)
        self.strict = strict                  # True: sync + raise on device-side status after each forward
        self.path = path                      # "auto" | "simt" | "tc" | "lw" (layer-at-a-time tensor-core path)
        if path not in ("auto", "simt", "tc", "lw"):
            raise ValueError("path must be 'auto', 'simt', 'tc' or 'lw'")
        # args.norm (models.py:935,945): PyG LayerNorm after every conv + after_embed_norm (models.py:480-483), statistics
        # over the whole batch -> always the layer-at-a-time path
        self.norm = bool(norm)
        if self.norm and path in ("simt", "tc"):
            raise ValueError("norm=True runs on the layer-at-a-time path only (path='auto' or 'lw')")
        # SURVEY 7.2(1) precision policy.  "tf32x3" (default): fp32 parity everywhere.  "tf32x1": forwards that keep no
        # stash (torch.no_grad(): acting, target evaluation) run ONE tensor-core pass per k-step (~1e-3 relative error,
        # opt-in; rainbow.py:214-239 only takes an arg-max).  Calls that can be differentiated always use 3xTF32.
        if inference_precision not in ("tf32x3", "tf32x1"):
            raise ValueError("inference_precision must be 'tf32x3' or 'tf32x1'")
        self.inference_precision = inference_precision
        self.capture_embed_grads = False      # True: fill final_conv_grads (rl_loop/explain.py:41)
        # > tile_rows: the batch may hold graphs that do not fit one tile of the fused kernels (board-size transfer up to
        # 19x19 = 363 nodes, evaluate_elo.py:229-251, plotting.py:61-80): every call, training included, then runs the
        # layer-at-a-time path, which has no graph-size limit (the number itself is not a limit any more); 0 = such graphs
        # are flagged (NaN outputs + status bit)
        self.max_graph_nodes = int(max_graph_nodes)
        self.gnn = _SageParams(2, hidden_channels, num_layers, norm=self.norm)
        self.noisy_dqn = bool(noisy_dqn)      # args.noisy_dqn / args.noisy_sigma0 (models.py:944-945)
        self.maker_head = _HeadParams(hidden_channels, num_head_layers, self.noisy_dqn, noisy_sigma0, norm=self.norm)
        self.breaker_head = _HeadParams(hidden_channels, num_head_layers, self.noisy_dqn, noisy_sigma0, norm=self.norm)
        self.after_embed_norm = _NormParams(hidden_channels) if self.norm else None
        self.final_conv_acts = None
        self.final_conv_grads = None
        self._desc = _lib.make_desc(hidden_channels, num_layers, num_head_layers)
        self._flat = None
        self._flat_version = 0
        self._slices = None
        self._ws = {}
        self._flat_grad = None
        self._grad_views = {}
        self._p2p = None
        self._noisy = self._noisy_grad = None
        self._noisy_params = []
        self._norm_flat = self._norm_grad = None
        self._norm_params = []
        self._norm_views = []
        self._shape_sig = self._shape_signature()
        self.last_csr = None

    def _uses_lw(self):
        return self.norm or self.path == "lw" or (self.path == "auto" and self.max_graph_nodes > self.tile_rows)

    def _ordered_norm_params(self):
        """hexgnn_lw_sizes().norm_floats block: per LayerNorm weight[H] then bias[H]; body layers, after_embed_norm, maker
        head layers, breaker head layers."""
        mods = list(self.gnn.norms) + [self.after_embed_norm] + list(self.maker_head.gnn.norms) + \
            list(self.breaker_head.gnn.norms)
        ps = []
        for m in mods:
            ps += [m.weight, m.bias]
        return ps

    # ---- parameter order == include/hexgnn.h flat layout ---------------------------------------------
    def _ordered_params(self):
        ps = []
        for c in self.gnn.convs:
            ps += [c.lin_l.weight, c.lin_l.bias, c.lin_r.weight]
        for head in (self.maker_head, self.breaker_head):
            for c in head.gnn.convs:
                ps += [c.lin_l.weight, c.lin_l.bias, c.lin_r.weight]
            ps += [head.linear._w_eff, head.linear._b_eff] if self.noisy_dqn else [head.linear.weight, head.linear.bias]
            ps += [head.value_head.layers[0].weight, head.value_head.layers[0].bias,
                   head.value_head.layers[1].weight, head.value_head.layers[1].bias]
        return ps

    def _flatten(self):
        """Move every parameter into one flat fp32 buffer (the C-ABI's `params`) and re-point the
        named parameters at views of it, so optimizers / state_dict keep working on the names."""
        ps = self._ordered_params()
        offs = _lib.param_offsets(self._desc)
        total = int(_lib.lib().hexgnn_param_count(self._desc))
        if len(offs) != len(ps):
            raise _lib.HexGnnError("parameter table mismatch between models.py and libhexgnn.so")
        dev = ps[0].device
        flat = torch.empty(total, dtype=torch.float32, device=dev)
        slices = []
        with torch.no_grad():
            for p, o in zip(ps, offs):
                n = p.numel()
                flat[o:o + n].copy_(p.detach().reshape(-1).to(torch.float32))
                p.data = flat[o:o + n].view(p.shape)
                slices.append((o, n, tuple(p.shape)))
        if slices[-1][0] + slices[-1][1] != total:
            raise _lib.HexGnnError("flat parameter size mismatch")
        self._flat, self._slices = flat, slices
        self._flat_version += 1
        self._ws = {}
        self._flat_grad = None
        self._grad_views = {}
        self._params_cache = ps
        self._norm_flat = self._norm_grad = None
        self._norm_params, self._norm_views = [], []
        if self.norm:
            nps = self._ordered_norm_params()
            H = self.hidden_channels
            nf = torch.empty(len(nps) * H, dtype=torch.float32, device=dev)
            with torch.no_grad():
                for i, q in enumerate(nps):
                    nf[i * H:(i + 1) * H].copy_(q.detach().to(torch.float32))
                    q.data = nf[i * H:(i + 1) * H]
            self._norm_flat, self._norm_params = nf, nps
        self._noisy = self._noisy_grad = None
        self._noisy_params = []
        if self.noisy_dqn:
            # mu / sigma / epsilon of both heads in one [2, 3, H+1] buffer (weight row, then the bias): the C-ABI's `noisy`
            H = self.hidden_channels
            nz = torch.empty(2, 3, H + 1, dtype=torch.float32, device=dev)
            with torch.no_grad():
                for h, head in enumerate((self.maker_head, self.breaker_head)):
                    lin = head.linear
                    for r, (w, b) in enumerate(((lin.weight_mu, lin.bias_mu), (lin.weight_sigma, lin.bias_sigma),
                                                (lin.weight_epsilon, lin.bias_epsilon))):
                        nz[h, r, :H].copy_(w.detach().reshape(-1))
                        nz[h, r, H:].copy_(b.detach().reshape(-1))
                        w.data = nz[h, r, :H].view(1, H)
                        b.data = nz[h, r, H:]
                    self._noisy_params += [(lin.weight_mu, h, 0, True), (lin.bias_mu, h, 0, False),
                                           (lin.weight_sigma, h, 1, True), (lin.bias_sigma, h, 1, False)]
            self._noisy = nz

    def _shape_signature(self):
        return (self.gnn.hidden_channels, len(self.gnn.convs), len(self.maker_head.gnn.convs),
                len(self.breaker_head.gnn.convs), self.maker_head.hidden_channels, self.breaker_head.hidden_channels)

    def _ensure_flat(self):
        sig = self._shape_signature()
        if sig != self._shape_sig:
            # grow_width / grow_depth changed the architecture (GN0/models.py:494-508, train.py:301-309)
            H, L, Lm, Lb, Hm, Hb = sig
            if Lm != Lb or Hm != H or Hb != H:
                raise _lib.HexGnnError(f"maker / breaker heads differ in shape ({Lm} vs {Lb} layers, widths {H}/{Hm}/{Hb}): "
                                       "grow both before the next forward")
            self.hidden_channels, self.num_layers, self.num_head_layers = H, L, Lm
            self._desc = _lib.make_desc(H, L, Lm)
            self._shape_sig = sig
            self._flat = None
        if self._flat is None or self._flat.device != self.gnn.convs[0].lin_l.weight.device or not self._aliased():
            self._flatten()

    def _aliased(self):
        """True while the named parameters still ARE views of the flat buffer.  `copy.deepcopy(model)` / pickling clone
        every Parameter (`Parameter.__deepcopy__`), after which `load_state_dict`, optimizers and target syncs would
        change the names while the kernels kept reading a stale buffer: detected here (two pointer compares per call)
        and repaired by re-flattening."""
        ps, sl, base = getattr(self, "_params_cache", None), self._slices, self._flat.data_ptr()
        if not ps or not sl or len(ps) != len(sl):
            return False
        now = self._ordered_params()
        for i in (0, len(ps) - 1):
            if now[i] is not ps[i] or ps[i].data_ptr() != base + 4 * sl[i][0]:
                return False
        if self._noisy is not None:
            lin = self.maker_head.linear
            if lin.weight_mu.data_ptr() != self._noisy.data_ptr():
                return False
        if self.norm:
            nps = self._ordered_norm_params()
            if self._norm_flat is None or len(nps) != len(self._norm_params) or nps[0] is not self._norm_params[0] or \
                    nps[0].data_ptr() != self._norm_flat.data_ptr() or \
                    nps[-1].data_ptr() != self._norm_flat.data_ptr() + 4 * (len(nps) - 1) * self.hidden_channels:
                return False
        return True

    def __deepcopy__(self, memo):
        """A deep copy starts un-flattened (its first call re-flattens its own parameter clones) and shares no workspace,
        gradient buffer or peer-memory handle with the original."""
        import copy
        cls = self.__class__
        new = cls.__new__(cls)
        memo[id(self)] = new
        skip = {"_flat", "_slices", "_ws", "_flat_grad", "_grad_views", "_p2p", "_noisy", "_noisy_grad", "_noisy_params",
                "_noisy_views", "_params_cache", "last_csr", "final_conv_acts", "final_conv_grads", "_norm_flat",
                "_norm_grad", "_norm_params", "_norm_views"}
        for k, v in self.__dict__.items():
            if k not in skip:
                new.__dict__[k] = copy.deepcopy(v, memo)
        new._flat = new._slices = new._flat_grad = new._p2p = new._noisy = new._noisy_grad = None
        new._ws, new._grad_views, new._noisy_params, new._params_cache = {}, {}, [], None
        new._norm_flat = new._norm_grad = None
        new._norm_params, new._norm_views = [], []
        new.last_csr = new.final_conv_acts = new.final_conv_grads = None
        for p in new.parameters():
            p.grad = None
        return new

    def _apply(self, fn, *args, **kwargs):
        out = super()._apply(fn, *args, **kwargs)
        self._flat = None  # .to()/.cuda()/.float() replace the parameter storages
        return out

    def reflatten(self):
        """Call after replacing parameter tensors by hand (e.g. `p.data = ...`)."""
        self._flat = None

    # ---- flat gradient: every p.grad is a view of one buffer (the C-ABI's `grad_params`) ----------------
    def _flat_grad_buffer(self):
        if self._flat_grad is None or self._flat_grad.device != self._flat.device:
            self._flat_grad = torch.zeros_like(self._flat)
        if self._p2p is not None:
            # peer-mapped, double-buffered: the backward writes where the peers will read; buffer k is reused two steps
            # later, after the next step's barrier has proven that every rank finished reading it
            self._flat_grad = self._p2p["views"][self._p2p["k"]]
        return self._flat_grad

    def _grad_scratch(self):
        w = self._ws.get("gscratch")
        if w is None or w.device != self._flat.device:
            w = torch.empty_like(self._flat)
            self._ws["gscratch"] = w
        return w

    def _grads_in_place(self):
        """True when every p.grad still is our slice of the flat gradient (zero_grad(set_to_none=False)
        or a second backward before the optimizer step): the new gradient must be ADDED."""
        if self._flat_grad is None:
            return False
        base = self._flat_grad.data_ptr()
        n_set = n_skip = 0
        for p, (o, n, _) in zip(self._params_cache, self._slices):
            if not p.requires_grad:     # frozen, or the composed noisy weights (a derived slot, not a parameter)
                n_skip += 1
                continue
            g = p.grad
            if g is None:
                continue
            if g.data_ptr() != base + 4 * o:
                # a foreign gradient tensor: fold it into the flat buffer semantics by copying it in
                self._flat_grad[o:o + n].copy_(g.reshape(-1))
            n_set += 1
        if n_set == 0:
            return False
        if n_set != len(self._slices) - n_skip:
            for p, (o, n, _) in zip(self._params_cache, self._slices):
                if p.grad is None and p.requires_grad:
                    self._flat_grad[o:o + n].zero_()
        self._point_grads()
        return True

    def _point_grads(self):
        fg = self._flat_grad
        key = (fg.data_ptr(), self._flat_version)
        views = self._grad_views.get(key)
        if views is None:
            # the views of a flat buffer are built once per buffer (P2P mode alternates between two buffers): a step
            # then costs one attribute store per parameter instead of two tensor constructions
            views = [fg[o:o + n].view(shp) for (o, n, shp) in self._slices]
            if len(self._grad_views) > 4:
                self._grad_views.clear()
            self._grad_views[key] = views
        for p, g in zip(self._params_cache, views):
            if p.requires_grad:
                p.grad = g

    def zero_grad(self, set_to_none: bool = True):
        """nn.Module.zero_grad over the cached parameter list (the generic version walks the module tree: ~0.1 ms of
        host time per learner step, more than the device spends on the CSR build)."""
        if self._flat is None or getattr(self, "_params_cache", None) is None:
            return super().zero_grad(set_to_none=set_to_none)
        if set_to_none:
            for p in self._params_cache:
                p.grad = None
            for p, *_ in self._noisy_params:
                p.grad = None
            for p in self._norm_params:
                p.grad = None
        elif self._grads_in_place():
            self._flat_grad.zero_()
            if self._noisy_grad is not None:
                self._noisy_grad.zero_()
            if self._norm_grad is not None:
                self._norm_grad.zero_()
        else:
            super().zero_grad(set_to_none=False)

    def reset_noise(self):
        """`Rainbow.reset_noise(net)` (rainbow.py:167-179) for this model: fresh factorised noise in both heads' advantage
        layers (the reference visits the two heads twice, i.e. draws twice; the distribution is the same)."""
        if self.noisy_dqn:
            self.maker_head.linear.reset_noise()
            self.breaker_head.linear.reset_noise()

    def disable_noise(self):
        """`Rainbow.disable_noise(net)` (rainbow.py:181-192): evaluation with the mean weights."""
        if self.noisy_dqn:
            self.maker_head.linear.disable_noise()
            self.breaker_head.linear.disable_noise()

    # ---- noisy-nets advantage layer (`noisy_dqn=True`): compose in front of the kernels, chain rule behind them -------
    def _compose_noisy(self):
        if self.noisy_dqn:
            ops.noisy_compose(self._desc, self._flat, self._noisy)

    def _noisy_backward(self, grad):
        """d_mu = g, d_sigma = g * epsilon from this backward's flat gradient; `+=` when the noisy parameters' .grad still
        are our views (a second backward before the optimizer step), like autograd."""
        if not self.noisy_dqn:
            return
        H = self.hidden_channels
        if self._noisy_grad is None or self._noisy_grad.device != self._noisy.device:
            self._noisy_grad = torch.zeros(2, 2, H + 1, dtype=torch.float32, device=self._noisy.device)
            self._noisy_views = [self._noisy_grad[h, r, :H].view(1, H) if is_w else self._noisy_grad[h, r, H:]
                                 for (_, h, r, is_w) in self._noisy_params]
        views = self._noisy_views
        acc = all(p.grad is not None and p.grad.data_ptr() == v.data_ptr() for (p, *_), v in zip(self._noisy_params, views))
        ops.noisy_grads(self._desc, grad, self._noisy, self._noisy_grad, acc)
        for (p, *_), v in zip(self._noisy_params, views):
            if p.requires_grad:
                p.grad = v

    @property
    def flat_gradient(self):
        """The flat fp32 gradient all `p.grad` alias (None before the first backward)."""
        return self._flat_grad

    def enable_p2p_allreduce(self, group=None, mode="auto"):
        """Put the flat gradient into NVLink peer-mapped (symmetric) memory so `allreduce_gradients` runs the library's own
        all-reduce instead of NCCL.  mode "mc" (or "auto" where the allocation has a multicast address): two-shot IN PLACE
        through the NVSwitch (`hexgnn_allreduce_mc`: multimem.ld_reduce of this rank's slice, multimem.st of the sums into
        every rank's buffer); mode "pull": the one-shot kernel (`hexgnn_allreduce_p2p`: every rank reads all peers and sums
        in rank order).  Collective: every rank of `group` must call it, after `.to(device)` and before the first
        backward.  Returns False (and keeps NCCL) where symmetric memory is unavailable."""
        import torch.distributed as dist
        self._ensure_flat()
        if not (dist.is_initialized() and dist.get_world_size(group) > 1 and self._flat.is_cuda):
            return False
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        n = self._flat.numel()
        npad = (n + 4 * world - 1) // (4 * world) * (4 * world)
        try:
            import torch.distributed._symmetric_memory as symm
            g = group if group is not None else dist.group.WORLD
            bufs, hdls = [], []
            for _ in range(2):
                buf = symm.empty(npad, dtype=torch.float32, device=self._flat.device)
                hdls.append(symm.rendezvous(buf, g))
                buf.zero_()
                bufs.append(buf)
            flags = symm.empty(max(world, 4), dtype=torch.int32, device=self._flat.device)
            fhdl = symm.rendezvous(flags, g)
            flags.zero_()
        except Exception as e:  # noqa: BLE001 -- any failure here means "no peer memory": stay on NCCL, say so
            import warnings
            warnings.warn(f"gnn_hex_b200: symmetric memory unavailable ({e!r}); gradients stay on NCCL all-reduce")
            return False
        has_mc = all(int(getattr(h, "multicast_ptr", 0) or 0) != 0 for h in hdls)
        if mode == "mc" and not has_mc:
            raise _lib.HexGnnError("enable_p2p_allreduce(mode='mc'): the symmetric allocation has no multicast address here")
        use_mc = has_mc and mode in ("auto", "mc")
        self._p2p = {"bufs": bufs, "views": [b[:n] for b in bufs], "hdls": hdls, "red": None if use_mc else torch.zeros_like(self._flat),
                     "k": 0, "mc": use_mc, "flags": flags, "fhdl": fhdl, "epoch": 0, "rank": rank, "npad": npad,
                     "counter": torch.zeros(1, dtype=torch.int32, device=self._flat.device)}
        torch.cuda.synchronize(self._flat.device)
        hdls[0].barrier(channel=0)      # every rank's buffers and flag words are zeroed before anybody signals
        self._flat_grad = None          # next backward adopts a peer-mapped buffer
        for p in self._params_cache:
            p.grad = None
        return True

    def allreduce_gradients(self, group=None):
        """Data-parallel learner step (BASELINE config 5): ONE all-reduce (sum) of the flat gradient; the caller
        scales its loss by 1/world_size.  After `enable_p2p_allreduce()`: one cross-rank barrier, then the library's own
        kernel over NVLink peer memory -- in place through the switch (`hexgnn_allreduce_mc`) or, in "pull" mode,
        `hexgnn_allreduce_p2p` into a private buffer the `p.grad` views are re-pointed to; the two peer-mapped buffers
        alternate so no second barrier is needed.  Gradients accumulated over several backwards (p.grad kept alive) fall
        back to NCCL.  No reference counterpart."""
        import torch.distributed as dist
        if self._flat_grad is None or not dist.is_initialized() or dist.get_world_size(group) <= 1:
            return
        p2p = self._p2p
        if p2p is not None and self._flat_grad.data_ptr() == p2p["bufs"][p2p["k"]].data_ptr():
            k = p2p["k"]
            hdl = p2p["hdls"][k]
            if p2p["mc"]:
                p2p["epoch"] += 1       # the kernel carries both barriers (flag words in peer-mapped memory)
                ops.allreduce_mc(hdl.multicast_ptr, p2p["fhdl"].buffer_ptrs_dev, p2p["rank"], hdl.world_size, p2p["npad"],
                                 p2p["epoch"], p2p["counter"])          # p.grad already alias buffer k: reduced in place
            else:
                hdl.barrier(channel=0)  # every rank's backward has written its gradient into its buffer k
                ops.allreduce_p2p(hdl.buffer_ptrs_dev, hdl.world_size, self._flat_grad.numel(), p2p["red"])
                self._flat_grad = p2p["red"]          # p.grad now alias the reduced (private) buffer
                self._point_grads()
            p2p["k"] = k ^ 1
            self._allreduce_noisy(group)
            return
        dist.all_reduce(self._flat_grad, op=dist.ReduceOp.SUM, group=group)
        self._allreduce_noisy(group)

    def _allreduce_noisy(self, group=None):
        """noisy_dqn: the advantage layers' d_mu / d_sigma live outside the flat gradient (`_noisy_grad`, formed from the
        LOCAL gradient and the rank's own epsilon by `_noisy_backward`), so they need their own (tiny) all-reduce --
        otherwise every rank would step mu / sigma with its shard's gradient and the replicas would drift apart."""
        import torch.distributed as dist
        if self.noisy_dqn and self._noisy_grad is not None:
            dist.all_reduce(self._noisy_grad, op=dist.ReduceOp.SUM, group=group)
        if self.norm and self._norm_grad is not None:   # LayerNorm affine gradients: a second small flat buffer
            dist.all_reduce(self._norm_grad, op=dist.ReduceOp.SUM, group=group)

    @property
    def flat_parameters(self):
        self._ensure_flat()
        return self._flat

    def _fwd_ws(self, csr):
        w = self._ws.get("fwd")
        need = _lib.sizes(self._desc, csr.N, csr.E, csr.B, csr.tile_rows).fwd_ws_bytes // 4
        if w is None or w.numel() < need or w.device != csr.blob.device:
            w = torch.empty(need, dtype=torch.float32, device=csr.blob.device)
            self._ws["fwd"] = w
        return w

    def _bwd_ws(self, csr):
        w = self._ws.get("bwd")
        need = _lib.sizes(self._desc, csr.N, csr.E, csr.B, csr.tile_rows).bwd_ws_bytes // 4
        if w is None or w.numel() < need or w.device != csr.blob.device:
            w = torch.empty(need, dtype=torch.float32, device=csr.blob.device)
            self._ws["bwd"] = w
        return w

    def _lw_ws(self, csr, which):
        sz = _lib.lw_sizes(self._desc, csr.N, csr.E, csr.B, self.norm)
        need = ((sz.fwd_ws_bytes if which == "fwd" else sz.bwd_ws_bytes) + 3) // 4
        w = self._ws.get("lw_" + which)
        if w is None or w.numel() < need or w.device != csr.blob.device:
            w = torch.empty(need, dtype=torch.float32, device=csr.blob.device)
            self._ws["lw_" + which] = w
        return w

    def _lw_backward(self, x, csr, flags, d_nodes, d_value, stash):
        """hexgnn_backward_lw into the flat gradient (+ the LayerNorm gradient block), with the same accumulate-or-assign
        semantics as the fused route."""
        accumulate = self._grads_in_place()
        out = self._grad_scratch() if accumulate else self._flat_grad_buffer()
        gn = None
        if self.norm:
            if self._norm_grad is None or self._norm_grad.device != self._norm_flat.device:
                self._norm_grad = torch.zeros_like(self._norm_flat)
                H = self.hidden_channels
                self._norm_views = [self._norm_grad[i * H:(i + 1) * H] for i in range(len(self._norm_params))]
            gn = self._ws.get("gnorm")
            if gn is None or gn.device != self._norm_flat.device or gn.numel() != self._norm_flat.numel():
                gn = self._ws["gnorm"] = torch.empty_like(self._norm_flat)
        grad, gnorm, d_emb = ops.lw_backward(self._desc, self._flat, self._norm_flat, x, csr, flags, d_nodes, d_value, stash,
                                             want_d_embeds=self.capture_embed_grads, ws=self._lw_ws(csr, "bwd"),
                                             grad_out=out, grad_norm=gn)
        if accumulate:
            self._flat_grad.add_(grad)
        else:
            self._point_grads()
        if self.norm:
            acc = all(p.grad is not None and p.grad.data_ptr() == v.data_ptr()
                      for p, v in zip(self._norm_params, self._norm_views))
            if acc:
                self._norm_grad.add_(gnorm)
            else:
                self._norm_grad.copy_(gnorm)
            for p, v in zip(self._norm_params, self._norm_views):
                if p.requires_grad:
                    p.grad = v
        self._noisy_backward(grad)
        if d_emb is not None:
            self.final_conv_grads = d_emb[:, :self.hidden_channels]
        return grad

    # ---- reference API ---------------------------------------------------------------------------------
    def build_csr(self, edge_index, graph_indices=None, ptr=None, num_nodes=None, num_graphs=None):
        if num_graphs is None and ptr is None and graph_indices is not None:
            # the reference does the same sync: batch_size = int(graph_indices.max()) + 1 (models.py:576)
            num_graphs = int(graph_indices.max()) + 1
        return ops.build_csr(edge_index, graph_indices, ptr, num_nodes=num_nodes, num_graphs=num_graphs,
                             tile_rows=self.tile_rows)

    def forward(self, x, edge_index, graph_indices: Optional[torch.Tensor] = None, ptr: Optional[torch.Tensor] = None,
                set_cache: bool = False, advantages_only=False, seperate=False, num_graphs: Optional[int] = None,
                csr: Optional[ops.CsrBlob] = None):
        self._ensure_flat()
        self._compose_noisy()
        if csr is None:
            csr = self.build_csr(edge_index, graph_indices, ptr, num_nodes=int(x.shape[0]), num_graphs=num_graphs)
        flags = 0
        if advantages_only:
            flags |= ops.HEXGNN_ADV_ONLY
        elif seperate:
            flags |= ops.HEXGNN_SEPARATE
        if self.path == "simt":
            flags |= ops.HEXGNN_PATH_SIMT
        elif self.path == "tc":
            flags |= ops.HEXGNN_PATH_TC
        self.final_conv_acts = None
        self.last_csr = csr
        anchor = self._params_cache[0]
        if self._uses_lw():
            flags &= ops.HEXGNN_ADV_ONLY | ops.HEXGNN_SEPARATE
            if torch.is_grad_enabled() and anchor.requires_grad:
                out_nodes, out_value = _LwQ.apply(self, x, csr, flags, anchor)
            else:
                out_nodes, out_value, _, _ = ops.lw_forward(self._desc, self._flat, self._norm_flat, x, csr, flags,
                                                            want_stash=False, ws=self._lw_ws(csr, "fwd"))
        elif torch.is_grad_enabled() and anchor.requires_grad:
            out_nodes, out_value = _FusedQ.apply(self, x, csr, flags, anchor)
        else:
            if self.inference_precision == "tf32x1" and self.path != "simt":
                flags |= ops.HEXGNN_TF32X1
            out_nodes, out_value, _, _ = ops.fused_forward(self._desc, self._flat, x, csr, flags, want_stash=False,
                                                           ws=self._fwd_ws(csr))
            flags &= ~ops.HEXGNN_TF32X1
            if self.max_graph_nodes > self.tile_rows:   # path == "tc" / "simt" forced: CUDA-core kernels for the oversize graphs
                ops.forward_big(self._desc, self._flat, x, csr, flags, self.max_graph_nodes, out_nodes, out_value)
        if self.strict:
            csr.raise_on_error()
        if advantages_only:
            return out_nodes.unsqueeze(1)           # [N,1] (models.py:567-569)
        if seperate:
            return out_value, out_nodes             # (value[B], adv - mean[N]) (models.py:580-581)
        return out_nodes                            # [N] (models.py:584)

    def train_step(self, x, edge_index, graph_indices=None, ptr=None, action=None, target=None, weight=None,
                   kind="mse", scale=1.0, num_graphs: Optional[int] = None, csr: Optional[ops.CsrBlob] = None):
        """The learner's ``loss.backward()`` block in one call (rainbow.py:301-316):

            q = model(x, edge_index, graph_indices, ptr); td_est = q[action + ptr[:-1]]
            loss = scale * mean(weight * l(target, td_est)); loss.backward()

        Returns ``(loss, td_err, q)``; every ``p.grad`` is set (accumulated when gradients are still in place) exactly as
        by the autograd route, bit for bit.  Because the loss gradient is derived inside the backward kernel, forward,
        dH chain and dW GEMM overlap (hexgnn_train_step).  ``ptr`` is required; tensor-core path, default Q output."""
        if ptr is None or action is None or target is None:
            raise _lib.HexGnnError("train_step: ptr, action and target are required")
        self._ensure_flat()
        self._compose_noisy()
        if csr is None:
            csr = self.build_csr(edge_index, graph_indices, ptr, num_nodes=int(x.shape[0]), num_graphs=num_graphs)
        self.last_csr = csr
        self.final_conv_acts = None
        if self._uses_lw():
            # same three results through the layer-at-a-time path: forward (stash), hexgnn_td_loss, hexgnn_backward_lw
            q, _, stash, _ = ops.lw_forward(self._desc, self._flat, self._norm_flat, x, csr, 0, want_stash=True,
                                            ws=self._lw_ws(csr, "fwd"))
            with torch.enable_grad():
                qd = q.detach().requires_grad_(True)
                loss, td_err = ops.td_loss(qd, ptr, action, target, weight, kind, scale)
                (d_q,) = torch.autograd.grad(loss, qd)
            self._lw_backward(x, csr, 0, d_q, None, stash)
            if self.strict:
                csr.raise_on_error()
            return loss.detach(), td_err, q
        accumulate = self._grads_in_place()
        out = self._grad_scratch() if accumulate else self._flat_grad_buffer()
        loss, td_err, q, grad = ops.fused_train_step(self._desc, self._flat, x, csr, ptr, action, target, weight, kind, scale,
                                                     grad_out=out, ws=self._bwd_ws(csr))
        if accumulate:
            self._flat_grad.add_(grad)
        else:
            self._point_grads()
        self._noisy_backward(grad)
        if self.strict:
            csr.raise_on_error()
        return loss, td_err, q

    @torch.no_grad()
    def copy_parameters_from(self, other):
        """`q_target.load_state_dict(q_policy.state_dict())` (rainbow.py:163-164) as ONE device-to-device copy of the flat
        parameter buffer."""
        self._ensure_flat()
        other._ensure_flat()
        if other._flat.numel() != self._flat.numel():
            raise _lib.HexGnnError("copy_parameters_from: different architectures")
        self._flat.copy_(other._flat, non_blocking=True)
        if self.noisy_dqn != other.noisy_dqn:
            raise _lib.HexGnnError("copy_parameters_from: noisy_dqn differs")
        if self.noisy_dqn:
            self._noisy.copy_(other._noisy, non_blocking=True)   # mu, sigma and the epsilon buffers (state_dict has all three)
        if self.norm != other.norm:
            raise _lib.HexGnnError("copy_parameters_from: norm differs")
        if self.norm:
            self._norm_flat.copy_(other._norm_flat, non_blocking=True)   # the LayerNorm weights live in their own flat block

    def simple_forward(self, data):
        if getattr(data, "batch", None) is not None:
            return self.forward(data.x, data.edge_index, data.batch, getattr(data, "ptr", None))
        return self.forward(data.x, data.edge_index)

    # no-op cache API: cached_norm=False in modern_two_headed (models.py:934,943)
    def export_norm_cache(self, *args):
        return []

    def import_norm_cache(self, *args):
        return None

    def grow_depth(self, additional_layers):
        """GN0/models.py:494-495 (train.py:302,304): more body layers, function unchanged.  The fused kernels pick the new
        shape up at the next call (`_ensure_flat` re-reads the module structure)."""
        self.gnn.grow_depth(additional_layers)

    def grow_width(self, new_width):
        """GN0/models.py:497-508 (train.py:301,303): wider body and heads."""
        self.gnn.grow_width(new_width)
        self.maker_head.grow_width(new_width, new_in_channels=new_width)
        self.breaker_head.grow_width(new_width, new_in_channels=new_width)


class FusedAdam:
    """`nn.utils.clip_grad_norm_` + `torch.optim.Adam(model.parameters(), lr, eps=...)` of the learner
    (rainbow.py:79,320,325-326) as two launches over the model's flat parameter / gradient buffers.

        opt = FusedAdam(model, lr=args.lr, eps=args.adam_eps, max_grad_norm=args.max_grad_norm)
        loss, td_err, q = model.train_step(...)        # or loss.backward()
        grad_norm = opt.step(x)                        # x: the batch features (x[0,2] tells which head was trained)
        model.zero_grad()

    `skip_untaken=False` reproduces the reference environment (torch 1.13: `opt.zero_grad()` keeps zero gradients, so a
    head that is not trained in a step still receives its Adam momentum); True gives torch >= 2 `set_to_none` behaviour.
    `state_dict()` / `load_state_dict()` carry exp_avg, exp_avg_sq and the three step counters."""

    def __init__(self, model, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, max_grad_norm=0.0, skip_untaken=False):
        self.model, self.lr, self.betas, self.eps = model, float(lr), tuple(betas), float(eps)
        self.max_grad_norm, self.skip_untaken = float(max_grad_norm), bool(skip_untaken)
        self._version = -1
        self.exp_avg = self.exp_avg_sq = self.state = self._ws = None

    def _ensure(self):
        m = self.model
        m._ensure_flat()
        if m.noisy_dqn:
            raise NotImplementedError("FusedAdam steps the flat parameter buffer; with noisy_dqn=True the advantage layer's "
                                      "mu / sigma live outside it: use torch.optim.Adam(model.parameters())")
        if m.norm:
            raise NotImplementedError("FusedAdam steps the flat parameter buffer; with norm=True the LayerNorm weights live "
                                      "outside it: use torch.optim.Adam(model.parameters())")
        if self._version != m._flat_version or self.exp_avg is None or self.exp_avg.device != m._flat.device:
            old = (self.exp_avg, self.exp_avg_sq, self.state)
            self.exp_avg = torch.zeros_like(m._flat)
            self.exp_avg_sq = torch.zeros_like(m._flat)
            self.state = torch.zeros(4, dtype=torch.int32, device=m._flat.device)
            if old[0] is not None and old[0].numel() == m._flat.numel():   # parameters moved (.to()): keep the moments
                self.exp_avg.copy_(old[0]); self.exp_avg_sq.copy_(old[1]); self.state.copy_(old[2])
            self._version = m._flat_version

    @torch.no_grad()
    def step(self, x):
        self._ensure()
        g = self.model.flat_gradient
        if g is None:
            raise _lib.HexGnnError("FusedAdam.step: no gradient (run train_step or backward first)")
        return ops.adam_step(self.model._desc, self.model._flat, g, self.exp_avg, self.exp_avg_sq, self.state, x, self.lr,
                             self.betas, self.eps, self.max_grad_norm, self.skip_untaken)

    def zero_grad(self, set_to_none=True):
        self.model.zero_grad(set_to_none=set_to_none)

    def state_dict(self):
        self._ensure()
        return {"exp_avg": self.exp_avg.clone(), "exp_avg_sq": self.exp_avg_sq.clone(), "steps": self.state.clone(),
                "lr": self.lr, "betas": self.betas, "eps": self.eps, "max_grad_norm": self.max_grad_norm}

    def load_state_dict(self, sd):
        self._ensure()
        self.exp_avg.copy_(sd["exp_avg"]); self.exp_avg_sq.copy_(sd["exp_avg_sq"]); self.state.copy_(sd["steps"])
        self.lr, self.betas, self.eps = float(sd["lr"]), tuple(sd["betas"]), float(sd["eps"])
        self.max_grad_norm = float(sd.get("max_grad_norm", self.max_grad_norm))


def get_pre_defined(name, args=None):
    """`GN0/models.py:892` for the one name this package implements."""
    if name != "modern_two_headed":
        raise NotImplementedError(f"gnn_hex_b200 implements only 'modern_two_headed', not {name!r}")
    return ModernTwoHeaded(args.hidden_channels, args.num_layers,
                           args.num_head_layers if hasattr(args, "num_head_layers") else 2,
                           noisy_dqn=bool(getattr(args, "noisy_dqn", False)),
                           noisy_sigma0=float(getattr(args, "noisy_sigma0", 0.5)),
                           norm=bool(getattr(args, "norm", False)))
