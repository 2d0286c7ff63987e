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


class _SageParams(nn.Module):
    """Holder named like `cachify_gnn(GraphSAGE)` (GN0/models.py:144-160): `.convs`, plus the
    attributes callers read (`hidden_channels` train.py:301, `has_cache` rainbow.py:344)."""
    supports_cache = True

    def __init__(self, in_channels, hidden, num_layers):
        super().__init__()
        self.in_channels = in_channels
        self.hidden_channels = hidden
        self.num_layers = num_layers
        self.has_cache = False
        self.convs = nn.ModuleList([_Conv(in_channels if i == 0 else hidden, hidden) for i in range(num_layers)])


class _MLPParams(nn.Module):
    """Holder named like `MLP(hidden//2, 1, 4*hidden, 1)` (GN0/models.py:36-49): `.layers.{0,1}`."""

    def __init__(self, hidden, num_input, num_output):
        super().__init__()
        self.layers = nn.ModuleList([nn.Linear(num_input, hidden), nn.Linear(hidden, num_output)])


class _HeadParams(nn.Module):
    """Holder named like `HeadNetwork` (GN0/models.py:318-334): gnn, linear, value_head."""
    supports_cache = True

    def __init__(self, hidden, num_layers):
        super().__init__()
        self.gnn = _SageParams(hidden, hidden, num_layers)
        self.linear = nn.Linear(hidden, 1)
        self.value_head = _MLPParams(hidden // 2, 4 * hidden, 1)
        self.hidden_channels = hidden


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
        grad, d_emb = ops.fused_backward(model._desc, model._flat, x, ctx.csr, ctx.flags, d_nodes,
                                         d_value if use_dv else None, ctx.stash,
                                         want_d_embeds=model.capture_embed_grads, ws=model._bwd_ws(ctx.csr),
                                         grad_out=out)
        if accumulate:
            model._flat_grad.add_(grad)
        else:
            model._point_grads()
        if d_emb is not None:
            model.final_conv_grads = d_emb[:, :model.hidden_channels]
        return None, None, None, None, None


class ModernTwoHeaded(nn.Module):
    """Drop-in for `get_pre_defined("modern_two_headed", args)` with norm=False, noisy_dqn=False.

    forward(x, edge_index, graph_indices=None, ptr=None, set_cache=False, advantages_only=False,
            seperate=False)            -- GN0/models.py:537
    simple_forward(data)               -- GN0/models.py:586-590
    Extra, non-reference keyword: `num_graphs` (avoids the host sync of `int(graph_indices.max())+1`,
    models.py:576, when `ptr` is not passed) and `csr` (reuse a CSR built by `build_csr`).
    """
    supports_cache = True  # class attribute in the reference (models.py:148,486); read at train.py:98

    def __init__(self, hidden_channels: int, num_layers: int, num_head_layers: int = 2, tile_rows: int = 128,
                 strict: bool = False, path: str = "auto", max_graph_nodes: int = 0):
        super().__init__()
        self.hidden_channels = int(hidden_channels)
        self.num_layers = int(num_layers)
        self.num_head_layers = int(num_head_layers)
        self.tile_rows = int(tile_rows)
        self.strict = strict                  # True: sync + raise on device-side status after each forward
        self.path = path                      # "auto" | "simt" | "tc"
        self.capture_embed_grads = False      # True: fill final_conv_grads (rl_loop/explain.py:41)
        # > tile_rows: inference calls also run the oversize-graph path (hexgnn_forward_big) for graphs of up to this
        # many nodes (board-size transfer: 13x13 -> 171, evaluate_elo.py:229-251); 0 = such graphs are flagged only
        self.max_graph_nodes = int(max_graph_nodes)
        self.gnn = _SageParams(2, hidden_channels, num_layers)
        self.maker_head = _HeadParams(hidden_channels, num_head_layers)
        self.breaker_head = _HeadParams(hidden_channels, num_head_layers)
        self.after_embed_norm = None
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
        self.last_csr = None

    # ---- parameter order == include/hexgnn.h flat layout ---------------------------------------------
    def _ordered_params(self):
        ps = []
        for c in self.gnn.convs:
            ps += [c.lin_l.weight, c.lin_l.bias, c.lin_r.weight]
        for head in (self.maker_head, self.breaker_head):
            for c in head.gnn.convs:
                ps += [c.lin_l.weight, c.lin_l.bias, c.lin_r.weight]
            ps += [head.linear.weight, head.linear.bias,
                   head.value_head.layers[0].weight, head.value_head.layers[0].bias,
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

    def _ensure_flat(self):
        if self._flat is None or self._flat.device != self.gnn.convs[0].lin_l.weight.device:
            self._flatten()

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
            self._flat_grad = self._p2p["bufs"][self._p2p["k"]]
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
        n_set = 0
        for p, (o, n, _) in zip(self._params_cache, self._slices):
            g = p.grad
            if g is None:
                continue
            if g.data_ptr() != base + 4 * o:
                # a foreign gradient tensor: fold it into the flat buffer semantics by copying it in
                self._flat_grad[o:o + n].copy_(g.reshape(-1))
            n_set += 1
        if n_set == 0:
            return False
        if n_set != len(self._slices):
            for p, (o, n, _) in zip(self._params_cache, self._slices):
                if p.grad is None:
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
        elif self._grads_in_place():
            self._flat_grad.zero_()
        else:
            super().zero_grad(set_to_none=False)

    @property
    def flat_gradient(self):
        """The flat fp32 gradient all `p.grad` alias (None before the first backward)."""
        return self._flat_grad

    def enable_p2p_allreduce(self, group=None):
        """Put the flat gradient into NVLink peer-mapped (symmetric) memory so `allreduce_gradients` can run the
        library's own one-shot all-reduce kernel (`hexgnn_allreduce_p2p`: every rank reads all peers over NVLink /
        NVSwitch and sums in rank order) instead of NCCL.  Collective: every rank of `group` must call it, after
        `.to(device)` and before the first backward.  Returns False (and keeps NCCL) where symmetric memory is
        unavailable."""
        import torch.distributed as dist
        self._ensure_flat()
        if not (dist.is_initialized() and dist.get_world_size(group) > 1 and self._flat.is_cuda):
            return False
        try:
            import torch.distributed._symmetric_memory as symm
            g = group if group is not None else dist.group.WORLD
            bufs, hdls = [], []
            for _ in range(2):
                buf = symm.empty(self._flat.numel(), dtype=torch.float32, device=self._flat.device)
                hdls.append(symm.rendezvous(buf, g))
                buf.zero_()
                bufs.append(buf)
        except Exception as e:  # noqa: BLE001 -- any failure here means "no peer memory": stay on NCCL, say so
            import warnings
            warnings.warn(f"gnn_hex_b200: symmetric memory unavailable ({e!r}); gradients stay on NCCL all-reduce")
            return False
        self._p2p = {"bufs": bufs, "hdls": hdls, "red": torch.zeros_like(self._flat), "k": 0}
        self._flat_grad = None          # next backward adopts a peer-mapped buffer
        for p in self._params_cache:
            p.grad = None
        return True

    def allreduce_gradients(self, group=None):
        """Data-parallel learner step (BASELINE config 5): ONE all-reduce (sum) of the flat gradient; the caller
        scales its loss by 1/world_size.  After `enable_p2p_allreduce()`: one cross-rank barrier, then
        `hexgnn_allreduce_p2p` (own kernel over NVLink peer memory, rank-ordered sum) into a private buffer the
        `p.grad` views are re-pointed to; the two peer-mapped buffers alternate so no second barrier is needed.
        Gradients accumulated over several backwards (p.grad kept alive) fall back to NCCL.  No reference counterpart."""
        import torch.distributed as dist
        if self._flat_grad is None or not dist.is_initialized() or dist.get_world_size(group) <= 1:
            return
        p2p = self._p2p
        if p2p is not None and self._flat_grad.data_ptr() == p2p["bufs"][p2p["k"]].data_ptr():
            k = p2p["k"]
            hdl = p2p["hdls"][k]
            hdl.barrier(channel=0)      # every rank's backward has written its gradient into its buffer k
            ops.allreduce_p2p(hdl.buffer_ptrs_dev, hdl.world_size, self._flat_grad.numel(), p2p["red"])
            self._flat_grad = p2p["red"]          # p.grad now alias the reduced (private) buffer
            self._point_grads()
            p2p["k"] = k ^ 1
            return
        dist.all_reduce(self._flat_grad, op=dist.ReduceOp.SUM, group=group)

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
        if torch.is_grad_enabled() and anchor.requires_grad:
            out_nodes, out_value = _FusedQ.apply(self, x, csr, flags, anchor)
        else:
            out_nodes, out_value, _, _ = ops.fused_forward(self._desc, self._flat, x, csr, flags, want_stash=False,
                                                           ws=self._fwd_ws(csr))
            if self.max_graph_nodes > self.tile_rows:
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
        if csr is None:
            csr = self.build_csr(edge_index, graph_indices, ptr, num_nodes=int(x.shape[0]), num_graphs=num_graphs)
        self.last_csr = csr
        self.final_conv_acts = None
        accumulate = self._grads_in_place()
        out = self._grad_scratch() if accumulate else self._flat_grad_buffer()
        loss, td_err, q, grad = ops.fused_train_step(self._desc, self._flat, x, csr, ptr, action, target, weight, kind, scale,
                                                     grad_out=out, ws=self._bwd_ws(csr))
        if accumulate:
            self._flat_grad.add_(grad)
        else:
            self._point_grads()
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

    def simple_forward(self, data):
        if getattr(data, "batch", None) is not None:
            return self.forward(data.x, data.edge_index, data.batch, getattr(data, "ptr", None))
        return self.forward(data.x, data.edge_index)

    # no-op cache API: cached_norm=False in modern_two_headed (models.py:934,943)
    def export_norm_cache(self, *args):
        return []

    def import_norm_cache(self, *args):
        return None

    def grow_width(self, *a, **k):
        raise NotImplementedError("grow_width/grow_depth (GN0/models.py:494-508) are outside the fused path")

    grow_depth = grow_width


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
    if getattr(args, "norm", False) or getattr(args, "noisy_dqn", False):
        raise NotImplementedError("norm=True / noisy_dqn=True are not on the fused path (SURVEY.md §8 f4)")
    return ModernTwoHeaded(args.hidden_channels, args.num_layers,
                           args.num_head_layers if hasattr(args, "num_head_layers") else 2)
