"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the ``modern_two_headed`` hot path of GNN_Hex.

Pure torch (no torch_geometric / torch_scatter: neither is installable here), fp32 or fp64.
Every function cites the reference lines it restates (paths relative to /root/reference).

Parity status.  The reference holds NO golden vector for the model numerics, so the
pin is made here: the reference's OWN module code (GN0/models.py, unmodified, where it lies) is executed on a
test-only stand-in for torch_geometric / torch_scatter (oracle/pyg_stub/) and (a) its outputs and gradients
on the committed positions are committed as golden vectors (tests/golden/model_*.npz, made by
tests/golden/make_model_fixtures.py) which this oracle must reproduce (tests/test_reference_pin.py, runs
anywhere), and (b) wherever /root/reference is mounted the same tests run the reference classes live
(`DuellingTwoHeaded`, `cachify_gnn(GraphSAGE)`, `HeadNetwork`, `MLP`, `FactorizedNoisyLinear`, growth).
What stays **unpinned** is only the third-party PRIMITIVES the stub restates: PyG 2.2.0
`SAGEConv.propagate` (the reference's own restatement: GN0/torch_script_models.py:52-73), `LayerNorm`
and torch_scatter 2.1.0 `scatter` -- those libraries (environment.yml:139,151) are not installable here.
Further self-checks: a hand-computed known answer on the reference's ``Debugging_manager`` graph
(graph_game/multi_env_manager.py:170-175), fp64 gradcheck, batched == per-graph, permutation
equivariance and edge-order invariance (tests/test_oracle.py).  The graph half (topology, edge_index,
collate) is pinned by fixtures from the reference's own C++ game compiled unchanged
(oracle/refgame/Makefile -> oracle/_ref/hexgen).
"""
from __future__ import annotations

import math
from types import SimpleNamespace
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

# --------------------------------------------------------------------------------------
# graph side
# --------------------------------------------------------------------------------------


def hex_board_undirected_edges(size: int) -> List[Tuple[int, int]]:
    """Empty-board Shannon graph of a size x size Hex board, red graph, no worthless edges.

    Restates graph_game/hex_board_game.py:214-232 (``graph_from_board(redgraph=True)``):
    vertices 0,1 are the terminals, square i is vertex i+2.
    """
    sq = size
    squares = size * size
    edges: List[Tuple[int, int]] = []
    for i in range(squares):
        v = i + 2
        if i < sq:
            edges.append((v, 0))
        if i // sq == sq - 1:
            edges.append((v, 1))
        if i % sq > 0 and (sq <= i <= squares - sq):
            edges.append((v, v - 1))
        if i >= sq:
            edges.append((v, v - sq))
            if i % sq != sq - 1:
                edges.append((v, v - sq + 1))
    return edges


class SimpleShannonGame:
    """Self-contained fallback position generator (no dead/captured removal).

    Restates the rule core of graph_game/shannon_node_switching_game.py:80-116: a maker move
    on v joins all of v's neighbours pairwise (unless both already touch the same terminal)
    and removes v; a breaker move removes v.  Terminal bookkeeping (``_fix_teminal_connections``)
    and dead/captured pruning are NOT restated -- fixtures for the benchmark come from the
    compiled reference game instead; this class exists so the GPU box (no /root/reference)
    can still synthesise arbitrary many valid symmetric graphs.
    """

    def __init__(self, size: int):
        self.size = size
        n = size * size + 2
        self.adj: Dict[int, set] = {v: set() for v in range(n)}
        for a, b in hex_board_undirected_edges(size):
            self.adj[a].add(b)
            self.adj[b].add(a)
        self.maker_turn = True

    def actions(self) -> List[int]:
        return [v for v in self.adj if v > 1]

    def make_move(self, v: int) -> None:
        neigh = sorted(self.adj[v])
        if self.maker_turn:
            for i, a in enumerate(neigh):
                for b in neigh[i + 1:]:
                    both_t0 = 0 in self.adj[a] and 0 in self.adj[b]
                    both_t1 = 1 in self.adj[a] and 1 in self.adj[b]
                    if a in (0, 1) and b in (0, 1):
                        continue
                    if not (both_t0 or both_t1):
                        self.adj[a].add(b)
                        self.adj[b].add(a)
        for a in neigh:
            self.adj[a].discard(v)
        del self.adj[v]
        self.maker_turn = not self.maker_turn

    def to_data(self) -> SimpleNamespace:
        verts = sorted(self.adj)
        vmap = {v: i for i, v in enumerate(verts)}
        und = [(vmap[a], vmap[b]) for a in verts for b in sorted(self.adj[a]) if a < b]
        return data_from_undirected(len(verts), und, int(self.maker_turn))


def data_from_undirected(n: int, und_edges: Sequence[Tuple[int, int]], maker_flag: int) -> SimpleNamespace:
    """``Data(x, edge_index)`` exactly as ``convert_node_switching_game(old_style=True)`` lays it out.

    Restates GN0/util/convert_graph.py:77-85 (features: total degree, is_terminal on rows 0,1,
    global maker flag) and :99-117 (``edge_index = cat(E, flip(E))``, int64).
    """
    if len(und_edges) == 0:
        e = torch.zeros((2, 0), dtype=torch.long)
    else:
        e = torch.tensor(list(und_edges), dtype=torch.long).t().contiguous()
    edge_index = torch.cat((e, torch.flip(e, dims=(0,))), dim=1)
    x = torch.zeros((n, 3), dtype=torch.float32)
    deg = torch.bincount(edge_index[1], minlength=n) if edge_index.numel() else torch.zeros(n, dtype=torch.long)
    x[:, 0] = deg.float()
    x[0, 1] = 1
    x[1, 1] = 1
    x[:, 2] = float(maker_flag)
    return SimpleNamespace(x=x, edge_index=edge_index)


def data_from_directed_csr(n: int, src: torch.Tensor, dst: torch.Tensor, maker_flag: int) -> SimpleNamespace:
    """Build the Python-style ``Data`` from the C++ game's symmetric directed edge list
    (cpp_hex/hex_graph_game/shannon_node_switching_game.cpp:718-720): keep s<t once, then flip."""
    keep = src < dst
    und = list(zip(src[keep].tolist(), dst[keep].tolist()))
    return data_from_undirected(n, und, maker_flag)


def collate(datas: Sequence[SimpleNamespace]) -> SimpleNamespace:
    """PyG ``Batch.from_data_list`` for (x, edge_index) == C++ ``collate_batch``.

    Restates cpp_hex/hex_graph_game/util.cpp:22-41: cat x; cat edge_index with per-graph node
    offsets; ``batch`` = graph id per node; ``ptr`` = node prefix sums.  All int64.
    """
    xs, eis, batch = [], [], []
    ptr = [0]
    off = 0
    for i, d in enumerate(datas):
        n = d.x.shape[0]
        xs.append(d.x)
        eis.append(d.edge_index + off)
        batch.append(torch.full((n,), i, dtype=torch.long))
        off += n
        ptr.append(off)
    return SimpleNamespace(
        x=torch.cat(xs, 0),
        edge_index=torch.cat(eis, 1),
        batch=torch.cat(batch, 0),
        ptr=torch.tensor(ptr, dtype=torch.long),
        num_graphs=len(datas),
    )


def csr_by_destination(edge_index: torch.Tensor, num_nodes: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """The CSR the device builder must reproduce bit-exactly: edges stably sorted by destination
    (ties keep input order), ``rowptr[N+1]`` int32 and ``col[E]`` = source node id int32.
    Message direction is edge_index[0] -> edge_index[1] (PyG ``flow='source_to_target'``)."""
    src, dst = edge_index[0], edge_index[1]
    order = torch.sort(dst, stable=True).indices
    col = src[order].to(torch.int32)
    deg = torch.bincount(dst, minlength=num_nodes)
    rowptr = torch.zeros(num_nodes + 1, dtype=torch.int64)
    rowptr[1:] = torch.cumsum(deg, 0)
    return rowptr.to(torch.int32), col


# --------------------------------------------------------------------------------------
# model side
# --------------------------------------------------------------------------------------


def _scatter_sum(src: torch.Tensor, index: torch.Tensor, dim_size: int) -> torch.Tensor:
    out = src.new_zeros((dim_size,) + tuple(src.shape[1:]))
    return out.index_add_(0, index, src)


class _ScatterExtremum(torch.autograd.Function):
    """torch_scatter 2.1.0 `scatter_max` / `scatter_min` (csrc/cpu/scatter_cpu.cpp + reducer.h):
    the CPU kernel scans rows in order and updates on strict `>` / `<`, so the FIRST row attaining
    the extremum is the arg; backward routes the whole gradient to that row
    (torch_scatter/scatter.py `scatter_max` -> `grad_in.scatter_(0, arg, grad_out)`).
    Empty groups output 0 and receive no gradient."""

    @staticmethod
    def forward(ctx, src, index, dim_size, is_max):
        idx = index.view(-1, *([1] * (src.dim() - 1))).expand_as(src)
        out = src.new_zeros((dim_size,) + tuple(src.shape[1:]))
        out = out.scatter_reduce(0, idx, src, reduce="amax" if is_max else "amin", include_self=False)
        hit = src == out.index_select(0, index)
        rows = torch.arange(src.shape[0], device=src.device).view(-1, *([1] * (src.dim() - 1))).expand_as(src)
        cand = torch.where(hit, rows, torch.full_like(rows, src.shape[0]))
        arg = torch.full(out.shape, src.shape[0], dtype=torch.long, device=src.device)
        arg = arg.scatter_reduce(0, idx, cand, reduce="amin", include_self=True)
        ctx.save_for_backward(arg)
        ctx.n = src.shape[0]
        return out

    @staticmethod
    def backward(ctx, grad_out):
        (arg,) = ctx.saved_tensors
        grad_in = grad_out.new_zeros((ctx.n + 1,) + tuple(grad_out.shape[1:]))
        grad_in.scatter_(0, arg, grad_out)  # empty groups have arg == n -> dropped row
        return grad_in[:ctx.n], None, None, None


def scatter(src: torch.Tensor, index: torch.Tensor, dim_size: int, reduce: str) -> torch.Tensor:
    """torch_scatter 2.1.0 ``scatter(src, index, dim=0, dim_size, reduce)`` semantics:
    empty groups give 0 for every reduce; mean divides by max(count, 1)."""
    if reduce == "sum":
        return _scatter_sum(src, index, dim_size)
    if reduce == "mean":
        cnt = torch.bincount(index, minlength=dim_size).clamp(min=1).to(src.dtype)
        return _scatter_sum(src, index, dim_size) / cnt.view(-1, *([1] * (src.dim() - 1)))
    if reduce in ("max", "min"):
        return _ScatterExtremum.apply(src, index, dim_size, reduce == "max")
    raise ValueError(reduce)


def sage_conv(x, edge_index, w_l, b_l, w_r):
    """PyG 2.2.0 ``SAGEConv(aggr='mean', root_weight=True, bias=True)``:
    ``lin_l(mean_{j->i} x_j) + lin_r(x_i)``; lin_l has the bias, lin_r none.
    In-repo restatement of the same math: GN0/torch_script_models.py:52-73."""
    src, dst = edge_index[0], edge_index[1]
    n = x.shape[0]
    agg = _scatter_sum(x.index_select(0, src), dst, n)
    cnt = torch.bincount(dst, minlength=n).clamp(min=1).to(x.dtype)
    mean = agg / cnt.unsqueeze(1)
    return F.linear(mean, w_l, b_l) + F.linear(x, w_r)


def layer_norm_graph(x, weight, bias, eps: float = 1e-5):
    """PyG 2.2.0 ``LayerNorm(mode='graph')`` called WITHOUT a batch vector, as the reference does
    (GN0/models.py:285 ``self.norms[i](x)``, :551 ``self.after_embed_norm(embeds)``): the statistics run over
    ALL nodes and channels of the batch; eps is added to the standard deviation, not the variance."""
    x = x - x.mean()
    out = x / (x.std(unbiased=False) + eps)
    return out * weight + bias


class _NormParams(torch.nn.Module):
    """Affine parameters named like PyG ``LayerNorm``: weight (ones), bias (zeros)."""

    def __init__(self, channels: int):
        super().__init__()
        self.weight = torch.nn.Parameter(torch.ones(channels))
        self.bias = torch.nn.Parameter(torch.zeros(channels))

    def forward(self, x):
        return layer_norm_graph(x, self.weight, self.bias)


class _SageStack(torch.nn.Module):
    """``cachify_gnn(GraphSAGE)`` with dropout=0, act='relu', act_first=False, out_channels=None
    (GN0/models.py:144-160,261-294): per layer conv -> [norm] -> ReLU, after EVERY layer incl. the last
    (with ``norm`` the subclass appends one more norm for the last layer, :158-162)."""

    def __init__(self, in_channels: int, hidden: int, num_layers: int, norm: bool = False):
        super().__init__()
        self.hidden_channels = hidden
        self.num_layers = num_layers
        self.has_cache = False
        self.supports_cache = True
        self.convs = torch.nn.ModuleList()
        for i in range(num_layers):
            c = torch.nn.Module()
            cin = in_channels if i == 0 else hidden
            c.lin_l = torch.nn.Linear(cin, hidden, bias=True)
            c.lin_r = torch.nn.Linear(cin, hidden, bias=False)
            self.convs.append(c)
        self.norms = torch.nn.ModuleList([_NormParams(hidden) for _ in range(num_layers)]) if norm else None

    def forward(self, x, edge_index):
        for i, c in enumerate(self.convs):
            x = sage_conv(x, edge_index, c.lin_l.weight, c.lin_l.bias, c.lin_r.weight)
            if self.norms is not None:
                x = self.norms[i](x)
            x = F.relu(x)
        return x

    @staticmethod
    def _new_conv(cin, cout):
        c = torch.nn.Module()
        c.lin_l = torch.nn.Linear(cin, cout, bias=True)
        c.lin_r = torch.nn.Linear(cin, cout, bias=False)
        return c

    def grow_depth(self, additional_layers):
        """GN0/models.py:166-185: identity layers (lin_l weight and bias 0, lin_r = eye)."""
        for _ in range(additional_layers):
            c = self._new_conv(self.hidden_channels, self.hidden_channels)
            c.lin_l.weight.data[:] = 0
            c.lin_l.bias.data[:] = 0
            c.lin_r.weight.data[:] = torch.eye(self.hidden_channels)
            self.convs.append(c)
        self.num_layers += additional_layers

    def grow_width(self, new_width, new_in_channels=None):
        """GN0/models.py:187-238 (has_output False, no norms).  Every layer is re-created at the new width with default
        init; then, per layer, the old matrix goes into the top-left block and the remainder of the OLD rows is zeroed
        (new input columns do not feed old channels), the new rows keep their fresh init; old lin_l biases are kept."""
        h, old = self.hidden_channels, list(self.convs)
        old_in = old[0].lin_l.weight.shape[1]
        cin0 = old_in if new_in_channels is None else new_in_channels
        widths_in = [cin0] + [new_width] * (self.num_layers - 1)          # :197-206: first, middle and last layer
        fresh = [self._new_conv(ci, new_width) for ci in widths_in]
        with torch.no_grad():
            for i, (c, o) in enumerate(zip(fresh, old)):                   # :207-224
                k = old_in if i == 0 else h
                for name in ("lin_l", "lin_r"):
                    w_new, w_old = getattr(c, name).weight, getattr(o, name).weight
                    w_new[:h, :k] = w_old
                    w_new[:h, k:] = 0
                c.lin_l.bias[:h] = o.lin_l.bias
        self.convs = torch.nn.ModuleList(fresh)
        self.hidden_channels = new_width


class _MLP(torch.nn.Module):
    """GN0/models.py:36-49,75-82 with num_hidden_layers=1, no output activation."""

    def __init__(self, hidden, num_input, num_output):
        super().__init__()
        self.layers = torch.nn.ModuleList([torch.nn.Linear(num_input, hidden), torch.nn.Linear(hidden, num_output)])

        self.hidden_channels, self.num_input = hidden, num_input

    def forward(self, x):
        return self.layers[1](F.relu(self.layers[0](x)))

    def grow_input_width(self, new_input_width, new_hidden_channels):
        """GN0/models.py:52-73 for two layers (the middle loop runs zero times): both layers become zero matrices of the new
        shape carrying the old weights in their top-left block; the first layer's bias is zero-extended, the output bias kept.
        The old input columns keep their positions 0 .. 4*old_H-1 although the pooled vector is laid out for the new width."""
        first = _zero_extended_linear(self.layers[0], new_hidden_channels, new_input_width, zero_bias=True)
        last = _zero_extended_linear(self.layers[1], 1, new_hidden_channels, zero_bias=False)
        self.layers[0], self.layers[1] = first, last
        self.num_input, self.hidden_channels = new_input_width, new_hidden_channels


def _zero_extended_linear(old: torch.nn.Linear, n_out: int, n_in: int, zero_bias: bool) -> torch.nn.Linear:
    """A fresh Linear(n_in, n_out) (default init is drawn, then discarded, as in the reference) whose weight is zero except
    for `old.weight` in the top-left block; bias = old bias, zero-extended when `zero_bias` else copied over the whole."""
    lin = torch.nn.Linear(n_in, n_out)
    o, i = old.weight.shape
    with torch.no_grad():
        lin.weight.zero_()
        lin.weight[:o, :i] = old.weight
        if zero_bias:
            lin.bias.zero_()
            lin.bias[:o] = old.bias
        else:
            lin.bias.copy_(old.bias)
    return lin


class NoisyLinearOracle(torch.nn.Module):
    """``FactorizedNoisyLinear`` (GN0/models.py:84-141): w = mu_w + sigma_w * eps_w, b = mu_b + sigma_b * eps_b with
    factorised noise eps_w = f(eps_out) f(eps_in)^T, eps_b = f(eps_out), f(x) = sgn(x) sqrt(|x|)."""

    def __init__(self, in_features: int, out_features: int, sigma_0: float):
        super().__init__()
        self.in_features, self.out_features, self.sigma_0 = in_features, out_features, sigma_0
        self.weight_mu = torch.nn.Parameter(torch.empty(out_features, in_features))
        self.weight_sigma = torch.nn.Parameter(torch.empty(out_features, in_features))
        self.register_buffer('weight_epsilon', torch.empty(out_features, in_features))
        self.bias_mu = torch.nn.Parameter(torch.empty(out_features))
        self.bias_sigma = torch.nn.Parameter(torch.empty(out_features))
        self.register_buffer('bias_epsilon', torch.empty(out_features))
        with torch.no_grad():                                                          # :105-114
            bound = 1 / math.sqrt(in_features)
            for mu, sigma in ((self.weight_mu, self.weight_sigma), (self.bias_mu, self.bias_sigma)):
                mu.uniform_(-bound, bound)
                sigma.fill_(sigma_0 * bound)
        self.reset_noise()

    @torch.no_grad()
    def reset_noise(self):                                                             # :116-128, z_in drawn first
        z_in, z_out = torch.randn(self.in_features), torch.randn(self.out_features)
        f_in, f_out = torch.copysign(z_in.abs().sqrt(), z_in), torch.copysign(z_out.abs().sqrt(), z_out)
        self.weight_epsilon.copy_(f_out.unsqueeze(1) * f_in.unsqueeze(0))
        self.bias_epsilon.copy_(f_out)

    @torch.no_grad()
    def disable_noise(self):                                                           # :130-133
        self.weight_epsilon.zero_()
        self.bias_epsilon.zero_()

    def forward(self, x):                                                              # :135-141
        return F.linear(x, self.weight_mu + self.weight_sigma * self.weight_epsilon,
                        self.bias_mu + self.bias_sigma * self.bias_epsilon)


class _Head(torch.nn.Module):
    """``HeadNetwork`` (GN0/models.py:318-334,368-384) with value_head_type='mlp',
    value_aggr_types=('sum','max','min','mean'); ``noisy_dqn`` swaps the advantage Linear (:331-334)."""

    def __init__(self, hidden: int, num_layers: int, noisy_dqn: bool = False, noise_sigma: float = 0.0,
                 norm: bool = False):
        super().__init__()
        self.gnn = _SageStack(hidden, hidden, num_layers, norm=norm)
        self.linear = NoisyLinearOracle(hidden, 1, noise_sigma) if noisy_dqn else torch.nn.Linear(hidden, 1)
        self.value_head = _MLP(hidden // 2, 4 * hidden, 1)
        self.hidden_channels = hidden

    def grow_width(self, new_width, new_in_channels=None):
        """GN0/models.py:336-358: wider head stack, advantage Linear zero-extended over the new channels, value MLP widened."""
        assert isinstance(self.linear, torch.nn.Linear)
        self.gnn.grow_width(new_width, new_in_channels=new_in_channels)
        self.linear = _zero_extended_linear(self.linear, 1, new_width, zero_bias=False)
        self.value_head.grow_input_width(new_width * 4, new_width // 2)
        self.hidden_channels = new_width

    def grow_depth(self, additional_layers):
        self.gnn.grow_depth(additional_layers)

    def forward(self, x, edge_index, graph_indices, num_graphs, advantages_only=False):
        x = self.gnn(x, edge_index)
        adv = self.linear(x)
        if advantages_only:
            return adv
        parts = [scatter(x, graph_indices, num_graphs, a) for a in ("sum", "max", "min", "mean")]
        return adv, self.value_head(torch.cat(parts, dim=1))


class ModernTwoHeadedOracle(torch.nn.Module):
    """``get_pre_defined('modern_two_headed', args)`` (GN0/models.py:928-947 -> DuellingTwoHeaded :477-590),
    incl. ``args.norm`` (PyG LayerNorm, whole-batch statistics) and ``args.noisy_dqn``.  Parameter names follow PyG 2.2.0 /
    the reference attribute accesses so state_dicts interchange with the product module."""

    def __init__(self, hidden_channels: int, num_layers: int, num_head_layers: int = 2, noisy_dqn: bool = False,
                 noisy_sigma0: float = 0.5, norm: bool = False):
        super().__init__()
        self.gnn = _SageStack(2, hidden_channels, num_layers, norm=norm)
        # models.py:480-483: with ``norm`` one more LayerNorm between body and heads (:550-551), no ReLU after it
        self.after_embed_norm = _NormParams(hidden_channels) if norm else None
        self.maker_head = _Head(hidden_channels, num_head_layers, noisy_dqn, noisy_sigma0, norm=norm)
        self.breaker_head = _Head(hidden_channels, num_head_layers, noisy_dqn, noisy_sigma0, norm=norm)
        self.final_conv_acts = None

    def grow_depth(self, additional_layers):                                   # models.py:494-495
        self.gnn.grow_depth(additional_layers)

    def grow_width(self, new_width):                                           # models.py:497-508 (no norm)
        self.gnn.grow_width(new_width)
        self.maker_head.grow_width(new_width, new_in_channels=new_width)
        self.breaker_head.grow_width(new_width, new_in_channels=new_width)

    def forward(self, x, edge_index, graph_indices=None, ptr=None, set_cache=False,
                advantages_only=False, seperate=False):
        assert torch.all(x[:, 2] == x[0, 2])                                   # models.py:538
        is_maker = x[0, 2]
        x = x[:, :2]
        if graph_indices is None:                                              # models.py:542-543
            graph_indices = torch.zeros(x.size(0), dtype=torch.long, device=x.device)
        num_graphs = int(graph_indices.max()) + 1 if graph_indices.numel() else 0
        embeds = self.gnn(x, edge_index)
        if self.after_embed_norm is not None:                                  # models.py:550-551
            embeds = self.after_embed_norm(embeds)
        self.final_conv_acts = embeds
        head = self.maker_head if is_maker == 1 else self.breaker_head        # models.py:556-565
        res = head(embeds, edge_index, graph_indices, num_graphs, advantages_only=advantages_only)
        if advantages_only:
            return 2 * torch.tanh(res)                                         # models.py:567-569
        adv = 2 * torch.tanh(res[0])
        value = torch.tanh(res[1])
        adv_means = scatter(adv, graph_indices, num_graphs, "mean")            # models.py:578
        centred = adv - adv_means.index_select(0, graph_indices)
        if seperate:
            return value.squeeze(), centred.squeeze()
        return (value.index_select(0, graph_indices) + centred).squeeze()     # models.py:584

    def simple_forward(self, data):
        if hasattr(data, "batch") and data.batch is not None:
            return self.forward(data.x, data.edge_index, data.batch, getattr(data, "ptr", None))
        return self.forward(data.x, data.edge_index)


def rainbow_loss(q: torch.Tensor, ptr: torch.Tensor, action: torch.Tensor, target: torch.Tensor,
                 weight: torch.Tensor) -> torch.Tensor:
    """The scalar the learner differentiates: Rainbow/common/rainbow.py:301-312 with MSE --
    ``td_est = Q[action + ptr[:-1]]``; ``loss = mean(w * (td_tgt - td_est)^2)``."""
    td_est = q[action + ptr[:-1]]
    return torch.mean(weight * (target - td_est) ** 2)


def flops_forward(n_nodes: int, num_graphs: int, hidden: int, num_layers: int, num_head_layers: int = 2,
                  advantages_only: bool = False) -> int:
    """Algorithmic FLOPs: node-linear contractions of the un-padded model."""
    H = hidden
    per_node = 4 * 2 * H + (num_layers - 1 + num_head_layers) * 4 * H * H + 2 * H
    f = n_nodes * per_node
    if not advantages_only:
        f += num_graphs * 2 * (4 * H * (H // 2) + (H // 2))
    return f


# ---- HexAra network: GN0/torch_script_models.py:286-379 -----------------------------------------------
class _AraStack(torch.nn.Module):
    """``ModifiedBaseNet(in, hidden, num_layers, out_channels=None|1, conv_class=ModifiedSAGEConv, norm=None)``
    (GN0/torch_script_models.py:96-189): convs in -> hidden, (num_layers - 2) x hidden -> hidden, hidden -> out (a one-layer
    stack is the single conv in -> out); ReLU after every conv EXCEPT the last (:178-183)."""

    def __init__(self, in_channels: int, hidden: int, num_layers: int, out_channels: Optional[int] = None):
        super().__init__()
        self.num_layers = num_layers
        self.convs = torch.nn.ModuleList()
        cin = in_channels
        for i in range(num_layers):
            cout = (out_channels if out_channels is not None else hidden) if i == num_layers - 1 else hidden
            self.convs.append(_SageStack._new_conv(cin, cout))
            cin = hidden

    def forward(self, x, edge_index):
        for i, c in enumerate(self.convs):
            x = sage_conv(x, edge_index, c.lin_l.weight, c.lin_l.bias, c.lin_r.weight)
            if i != self.num_layers - 1:
                x = F.relu(x)
        return x


class _AraMLP(torch.nn.Module):
    """``MLP(hidden//2, 1, 4*hidden, 1)`` of torch_script_models.py:263-284: Linear -> ReLU -> Linear, no output activation."""

    def __init__(self, hidden: int, num_input: int, num_output: int):
        super().__init__()
        self.layers = torch.nn.ModuleList([torch.nn.Linear(num_input, hidden), torch.nn.Linear(hidden, num_output)])

    def forward(self, x):
        return self.layers[1](F.relu(self.layers[0](x)))


def scatter_log_softmax(src: torch.Tensor, index: torch.Tensor, dim_size: int, eps: float = 1e-12) -> torch.Tensor:
    """torch_scatter 2.1.0 ``composite.scatter_log_softmax``: x - max_g(x) - log(sum_g exp(x - max_g) + eps)."""
    mx = scatter(src, index, dim_size, "max")
    rec = src - mx.index_select(0, index)
    norm = _scatter_sum(rec.exp(), index, dim_size)
    return rec - (norm + eps).log().index_select(0, index)


class HexAraOracle(torch.nn.Module):
    """``SAGE_torch_script(hidden_channels, hidden_layers, policy_layers, value_layers, in_channels=3, swap_allowed)``
    (GN0/torch_script_models.py:286-379, factory ``get_current_model`` :495-500; called from C++ at
    cpp_hex/CrazyAra/nn_api.cpp:14-19).  x follows the C++ convention (shannon_node_switching_game.cpp:690-726): column 0 =
    IS_TERMINAL, columns 1 / 2 = swap flags; all three are network inputs.  Returns (log-policy over the non-terminal nodes
    of every graph [+ one swap entry per graph whose last node has x[:,2] != 0], value [B], output graph ids, output ptr)."""

    def __init__(self, hidden_channels: int, hidden_layers: int, policy_layers: int, value_layers: int, in_channels: int = 3,
                 swap_allowed: bool = False):
        super().__init__()
        self.swap_allowed = swap_allowed
        self.gnn = _AraStack(in_channels, hidden_channels, hidden_layers)
        self.my_modules = torch.nn.ModuleDict()
        self.my_modules["value_head"] = _AraStack(hidden_channels, hidden_channels, value_layers)
        self.my_modules["policy_head"] = _AraStack(hidden_channels, hidden_channels, policy_layers, out_channels=1)
        self.my_modules["value_linear"] = _AraMLP(hidden_channels // 2, hidden_channels * 4, 1)
        self.my_modules["swap_linear"] = _AraMLP(hidden_channels // 2, hidden_channels * 4, 1)

    def forward(self, x, edge_index, graph_indices, batch_ptr):
        B = batch_ptr.numel() - 1
        n = batch_ptr[1:] - batch_ptr[:-1]
        assert bool((n > 2).all())                                                          # :311
        embeds = self.gnn(x, edge_index)                                                    # :312 (norm=None)
        pi = self.my_modules["policy_head"](embeds, edge_index).reshape(-1)                 # :318, :330
        ve = self.my_modules["value_head"](embeds, edge_index)                              # :319
        parts = torch.cat([scatter(ve, graph_indices, B, r) for r in ("sum", "max", "min", "mean")], dim=1)   # :320-324
        value = torch.tanh(self.my_modules["value_linear"](parts)).reshape(-1)              # :326-327
        # terminal-row removal (+ swap rule), :337-373, restated per graph: graph g keeps its non-terminal nodes in order and,
        # when swapping is allowed in it (x[last node of g, 2] != 0), one more entry -- the swap logit -- right after them
        # (the reference writes it into the slot of the NEXT graph's first terminal node and re-labels it, :353-357, or
        # appends it for the last graph, :364-366: the same position)
        swap = torch.zeros(B, dtype=torch.bool)
        if self.swap_allowed:
            swap = x[batch_ptr[1:] - 1, 2] != 0                                             # :340, :349
            logit = self.my_modules["swap_linear"](parts).reshape(-1)                       # :338-339
        pieces, gids = [], []
        for g in range(B):
            pieces.append(pi[int(batch_ptr[g]) + 2:int(batch_ptr[g + 1])])
            cnt = int(n[g]) - 2
            if bool(swap[g]):
                pieces.append(logit[g:g + 1])
                cnt += 1
            gids.append(torch.full((cnt,), g, dtype=torch.long))
        out_gi = torch.cat(gids)
        out_ptr = batch_ptr - 2 * torch.arange(B + 1) + torch.cat([torch.zeros(1, dtype=torch.long), swap.long().cumsum(0)])  # :351, :371
        return scatter_log_softmax(torch.cat(pieces), out_gi, B), value, out_gi, out_ptr    # :378-379
