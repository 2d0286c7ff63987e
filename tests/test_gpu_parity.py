"""GPU (B200): parity of the CUDA path, called through the C-ABI, against the CPU oracle.

Bars (BASELINE.json north_star): CSR / batching bit-exact; Q-values and gradients within 1e-5
relative (fp32).  Nothing here reads /root/reference."""
import pytest
import torch

from gnn_hex_b200 import ModernTwoHeaded, data, ops, _lib
import positions
from oracle import hexgnn_oracle as orc
from helpers import REL_TOL, grad_errs, loss_inputs, make_pair, oracle64, rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _csr_check(b, **kw):
    d = b.to(DEV)
    csr = ops.build_csr(d.edge_index, **{k: (getattr(d, v) if isinstance(v, str) else v) for k, v in kw.items()})
    rowptr, col = orc.csr_by_destination(b.edge_index, b.x.shape[0])
    assert int(csr.status.item()) == 0
    assert torch.equal(csr.rowptr.cpu(), rowptr)
    assert torch.equal(csr.col.cpu(), col)
    assert torch.equal(csr.gptr.cpu().long(), b.ptr)
    return csr


@pytest.mark.parametrize("fixture", ["pos_11x11_seed1234.npz", "pos_7x7_seed1234.npz"])
def test_csr_bit_exact(fixture):
    gs = positions.load_positions(fixture, maker=1)
    b = data.collate(gs)
    _csr_check(b, batch="batch", ptr="ptr", num_nodes=b.x.shape[0])
    _csr_check(b, batch="batch", num_nodes=b.x.shape[0], num_graphs=b.num_graphs)
    _csr_check(b, ptr="ptr", num_nodes=b.x.shape[0])
    # shuffled edge order: still == stable sort by destination of THAT order
    perm = torch.randperm(b.edge_index.shape[1], generator=torch.Generator().manual_seed(5))
    b2 = data.HexBatch(x=b.x, edge_index=b.edge_index[:, perm].contiguous(), batch=b.batch, ptr=b.ptr,
                       num_graphs=b.num_graphs)
    csr = _csr_check(b2, batch="batch", ptr="ptr", num_nodes=b.x.shape[0])
    # tiles: whole graphs, <= 128 rows, cover everything
    t = csr.tile_g0.cpu()[: int(csr.ntiles.item()) + 1]
    assert t[0] == 0 and t[-1] == b.num_graphs and torch.all(t[1:] > t[:-1])
    rows = b.ptr[t[1:]] - b.ptr[t[:-1]]
    assert rows.max() <= 128


def test_csr_single_graph_and_empty_graphs():
    g = positions.empty_board(11)
    d = g.to(DEV)
    csr = ops.build_csr(d.edge_index, num_nodes=123)
    rowptr, col = orc.csr_by_destination(g.edge_index, 123)
    assert torch.equal(csr.rowptr.cpu(), rowptr) and torch.equal(csr.col.cpu(), col)
    assert csr.gptr.cpu().tolist() == [0, 123] and int(csr.ntiles.item()) == 1
    # batch vector that skips graph ids (empty graphs 1 and 3)
    batch = torch.cat([torch.zeros(123, dtype=torch.long), torch.full((123,), 2, dtype=torch.long)])
    ei = torch.cat([g.edge_index, g.edge_index + 123], 1)
    csr = ops.build_csr(ei.to(DEV), batch.to(DEV), num_nodes=246, num_graphs=4)
    assert csr.gptr.cpu().tolist() == [0, 123, 123, 246, 246]
    assert int(csr.status.item()) == 0


def test_csr_flags_bad_input():
    g = positions.empty_board(5)
    ei = g.edge_index.clone()
    ei[0, 0] = 10_000
    csr = ops.build_csr(ei.to(DEV), num_nodes=27)
    assert int(csr.status.item()) & _lib.ST_EDGE_CROSSES
    batch = torch.tensor([0] * 20 + [1] * 3 + [0] * 4)
    csr = ops.build_csr(g.edge_index.to(DEV), batch.to(DEV), num_nodes=27, num_graphs=2)
    assert int(csr.status.item()) & _lib.ST_BATCH_UNSORTED
    with pytest.raises(_lib.HexGnnError):
        csr.raise_on_error()


CONFIGS = {
    "S": dict(hidden=35, layers=10, fixture="pos_7x7_seed1234.npz", board=7),
    "L": dict(hidden=110, layers=15, fixture="pos_11x11_seed1234.npz", board=11),
}


def _batch(cfg, kind, maker, count):
    if kind == "empty":
        gs = [positions.empty_board(cfg["board"], maker)] * count
    else:
        gs = positions.load_positions(cfg["fixture"], maker=maker)[:count]
    return data.collate(gs)


@pytest.mark.parametrize("name,kind,maker,count", [
    ("S", "play", 1, 256), ("S", "play", 0, 64), ("S", "empty", 1, 32),
    ("L", "play", 1, 64), ("L", "play", 0, 32), ("L", "empty", 0, 16),
])
def test_forward_parity(name, kind, maker, count):
    cfg = CONFIGS[name]
    o, m = make_pair(cfg["hidden"], cfg["layers"])
    m = m.to(DEV)
    b = _batch(cfg, kind, maker, count)
    d = b.to(DEV)
    with torch.no_grad():
        q_ref = o(b.x, b.edge_index, b.batch, b.ptr)
        a_ref = o(b.x, b.edge_index, b.batch, advantages_only=True)
        v_ref, c_ref = o(b.x, b.edge_index, b.batch, seperate=True)
        q = m(d.x, d.edge_index, d.batch, d.ptr)
        a = m(d.x, d.edge_index, d.batch, d.ptr, advantages_only=True)
        v, c = m(d.x, d.edge_index, d.batch, d.ptr, seperate=True)
    assert q.shape == q_ref.shape and a.shape == a_ref.shape and v.shape == v_ref.shape
    assert rel_err(q, q_ref) <= REL_TOL
    assert rel_err(a, a_ref) <= REL_TOL
    assert rel_err(v, v_ref) <= REL_TOL and rel_err(c, c_ref) <= REL_TOL * max(1.0, float(a_ref.abs().max() / c_ref.abs().max()))
    # simple_forward == forward (models.py:586-590) and single-graph call (graph_indices=None)
    with torch.no_grad():
        assert torch.equal(m.simple_forward(d), q)
        g0 = data.HexGraph(x=b.x[: b.ptr[1]], edge_index=b.edge_index[:, b.edge_index[1] < b.ptr[1]])
        q0 = m(g0.x.to(DEV), g0.edge_index.to(DEV))
    assert rel_err(q0, q_ref[: b.ptr[1]]) <= REL_TOL


@pytest.mark.parametrize("name,kind,maker,count", [
    ("S", "play", 1, 256), ("S", "empty", 0, 16), ("L", "play", 1, 48), ("L", "empty", 1, 8),
])
def test_backward_parity(name, kind, maker, count):
    cfg = CONFIGS[name]
    o, m = make_pair(cfg["hidden"], cfg["layers"])
    m = m.to(DEV)
    b = _batch(cfg, kind, maker, count)
    d = b.to(DEV)
    action, target, weight = loss_inputs(b)
    loss_ref = orc.rainbow_loss(o(b.x, b.edge_index, b.batch, b.ptr), b.ptr, action, target, weight)
    loss_ref.backward()
    o64 = oracle64(o)
    orc.rainbow_loss(o64(b.x.double(), b.edge_index, b.batch, b.ptr), b.ptr, action, target.double(),
                     weight.double()).backward()
    q = m(d.x, d.edge_index, d.batch, d.ptr)
    loss = orc.rainbow_loss(q, d.ptr, action.to(DEV), target.to(DEV), weight.to(DEV))
    loss.backward()
    assert rel_err(loss, loss_ref) <= REL_TOL
    glob, worst, wname = grad_errs(m, o, o64)
    assert glob <= REL_TOL, glob                # 1e-5 relative on the gradient (BASELINE.json)
    assert worst <= 1.0, (worst, wname)         # per tensor: within fp32 round-off of the fp64 truth
    # the head not taken gets exactly zero gradient, like autograd leaves it untouched
    other = m.breaker_head if maker == 1 else m.maker_head
    assert all(float(p.grad.abs().max()) == 0.0 for p in other.parameters())
    # embedding tap (models.py:553)
    assert rel_err(m.final_conv_acts, o.final_conv_acts) <= REL_TOL


def test_backward_other_modes_and_embed_grads():
    o, m = make_pair(35, 10)
    m = m.to(DEV)
    m.capture_embed_grads = True
    b = _batch(CONFIGS["S"], "play", 1, 40)
    d = b.to(DEV)
    gen = torch.Generator().manual_seed(3)
    wn = torch.randn(b.x.shape[0], generator=gen)
    wv = torch.randn(b.num_graphs, generator=gen)
    # seperate=True
    v_ref, c_ref = o(b.x, b.edge_index, b.batch, seperate=True)
    o.final_conv_acts.retain_grad()
    ((c_ref * wn).sum() + (v_ref * wv).sum()).backward()
    o64 = oracle64(o)
    v64, c64 = o64(b.x.double(), b.edge_index, b.batch, seperate=True)
    ((c64 * wn.double()).sum() + (v64 * wv.double()).sum()).backward()
    v, c = m(d.x, d.edge_index, d.batch, d.ptr, seperate=True)
    ((c * wn.to(DEV)).sum() + (v * wv.to(DEV)).sum()).backward()
    glob, worst, wname = grad_errs(m, o, o64)
    assert glob <= REL_TOL and worst <= 1.0, (glob, worst, wname)
    # Embedding-gradient tap (models.py:534-535).  Twin / mirror-symmetric nodes tie in the max/min
    # pooling up to the last ulp, so WHICH of two equivalent rows receives the pooled gradient depends
    # on rounding (torch_scatter's CPU and CUDA kernels differ the same way): with pooling in the
    # graph, compare the permutation-invariant per-graph column sums; row by row below (no pooling).
    gk, gr = m.final_conv_grads.cpu(), o.final_conv_acts.grad
    sk = torch.zeros(b.num_graphs, gk.shape[1]).index_add_(0, b.batch, gk)
    sr = torch.zeros(b.num_graphs, gr.shape[1]).index_add_(0, b.batch, gr)
    assert rel_err(sk, sr) <= 10 * REL_TOL
    # advantages_only=True
    o.zero_grad(); m.zero_grad(); o64.zero_grad()
    (o64(b.x.double(), b.edge_index, b.batch, advantages_only=True).squeeze(1) * wn.double()).sum().backward()
    ra = o(b.x, b.edge_index, b.batch, advantages_only=True)
    o.final_conv_acts.retain_grad()
    (ra.squeeze(1) * wn).sum().backward()
    (m(d.x, d.edge_index, d.batch, d.ptr, advantages_only=True).squeeze(1) * wn.to(DEV)).sum().backward()
    glob, worst, wname = grad_errs(m, o, o64)
    assert glob <= REL_TOL and worst <= 1.0, (glob, worst, wname)
    assert rel_err(m.final_conv_grads, o.final_conv_acts.grad) <= REL_TOL


def test_deterministic_bitwise():
    _, m = make_pair(110, 15)
    m = m.to(DEV)
    b = _batch(CONFIGS["L"], "play", 1, 96).to(DEV)
    action, target, weight = [t.to(DEV) for t in loss_inputs(data.collate(positions.load_positions(CONFIGS["L"]["fixture"], maker=1)[:96]))]
    outs = []
    for _ in range(3):
        m.zero_grad()
        q = m(b.x, b.edge_index, b.batch, b.ptr)
        orc.rainbow_loss(q, b.ptr, action, target, weight).backward()
        outs.append((q.detach().clone(), torch.cat([p.grad.reshape(-1) for p in m.parameters()]).clone()))
    for q, g in outs[1:]:
        assert torch.equal(q, outs[0][0]) and torch.equal(g, outs[0][1])


def test_ragged_mix_forward_and_oversize_flag():
    o, m = make_pair(110, 15)
    m = m.to(DEV)
    gs = positions.load_positions("pos_mix_5_13_seed1234.npz", maker=1)
    small = [g for g in gs if g.x.shape[0] <= 128 and g.edge_index.shape[1] <= 1024][:96]
    b = data.collate(small)
    d = b.to(DEV)
    with torch.no_grad():
        q = m(d.x, d.edge_index, d.batch, d.ptr)
        q_ref = o(b.x, b.edge_index, b.batch, b.ptr)
    assert rel_err(q, q_ref) <= REL_TOL
    big = [g for g in gs if g.x.shape[0] > 128][:2] + small[:2]
    d = data.collate(big).to(DEV)
    m.strict = True
    with torch.no_grad(), pytest.raises(_lib.HexGnnError):
        m(d.x, d.edge_index, d.batch, d.ptr)


def test_cpu_tensors_are_rejected_loudly():
    _, m = make_pair(35, 10)
    b = _batch(CONFIGS["S"], "play", 1, 2)
    with pytest.raises(_lib.HexGnnError):
        m(b.x, b.edge_index, b.batch, b.ptr)


@pytest.mark.parametrize("kind", ["mse", "smooth_l1"])
@pytest.mark.parametrize("weighted", [True, False])
def test_td_loss_matches_torch(kind, weighted):
    """hexgnn_td_loss (rainbow.py:301-312 on device) vs the same expression in torch: loss, td errors, dloss/dq."""
    from gnn_hex_b200 import ops
    b = data.collate(positions.load_positions("pos_7x7_seed1234.npz", maker=1)[:97])
    action, target, weight = loss_inputs(b)
    g = torch.Generator().manual_seed(5)
    q = (torch.randn(b.x.shape[0], generator=g) * 1.5).requires_grad_(True)   # |err| on both sides of 1
    w = weight if weighted else None
    td_est = q[action + b.ptr[:-1]]
    fn = torch.nn.MSELoss(reduction="none") if kind == "mse" else torch.nn.SmoothL1Loss(reduction="none")
    losses = fn(target, td_est)
    ref = torch.mean(w * losses) if weighted else torch.mean(losses)
    (ref * 0.5).backward()
    qd = q.detach().to(DEV).requires_grad_(True)
    loss, err = ops.td_loss(qd, b.ptr.to(DEV), action.to(DEV), target.to(DEV), None if w is None else w.to(DEV), kind, 0.5)
    loss.backward()
    assert abs(float(loss) - 0.5 * float(ref)) <= 1e-6 * max(1.0, abs(float(ref)))
    assert torch.allclose(err.cpu(), (td_est - target).detach(), atol=1e-6)
    assert torch.allclose(qd.grad.cpu(), q.grad, atol=1e-7, rtol=1e-5)
    assert int((qd.grad != 0).sum()) <= b.num_graphs


def test_mask_argmax_and_td_target_match_torch():
    """hexgnn_mask_argmax / hexgnn_td_target vs rainbow.py:222-225,263-269 written with torch ops."""
    b = data.collate(positions.load_positions("pos_mix_5_13_seed1234.npz", maker=1)[:120])
    g = torch.Generator().manual_seed(9)
    q = torch.randn(b.x.shape[0], generator=g)
    ref = q.clone()
    ref[b.ptr[:-1]] = -5
    ref[b.ptr[:-1] + 1] = -5
    mx_ref = torch.stack([ref[b.ptr[i]:b.ptr[i + 1]].max() for i in range(b.num_graphs)])
    arg_ref = torch.stack([ref[b.ptr[i]:b.ptr[i + 1]].argmax() + b.ptr[i] for i in range(b.num_graphs)])
    qd = q.to(DEV)
    mx, arg = ops.mask_argmax(qd, b.ptr.to(DEV))
    assert torch.equal(qd.cpu(), ref) and torch.equal(mx.cpu(), mx_ref) and torch.equal(arg.cpu(), arg_ref)
    qt = torch.randn(b.x.shape[0], generator=g)
    reward = torch.randn(b.num_graphs, generator=g)
    done = (torch.rand(b.num_graphs, generator=g) < 0.3).float()
    tgt = ops.td_target(qt.to(DEV), arg, reward.to(DEV), done.to(DEV), 0.97)
    assert torch.allclose(tgt.cpu(), reward + 0.97 * qt[arg_ref] * (1 - done), atol=1e-6)
    # ties resolve to the lowest row; all-masked two-node graphs return the mask value
    z = torch.zeros(7, device=DEV)
    mx2, arg2 = ops.mask_argmax(z, torch.tensor([0, 2, 7]), mask_value=-5.0)
    assert mx2.tolist() == [-5.0, 0.0] and arg2.tolist() == [0, 4]


@pytest.mark.parametrize("mode", ["q", "adv", "sep"])
def test_oversize_graphs_board_size_transfer(mode):
    """BASELINE configs[3]: a mixed 5x5..13x13 batch; the 12x12 / 13x13 positions (up to 171 nodes) do not fit one
    128-row tile and -- with the fused path forced (path="tc"; the default route for such batches is the layer-at-a-time
    tensor-core path, tests/test_gpu_lw.py) -- go through hexgnn_forward_big.  Every graph must match the oracle; the status
    word ends clean."""
    o, m = make_pair(110, 15)
    m = m.to(DEV)
    m.path = "tc"
    m.max_graph_nodes = 256
    m.strict = True
    gs = positions.load_positions("pos_mix_5_13_seed1234.npz", maker=1)
    big = [g for g in gs if g.x.shape[0] > 128]
    assert len(big) >= 4
    mix = big[:6] + [g for g in gs if g.x.shape[0] <= 128][:26] + big[6:10]
    b = data.collate(mix)
    d = b.to(DEV)
    kw = {"advantages_only": mode == "adv", "seperate": mode == "sep"}
    with torch.no_grad():
        out = m(d.x, d.edge_index, d.batch, d.ptr, **kw)
        ref = o(b.x, b.edge_index, b.batch, b.ptr, **kw)
    if mode == "sep":
        # adv - mean(adv) is a difference of nearly equal numbers for random-init weights: scale by the value range
        scale = float(max(ref[0].abs().max(), ref[1].abs().max()))
        assert rel_err(out[0], ref[0]) <= REL_TOL
        assert float((out[1].cpu() - ref[1]).abs().max()) <= REL_TOL * scale
    else:
        assert out.shape == ref.shape and rel_err(out, ref) <= REL_TOL
    # a graph beyond max_graph_nodes is still refused loudly
    m.max_graph_nodes = 140
    with torch.no_grad(), pytest.raises(_lib.HexGnnError):
        m(d.x, d.edge_index, d.batch, d.ptr)


@pytest.mark.gpu
@pytest.mark.parametrize("size,count,kind,weighted", [
    ("L", 256, "mse", True),          # 209 tiles on 148 CTAs: forward / chain / dW overlap through the tile flags
    ("L", 40, "smooth_l1", False),    # fewer tiles than CTAs: no phase-A items, nothing to overlap
    ("S", 256, "mse", True),
])
def test_train_step_equals_autograd_route(size, count, kind, weighted):
    """hexgnn_train_step (forward + TD loss + backward chained by programmatic dependent launch, the loss gradient derived
    inside the chain kernel) against forward -> td_loss -> backward as separate calls: the same bits (rainbow.py:301-316)."""
    cfg = CONFIGS[size]
    _, m = make_pair(cfg["hidden"], cfg["layers"])
    m = m.to(DEV)
    gs = positions.load_positions(cfg["fixture"], maker=1)[:count]
    hb = data.collate(gs)
    b = hb.to(DEV)
    action, target, weight = [t.to(DEV) for t in loss_inputs(hb)]
    if not weighted:
        weight = None
    m.zero_grad(set_to_none=True)
    q_ref = m(b.x, b.edge_index, b.batch, b.ptr)
    loss_ref, err_ref = ops.td_loss(q_ref, b.ptr, action, target, weight, kind=kind, scale=0.5)
    loss_ref.backward()
    g_ref = m.flat_gradient.clone()
    for _ in range(3):   # repeated: the flags / tickets are reset per call, results are bit-reproducible
        m.zero_grad(set_to_none=True)
        loss, err, q = m.train_step(b.x, b.edge_index, b.batch, b.ptr, action=action, target=target, weight=weight,
                                    kind=kind, scale=0.5)
        assert int(m.last_csr.status.item()) == 0
        assert torch.equal(q, q_ref.detach())
        assert torch.equal(loss, loss_ref.detach()) and torch.equal(err, err_ref)
        assert torch.equal(m.flat_gradient, g_ref)
    # gradients still in place: the next step accumulates, like autograd's +=
    loss, err, q = m.train_step(b.x, b.edge_index, b.batch, b.ptr, action=action, target=target, weight=weight,
                                kind=kind, scale=0.5)
    assert torch.allclose(m.flat_gradient, 2 * g_ref, rtol=1e-6, atol=0)


@pytest.mark.gpu
@pytest.mark.parametrize("skip_untaken", [False, True])
def test_fused_adam_matches_torch_adam_with_clipping(skip_untaken):
    """hexgnn_adam_step against nn.utils.clip_grad_norm_ + torch.optim.Adam on the same gradients (rainbow.py:79,320-326),
    over steps that alternate the trained head: the untrained head keeps receiving zero-gradient Adam updates once it
    has been trained (torch 1.13 zero_grad, the reference environment) or is left alone (set_to_none semantics)."""
    from gnn_hex_b200 import FusedAdam
    cfg = CONFIGS["S"]
    _, m = make_pair(cfg["hidden"], cfg["layers"])
    m = m.to(DEV)
    # torch twin: the same parameter values as separate leaf tensors
    ref_params = [p.detach().clone().requires_grad_(True) for p in m._ordered_params()]
    ref_opt = torch.optim.Adam(ref_params, lr=3e-3, eps=1.5e-4)
    opt = FusedAdam(m, lr=3e-3, eps=1.5e-4, max_grad_norm=0.05, skip_untaken=skip_untaken)
    for step, maker in enumerate([1, 1, 0, 1, 0, 0]):
        hb = data.collate(positions.load_positions(cfg["fixture"], maker=maker)[step * 8:step * 8 + 24])
        b = hb.to(DEV)
        action, target, weight = [t.to(DEV) for t in loss_inputs(hb)]
        m.zero_grad(set_to_none=True)
        m.train_step(b.x, b.edge_index, b.batch, b.ptr, action=action, target=target, weight=weight)
        g = m.flat_gradient.clone()
        n_body, n_head = 3 * cfg["layers"], 3 * 2 + 6          # tensors of the body / of one head
        for j, (rp, (o, n, shp)) in enumerate(zip(ref_params, m._slices)):
            gi = g[o:o + n].view(shp).clone()
            taken = j < n_body or ((j - n_body) // n_head == (0 if maker else 1))
            if taken:
                rp.grad = gi
            elif skip_untaken or rp.grad is None:
                rp.grad = None                                 # never trained yet, or torch >= 2 set_to_none
            else:
                rp.grad = torch.zeros_like(rp)                 # torch 1.13 opt.zero_grad(): the tensor stays, zeroed
        live = [rp for rp in ref_params if rp.grad is not None]
        norm_ref = torch.nn.utils.clip_grad_norm_(live, 0.05)
        ref_opt.step()
        norm = opt.step(b.x)
        assert abs(float(norm) - float(norm_ref)) <= 2e-6 * float(norm_ref)
        flat_ref = torch.cat([rp.detach().reshape(-1) for rp in ref_params])
        err = (m.flat_parameters - flat_ref).abs().max().item()
        assert err <= 2e-6, (step, err)
    # target sync: one D2D copy
    _, tgt = make_pair(cfg["hidden"], cfg["layers"], seed=5)
    tgt = tgt.to(DEV)
    tgt.copy_parameters_from(m)
    assert torch.equal(tgt.flat_parameters, m.flat_parameters)
    assert torch.equal(tgt.gnn.convs[3].lin_l.weight, m.gnn.convs[3].lin_l.weight)


@pytest.mark.gpu
def test_device_collate_bit_exact():
    """hexgnn_collate (one launch over device-resident graphs) == host collate == Batch.from_data_list layout
    (replay_buffer.py:215-216, util.cpp:22-41), including an edgeless single-node graph and repeated use of the slots."""
    col = data.DeviceCollator(DEV, max_graphs=512, slots=2)
    gs = positions.load_positions("pos_mix_5_13_seed1234.npz", maker=1)[:200]
    lone = data.HexGraph(x=torch.tensor([[0.0, 1.0, 1.0]]), edge_index=torch.zeros((2, 0), dtype=torch.long))
    for lo, hi in [(0, 200), (3, 4), (50, 97), (100, 101), (7, 150)]:
        part = gs[lo:hi] + ([lone] if hi - lo > 1 else [])
        ref = data.collate(part)
        out = col([g.to(DEV) for g in part])
        assert out.num_graphs == ref.num_graphs
        for name in ("x", "edge_index", "batch", "ptr"):
            a, b = getattr(out, name).cpu(), getattr(ref, name)
            assert a.dtype == b.dtype and a.shape == b.shape and torch.equal(a, b), name
    with pytest.raises(_lib.HexGnnError):
        col([gs[0]])          # host tensors are refused: no CPU path


@pytest.mark.gpu
@pytest.mark.parametrize("double_dqn", [True, False])
def test_rainbow_learner_step_matches_reference_formulas(double_dqn):
    """RainbowLearner.train (CSR x2, 3 forwards, mask/argmax, TD target, train_step, clip + Adam: all on the device) against
    the same step written with the oracle modules exactly as rainbow.py:259-277,301-326 writes it."""
    from gnn_hex_b200 import RainbowLearner
    cfg = CONFIGS["S"]
    o_pol, m_pol = make_pair(cfg["hidden"], cfg["layers"], seed=0)
    o_tgt, m_tgt = make_pair(cfg["hidden"], cfg["layers"], seed=5)
    m_pol, m_tgt = m_pol.to(DEV), m_tgt.to(DEV)
    gs = positions.load_positions(cfg["fixture"], maker=1)
    hs, hn = data.collate(gs[:64]), data.collate(gs[64:128])
    g = torch.Generator().manual_seed(3)
    action = torch.stack([torch.randint(2, int(n), (1,), generator=g)[0] for n in (hs.ptr[1:] - hs.ptr[:-1])])
    reward = (torch.rand(64, generator=g) < 0.2).float() * torch.sign(torch.randn(64, generator=g))
    done = (torch.rand(64, generator=g) < 0.3).float()
    weights = torch.rand(64, generator=g) * 0.5 + 0.5
    lr, eps, max_norm, gamma, n_step = 1e-3, 1.5e-4, 0.02, 0.97, 2
    # ---- reference formulas on the oracle (CPU) ----
    with torch.no_grad():
        if double_dqn:
            adv = o_pol(hn.x, hn.edge_index, hn.batch, hn.ptr, advantages_only=True).clone()
            adv[hn.ptr[:-1]] = -5
            adv[hn.ptr[:-1] + 1] = -5
            best = torch.stack([hn.ptr[i] + adv[hn.ptr[i]:hn.ptr[i + 1]].reshape(-1).argmax() for i in range(64)])
            next_q = o_tgt(hn.x, hn.edge_index, hn.batch, hn.ptr).squeeze()[best]
        else:
            qt = o_tgt(hn.x, hn.edge_index, hn.batch, hn.ptr).clone()
            qt[hn.ptr[:-1]] = -5
            qt[hn.ptr[:-1] + 1] = -5
            next_q = torch.stack([qt[hn.ptr[i]:hn.ptr[i + 1]].max() for i in range(64)])
        td_tgt = reward + gamma ** n_step * next_q * (1 - done)
    opt = torch.optim.Adam(o_pol.parameters(), lr=lr, eps=eps)
    opt.zero_grad()
    p_res = o_pol(hs.x, hs.edge_index, hs.batch, hs.ptr)
    td_est = p_res[action + hs.ptr[:-1]]
    loss_ref = torch.mean(weights * (td_tgt - td_est) ** 2)
    loss_ref.backward()
    live = [p for p in o_pol.parameters() if p.grad is not None]
    norm_ref = torch.nn.utils.clip_grad_norm_(live, max_norm)
    opt.step()
    # ---- the device step ----
    lrn = RainbowLearner(m_pol, m_tgt, lr=lr, adam_eps=eps, max_grad_norm=max_norm, gamma=gamma, n_step=n_step,
                         double_dqn=double_dqn, loss_fn="mse")
    out = lrn.train(hs.to(DEV), hn.to(DEV), action.to(DEV), reward.to(DEV), done.to(DEV), weights.to(DEV))
    assert rel_err(out["td_target"], td_tgt) <= REL_TOL
    assert abs(float(out["loss"]) - float(loss_ref)) <= 1e-5 * max(abs(float(loss_ref)), 1e-3)
    assert abs(float(out["grad_norm"]) - float(norm_ref)) <= 2e-5 * float(norm_ref)
    assert torch.allclose(out["priorities"].cpu(), (td_est - td_tgt).detach().abs() + 1e-6, atol=2e-6)
    ref_sd = dict(o_pol.named_parameters())
    for name, p in m_pol.named_parameters():
        assert torch.allclose(p.detach().cpu(), ref_sd[name].detach(), atol=1e-5, rtol=0), name
    lrn.sync_Q_target()
    assert torch.equal(m_tgt.flat_parameters, m_pol.flat_parameters)


@pytest.mark.gpu
@pytest.mark.parametrize("count", [1, 3, 700])
def test_train_step_batch_extremes(count):
    """hexgnn_train_step on a single graph, on three graphs (fewer tiles than CTAs: every CTA but a few idles, the dW kernel
    has no phase-A items) and on 700 graphs (5 rounds of tiles: multi-round phase A, ranked tile order): loss against the
    fp32 oracle, gradients bit-identical to the forward -> td_loss -> backward route and repeatable."""
    cfg = CONFIGS["L"]
    o, m = make_pair(cfg["hidden"], cfg["layers"])
    m = m.to(DEV)
    gs = positions.load_positions(cfg["fixture"], maker=1)
    gs = (gs * 3)[:count]
    hb = data.collate(gs)
    b = hb.to(DEV)
    action, target, weight = loss_inputs(hb)
    with torch.no_grad():
        loss_ref = orc.rainbow_loss(o(hb.x, hb.edge_index, hb.batch, hb.ptr), hb.ptr, action, target, weight)
    action, target, weight = action.to(DEV), target.to(DEV), weight.to(DEV)
    m.zero_grad(set_to_none=True)
    q_ref = m(b.x, b.edge_index, b.batch, b.ptr)
    ops.td_loss(q_ref, b.ptr, action, target, weight)[0].backward()
    g_ref = m.flat_gradient.clone()
    for _ in range(2):
        m.zero_grad(set_to_none=True)
        loss, err, q = m.train_step(b.x, b.edge_index, b.batch, b.ptr, action=action, target=target, weight=weight)
        assert int(m.last_csr.status.item()) == 0
        assert torch.equal(q, q_ref.detach()) and torch.equal(m.flat_gradient, g_ref)
        assert abs(float(loss) - float(loss_ref)) <= 1e-5 * max(abs(float(loss_ref)), 1e-3)
    assert torch.isfinite(g_ref).all() and float(g_ref.abs().max()) > 0


def _loss_backward_pair(o, m, b, maker_unused=None):
    """Same rainbow loss through the oracle (fp32 + fp64) and the product module; returns (o64, loss, loss_ref)."""
    d = b.to(DEV)
    action, target, weight = loss_inputs(b)
    loss_ref = orc.rainbow_loss(o(b.x, b.edge_index, b.batch, b.ptr), b.ptr, action, target, weight)
    loss_ref.backward()
    o64 = oracle64(o)
    orc.rainbow_loss(o64(b.x.double(), b.edge_index, b.batch, b.ptr), b.ptr, action, target.double(),
                     weight.double()).backward()
    q = m(d.x, d.edge_index, d.batch, d.ptr)
    loss = orc.rainbow_loss(q, d.ptr, action.to(DEV), target.to(DEV), weight.to(DEV))
    loss.backward()
    return o64, loss, loss_ref


@pytest.mark.parametrize("maker", [1, 0])
def test_noisy_dqn_forward_backward_parity(maker):
    """`--noisy_dqn=True` (models.py:84-141,331-334): composed advantage weights in front of the fused kernels, d_mu / d_sigma
    behind them; reset_noise / disable_noise (rainbow.py:167-192) take effect at the next call."""
    o, m = make_pair(35, 10, noisy_dqn=True)
    m = m.to(DEV)
    b = _batch(CONFIGS["S"], "play", maker, 48)
    d = b.to(DEV)
    with torch.no_grad():
        q_ref = o(b.x, b.edge_index, b.batch, b.ptr)
        a_ref = o(b.x, b.edge_index, b.batch, advantages_only=True)
        q = m(d.x, d.edge_index, d.batch, d.ptr)
        a = m(d.x, d.edge_index, d.batch, d.ptr, advantages_only=True)
    assert rel_err(q, q_ref) <= REL_TOL and rel_err(a, a_ref) <= REL_TOL
    o64, loss, loss_ref = _loss_backward_pair(o, m, b)
    assert rel_err(loss, loss_ref) <= REL_TOL
    glob, worst, wname = grad_errs(m, o, o64)
    assert glob <= REL_TOL, glob
    assert worst <= 1.0, (worst, wname)
    head = m.maker_head if maker == 1 else m.breaker_head
    head_o = o.maker_head if maker == 1 else o.breaker_head
    assert float(head.linear.weight_sigma.grad.abs().max()) > 0
    # d_sigma = d_mu * epsilon exactly (one fp32 product)
    assert torch.equal(head.linear.weight_sigma.grad, head.linear.weight_mu.grad * head.linear.weight_epsilon)
    # a second backward without zero_grad accumulates, like autograd
    g1 = head.linear.weight_sigma.grad.clone()
    b1 = head.gnn.convs[0].lin_l.weight.grad.clone()
    action, target, weight = loss_inputs(b)
    orc.rainbow_loss(m(d.x, d.edge_index, d.batch, d.ptr), d.ptr, action.to(DEV), target.to(DEV), weight.to(DEV)).backward()
    assert torch.allclose(head.linear.weight_sigma.grad, 2 * g1, rtol=1e-6, atol=0)
    assert torch.allclose(head.gnn.convs[0].lin_l.weight.grad, 2 * b1, rtol=1e-6, atol=0)
    m.zero_grad()
    assert head.linear.weight_sigma.grad is None and head.linear.bias_mu.grad is None
    # train_step == the autograd route, bit for bit, noisy parameters included
    m.zero_grad()
    ops.td_loss(m(d.x, d.edge_index, d.batch, d.ptr), d.ptr, action.to(DEV), target.to(DEV), weight.to(DEV))[0].backward()
    want = {k: p.grad.clone() for k, p in m.named_parameters()}
    m.zero_grad()
    m.train_step(d.x, d.edge_index, d.batch, d.ptr, action=action.to(DEV), target=target.to(DEV), weight=weight.to(DEV))
    for k, p in m.named_parameters():
        assert torch.equal(p.grad, want[k]), k
    # new noise -> different advantages, still equal to the oracle with the same noise; no noise -> plain mu
    torch.manual_seed(9)
    head.linear.reset_noise()
    head_o.linear.weight_epsilon.copy_(head.linear.weight_epsilon.cpu())
    head_o.linear.bias_epsilon.copy_(head.linear.bias_epsilon.cpu())
    with torch.no_grad():
        a2 = m(d.x, d.edge_index, d.batch, d.ptr, advantages_only=True)
        assert rel_err(a2, o(b.x, b.edge_index, b.batch, advantages_only=True)) <= REL_TOL
        assert not torch.equal(a2, a)
        head.linear.disable_noise()
        head_o.linear.disable_noise()
        a3 = m(d.x, d.edge_index, d.batch, d.ptr, advantages_only=True)
        assert rel_err(a3, o(b.x, b.edge_index, b.batch, advantages_only=True)) <= REL_TOL
    # target-network sync carries mu, sigma and the noise (state_dict has all three, rainbow.py:163-164)
    _, t = make_pair(35, 10, seed=3, noisy_dqn=True)
    t = t.to(DEV)
    t.copy_parameters_from(m)
    with torch.no_grad():
        assert torch.equal(t(d.x, d.edge_index, d.batch, d.ptr, advantages_only=True), a3)
    # the flat-buffer optimizer does not cover mu / sigma: it must say so, torch.optim.Adam works on the named parameters
    from gnn_hex_b200 import FusedAdam
    with pytest.raises(NotImplementedError):
        FusedAdam(m).step(d.x)
    opt = torch.optim.Adam(m.parameters(), lr=1e-3)
    m.zero_grad()
    m.train_step(d.x, d.edge_index, d.batch, d.ptr, action=action.to(DEV), target=target.to(DEV), weight=weight.to(DEV))
    mu0 = head.linear.weight_mu.detach().clone()
    opt.step()
    assert not torch.equal(head.linear.weight_mu.detach(), mu0)


def test_grown_model_parity():
    """grow_depth / grow_width (models.py:166-238,336-361,494-508; train.py:301-309): the fused path follows the new shape."""
    o, m = make_pair(20, 4)
    m = m.to(DEV)
    b = _batch(CONFIGS["S"], "play", 1, 40)
    d = b.to(DEV)
    with torch.no_grad():
        q0 = m(d.x, d.edge_index, d.batch, d.ptr)
    torch.manual_seed(21)
    o.grow_depth(2)
    m.grow_depth(2)
    with torch.no_grad():
        q1 = m(d.x, d.edge_index, d.batch, d.ptr)
    assert rel_err(q1, q0) <= REL_TOL                      # identity layers: same function
    torch.manual_seed(22)
    o.grow_width(35)
    o.maker_head.grow_depth(1); o.breaker_head.grow_depth(1)
    torch.manual_seed(22)
    m.grow_width(35)
    m.maker_head.grow_depth(1)
    with pytest.raises(_lib.HexGnnError):                  # heads of different depth: refused, not silently wrong
        m(d.x, d.edge_index, d.batch, d.ptr)
    m.breaker_head.grow_depth(1)
    m.load_state_dict(o.state_dict())                      # (grow_width draws on the device RNG here, on the CPU there)
    with torch.no_grad():
        q_ref = o(b.x, b.edge_index, b.batch, b.ptr)
        q2 = m(d.x, d.edge_index, d.batch, d.ptr)
    assert (m.hidden_channels, m.num_layers, m.num_head_layers) == (35, 6, 3)
    assert rel_err(q2, q_ref) <= REL_TOL
    o64, loss, loss_ref = _loss_backward_pair(o, m, b)
    assert rel_err(loss, loss_ref) <= REL_TOL
    glob, worst, wname = grad_errs(m, o, o64)
    assert glob <= REL_TOL, glob
    assert worst <= 1.0, (worst, wname)


def test_custom_ops_match_the_module():
    """torch.ops.hexgnn.* (functional, flat parameters) == the nn.Module route, bit for bit, gradient included."""
    _, m = make_pair(35, 10)
    m = m.to(DEV)
    hb = _batch(CONFIGS["S"], "play", 1, 64)
    b = hb.to(DEV)
    action, target, weight = [t.to(DEV) for t in loss_inputs(hb)]
    N, E, B = b.x.shape[0], b.edge_index.shape[1], hb.num_graphs
    m.zero_grad()
    q_ref = m(b.x, b.edge_index, b.batch, b.ptr)
    ops.td_loss(q_ref, b.ptr, action, target, weight)[0].backward()
    g_ref = m.flat_gradient.clone()
    csr = torch.ops.hexgnn.build_csr(b.edge_index, b.batch, b.ptr, N, B, 128)
    mine = ops.CsrBlob(csr, N, E, B, 128)         # (compare the CSR's fields: padding / builder scratch words are unset)
    assert csr.numel() == m.last_csr.blob.numel()
    for f in ("status", "ntiles", "gptr", "rowptr", "col"):
        assert torch.equal(getattr(mine, f), getattr(m.last_csr, f)), f
    params = m.flat_parameters.detach().clone().requires_grad_(True)
    q, v, stash = torch.ops.hexgnn.fused_forward(params, b.x, csr, E, B, 35, 10, 2, 128, 0, True)
    assert torch.equal(q, q_ref.detach())
    ops.td_loss(q, b.ptr, action, target, weight)[0].backward()
    assert torch.equal(params.grad, g_ref)
    with torch.no_grad():
        a, _, s0 = torch.ops.hexgnn.fused_forward(params, b.x, csr, E, B, 35, 10, 2, 128, ops.HEXGNN_ADV_ONLY, False)
        assert s0.numel() == 0
        assert torch.equal(a, m(b.x, b.edge_index, b.batch, b.ptr, advantages_only=True).reshape(-1))
    torch.library.opcheck(torch.ops.hexgnn.fused_forward, (params, b.x, csr, E, B, 35, 10, 2, 128, 0, True),
                          test_utils=("test_schema", "test_faketensor", "test_autograd_registration"))
    torch.library.opcheck(torch.ops.hexgnn.build_csr, (b.edge_index, b.batch, b.ptr, N, B, 128),
                          test_utils=("test_schema", "test_faketensor"))


def test_packed_pinned_batch_is_one_copy_and_bit_exact():
    """HexBatch.pin_memory() packs the four fields into one pinned buffer; .to(cuda) is a single copy whose views feed the
    same kernels: identical CSR and Q-values."""
    _, m = make_pair(35, 10)
    m = m.to(DEV)
    hb = _batch(CONFIGS["S"], "play", 0, 33)
    pb = hb.pin_memory()
    assert pb._packed.is_pinned() and pb.x.data_ptr() >= pb._packed.data_ptr()
    d0, d1 = hb.to(DEV), pb.to(DEV, non_blocking=True)
    for f in ("x", "edge_index", "batch", "ptr"):
        assert torch.equal(getattr(d0, f), getattr(d1, f)) and getattr(d1, f).data_ptr() % 256 == 0
    assert d1._packed.data_ptr() == d1.edge_index.data_ptr()
    with torch.no_grad():
        q0 = m(d0.x, d0.edge_index, d0.batch, d0.ptr)
        csr0 = m.last_csr.blob.clone()
        q1 = m(d1.x, d1.edge_index, d1.batch, d1.ptr)
    assert torch.equal(q0, q1)
    c0 = ops.CsrBlob(csr0, m.last_csr.N, m.last_csr.E, m.last_csr.B, 128)
    for f in ("status", "ntiles", "gptr", "rowptr", "col"):
        assert torch.equal(getattr(c0, f), getattr(m.last_csr, f)), f


@pytest.mark.parametrize("name,maker,count", [("S", 1, 64), ("L", 0, 24)])
def test_input_gradient_parity(name, maker, count):
    """dL/dx (`input.grad` after loss.backward(), rl_loop/explain.py:26-36) through hexgnn_backward_dx, against the fp64 oracle;
    column 2 (head selector) gets zero; the parameter gradient is the same bits as without the input gradient."""
    cfg = CONFIGS[name]
    o, m = make_pair(cfg["hidden"], cfg["layers"])
    m = m.to(DEV)
    b = _batch(cfg, "play", maker, count)
    d = b.to(DEV)
    action, target, weight = loss_inputs(b)
    o64 = oracle64(o)
    x64 = b.x.double().requires_grad_(True)
    orc.rainbow_loss(o64(x64, b.edge_index, b.batch, b.ptr), b.ptr, action, target.double(), weight.double()).backward()
    x32 = b.x.clone().requires_grad_(True)
    orc.rainbow_loss(o(x32, b.edge_index, b.batch, b.ptr), b.ptr, action, target, weight).backward()
    m.zero_grad()
    ops.td_loss(m(d.x, d.edge_index, d.batch, d.ptr), d.ptr, action.to(DEV), target.to(DEV), weight.to(DEV))[0].backward()
    g_plain = m.flat_gradient.clone()
    m.zero_grad()
    xg = d.x.clone().requires_grad_(True)
    ops.td_loss(m(xg, d.edge_index, d.batch, d.ptr), d.ptr, action.to(DEV), target.to(DEV), weight.to(DEV))[0].backward()
    assert xg.grad is not None and xg.grad.shape == d.x.shape
    assert torch.equal(m.flat_gradient, g_plain)
    assert float(xg.grad[:, 2].abs().max()) == 0.0 and float(x64.grad[:, 2].abs().max()) == 0.0
    err = rel_err(xg.grad, x64.grad)
    err32 = rel_err(x32.grad, x64.grad)
    assert err <= max(4e-5, 16 * err32), (err, err32)


def test_rainbow_learner_noisy_nets_and_act():
    """RainbowLearner with `noisy_dqn=True`: td_target resets the target net's noise (rainbow.py:261), the step uses the
    reference's clip_grad_norm_ + torch Adam over the named parameters (mu / sigma included); `act` (rainbow.py:214-239)."""
    from gnn_hex_b200 import RainbowLearner
    cfg = CONFIGS["S"]
    o_pol, m_pol = make_pair(cfg["hidden"], cfg["layers"], seed=0, noisy_dqn=True)
    o_tgt, m_tgt = make_pair(cfg["hidden"], cfg["layers"], seed=5, noisy_dqn=True)
    m_pol, m_tgt = m_pol.to(DEV), m_tgt.to(DEV)
    gs = positions.load_positions(cfg["fixture"], maker=1)
    hs, hn = data.collate(gs[:48]), data.collate(gs[48:96])
    g = torch.Generator().manual_seed(3)
    action = torch.stack([torch.randint(2, int(n), (1,), generator=g)[0] for n in (hs.ptr[1:] - hs.ptr[:-1])])
    reward = (torch.rand(48, generator=g) < 0.2).float()
    done = (torch.rand(48, generator=g) < 0.3).float()
    weights = torch.rand(48, generator=g) * 0.5 + 0.5
    lr, eps, max_norm, gamma, n_step = 1e-3, 1.5e-4, 0.02, 0.97, 2
    lrn = RainbowLearner(m_pol, m_tgt, lr=lr, adam_eps=eps, max_grad_norm=max_norm, gamma=gamma, n_step=n_step)
    eps_before = m_tgt.maker_head.linear.weight_epsilon.clone()
    out = lrn.train(hs.to(DEV), hn.to(DEV), action.to(DEV), reward.to(DEV), done.to(DEV), weights.to(DEV))
    assert not torch.equal(m_tgt.maker_head.linear.weight_epsilon, eps_before)      # the target's noise was redrawn
    for ho, hm in ((o_tgt.maker_head, m_tgt.maker_head), (o_tgt.breaker_head, m_tgt.breaker_head)):
        ho.linear.weight_epsilon.copy_(hm.linear.weight_epsilon.cpu())
        ho.linear.bias_epsilon.copy_(hm.linear.bias_epsilon.cpu())
    with torch.no_grad():
        adv = o_pol(hn.x, hn.edge_index, hn.batch, hn.ptr, advantages_only=True).clone()
        adv[hn.ptr[:-1]] = -5
        adv[hn.ptr[:-1] + 1] = -5
        best = torch.stack([hn.ptr[i] + adv[hn.ptr[i]:hn.ptr[i + 1]].reshape(-1).argmax() for i in range(48)])
        td_tgt = reward + gamma ** n_step * o_tgt(hn.x, hn.edge_index, hn.batch, hn.ptr).squeeze()[best] * (1 - done)
    opt = torch.optim.Adam(o_pol.parameters(), lr=lr, eps=eps)
    opt.zero_grad()
    td_est = o_pol(hs.x, hs.edge_index, hs.batch, hs.ptr)[action + hs.ptr[:-1]]
    loss_ref = torch.mean(weights * (td_tgt - td_est) ** 2)
    loss_ref.backward()
    norm_ref = torch.nn.utils.clip_grad_norm_([p for p in o_pol.parameters() if p.grad is not None], max_norm)
    opt.step()
    assert rel_err(out["td_target"], td_tgt) <= REL_TOL
    assert abs(float(out["loss"]) - float(loss_ref)) <= 1e-5 * max(abs(float(loss_ref)), 1e-3)
    assert abs(float(out["grad_norm"]) - float(norm_ref)) <= 2e-5 * float(norm_ref)
    ref_sd = dict(o_pol.named_parameters())
    for name, p in m_pol.named_parameters():
        assert torch.allclose(p.detach().cpu(), ref_sd[name].detach(), atol=1e-5, rtol=0), name
    assert not torch.equal(m_pol.maker_head.linear.weight_sigma.detach().cpu(),
                           torch.full_like(ref_sd["maker_head.linear.weight_sigma"], 0.5 / cfg["hidden"] ** 0.5))
    # act: greedy == arg-max of the oracle's masked advantages (noise disabled on both sides); eps = 1 explores everywhere
    m_pol.disable_noise()
    for h in (o_pol.maker_head, o_pol.breaker_head):
        h.linear.disable_noise()
    dn = hn.to(DEV)
    actions, expl = lrn.act(dn, eps=0.0)
    with torch.no_grad():
        adv = o_pol(hn.x, hn.edge_index, hn.batch, hn.ptr, advantages_only=True).clone().reshape(-1)
    adv[hn.ptr[:-1]] = -5
    adv[hn.ptr[:-1] + 1] = -5
    want = torch.stack([adv[hn.ptr[i]:hn.ptr[i + 1]].argmax() for i in range(48)])
    top2 = [adv[hn.ptr[i]:hn.ptr[i + 1]].topk(2).values for i in range(48)]
    clear = torch.tensor([float(t[0] - t[1]) > 1e-4 for t in top2])
    assert torch.equal(actions.cpu()[clear], want[clear]) and not expl.any() and int(actions.min()) >= 2
    actions, expl = lrn.act(dn, eps=1.0)
    n = (hn.ptr[1:] - hn.ptr[:-1])
    assert expl.all() and bool(((actions.cpu() >= 2) & (actions.cpu() < n)).all())


def test_full_size_headline_batch_properties():
    """BASELINE configs[1] at its full size (GNN-L, 256 11x11 positions), through the properties the domain offers:
    Q and loss against the fp32 oracle and the whole gradient against the fp64 arbiter; graphs are independent, so permuting
    the GRAPH order of the batch permutes Q (to round-off: tiles are re-packed) and moves the gradient only by summation
    order; the gradient is linear in the loss scale (a power of two: bit-exact)."""
    cfg = CONFIGS["L"]
    o, m = make_pair(cfg["hidden"], cfg["layers"])
    m = m.to(DEV)
    gs = positions.load_positions(cfg["fixture"], maker=1)[:256]
    hb = data.collate(gs)
    b = hb.to(DEV)
    action, target, weight = loss_inputs(hb)
    dev_in = [t.to(DEV) for t in (action, target, weight)]
    m.zero_grad()
    loss, td_err, q = m.train_step(b.x, b.edge_index, b.batch, b.ptr, action=dev_in[0], target=dev_in[1], weight=dev_in[2])
    g1 = m.flat_gradient.clone()
    assert int(m.last_csr.status.item()) == 0
    q_ref = o(hb.x, hb.edge_index, hb.batch, hb.ptr)
    loss_ref = orc.rainbow_loss(q_ref, hb.ptr, action, target, weight)
    loss_ref.backward()
    o64 = oracle64(o)
    orc.rainbow_loss(o64(hb.x.double(), hb.edge_index, hb.batch, hb.ptr), hb.ptr, action, target.double(),
                     weight.double()).backward()
    assert rel_err(q, q_ref) <= REL_TOL and rel_err(loss, loss_ref) <= REL_TOL
    glob, worst, wname = grad_errs(m, o, o64)
    assert glob <= REL_TOL, glob
    assert worst <= 1.0, (worst, wname)
    # linearity in the loss scale
    m.zero_grad()
    m.train_step(b.x, b.edge_index, b.batch, b.ptr, action=dev_in[0], target=dev_in[1], weight=dev_in[2], scale=0.25)
    assert torch.equal(m.flat_gradient, 0.25 * g1)
    # permutation of the graph order
    perm = torch.randperm(256, generator=torch.Generator().manual_seed(11)).tolist()
    hp = data.collate([gs[i] for i in perm])
    p = hp.to(DEV)
    pa, pt, pw = [t[perm].to(DEV) for t in (action, target, weight)]
    m.zero_grad()
    loss_p, _, q_p = m.train_step(p.x, p.edge_index, p.batch, p.ptr, action=pa, target=pt, weight=pw)
    q_back = torch.cat([q[hb.ptr[i]:hb.ptr[i + 1]] for i in perm])
    assert rel_err(q_p, q_back) <= 2e-6           # (not bit-exact: a row's position in its tile sets the order of its sums)
    assert rel_err(m.flat_gradient, g1) <= REL_TOL
    assert abs(float(loss_p) - float(loss)) <= 1e-6 * abs(float(loss))


def test_csr_five_launch_path_in_subprocess():
    """The builder's fallback (five phase kernels, used when a cooperative launch is refused or the grid cannot be
    co-resident) stays bit-exact: HEXGNN_CSR_BLOCKS_PER_SM=0 disables the fused kernel for one process."""
    import os, subprocess, sys
    code = (
        "import sys, torch; sys.path.insert(0, 'tests')\n"
        "from gnn_hex_b200 import data, ops\n"
        "import positions\n"
        "from oracle import hexgnn_oracle as orc\n"
        "for f in ('pos_11x11_seed1234.npz', 'pos_mix_5_13_seed1234.npz'):\n"
        "    b = data.collate(positions.load_positions(f, maker=1)); d = b.to('cuda:0')\n"
        "    c0 = ops.launch_counter\n"
        "    csr = ops.build_csr(d.edge_index, d.batch, d.ptr, num_nodes=b.x.shape[0])\n"
        "    rowptr, col = orc.csr_by_destination(b.edge_index, b.x.shape[0])\n"
        "    assert torch.equal(csr.rowptr.cpu(), rowptr) and torch.equal(csr.col.cpu(), col)\n"
        "    assert torch.equal(csr.gptr.cpu().long(), b.ptr) and int(csr.ntiles.item()) > 0\n"
        "print('FIVE_LAUNCH_OK')\n")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, HEXGNN_CSR_BLOCKS_PER_SM="0", PYTHONPATH=root)
    r = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=300)
    assert "FIVE_LAUNCH_OK" in r.stdout, r.stdout + r.stderr


def test_tf32x1_inference_mode_is_opt_in_and_bounded():
    """SURVEY 7.2(1) precision policy: `inference_precision="tf32x1"` runs ONE tensor-core pass per k-step in forwards that
    keep no stash (acting / target evaluation, rainbow.py:214-239,259-275): measurably coarser than the 3xTF32 default but
    within 1e-2, same arg-max as the default on clear-cut positions; differentiable calls ignore it (bit-identical to
    tf32x3); the C-ABI refuses the flag together with a stash."""
    import ctypes as C
    cfg = CONFIGS["L"]
    o, m = make_pair(cfg["hidden"], cfg["layers"])
    m = m.to(DEV)
    b = _batch(cfg, "play", 1, 48)
    d = b.to(DEV)
    with torch.no_grad():
        a_ref = o(b.x, b.edge_index, b.batch, advantages_only=True)
        a3 = m(d.x, d.edge_index, d.batch, d.ptr, advantages_only=True)
        m.inference_precision = "tf32x1"
        a1 = m(d.x, d.edge_index, d.batch, d.ptr, advantages_only=True)
        q1 = m(d.x, d.edge_index, d.batch, d.ptr)
    e3, e1 = rel_err(a3, a_ref), rel_err(a1, a_ref)
    assert e3 <= REL_TOL and REL_TOL < e1 <= 1e-2, (e3, e1)
    assert rel_err(q1, o(b.x, b.edge_index, b.batch, b.ptr).detach()) <= 1e-2
    # a call that can be differentiated always runs 3xTF32
    q_grad = m(d.x, d.edge_index, d.batch, d.ptr)
    m.inference_precision = "tf32x3"
    assert torch.equal(q_grad.detach(), m(d.x, d.edge_index, d.batch, d.ptr).detach())
    with pytest.raises(ValueError):
        ModernTwoHeaded(12, 3, 2, inference_precision="bf16")
    csr = m.last_csr
    sz = _lib.sizes(m._desc, csr.N, csr.E, csr.B, 128)
    stash = torch.empty(sz.stash_bytes // 4, device=DEV)
    ws = torch.empty(sz.fwd_ws_bytes // 4, device=DEV)
    out = torch.empty(csr.N, device=DEV)
    rc = _lib.lib().hexgnn_forward(C.byref(m._desc), C.c_void_p(m._flat.data_ptr()), C.c_void_p(d.x.data_ptr()), csr.N, csr.E,
                                   csr.B, C.c_void_p(csr.blob.data_ptr()), 128, ops.HEXGNN_TF32X1, C.c_void_p(out.data_ptr()),
                                   None, C.c_void_p(stash.data_ptr()), C.c_void_p(ws.data_ptr()), ws.numel() * 4, None)
    assert rc == -1 and b"TF32X1" in _lib.lib().hexgnn_last_error()


def test_fp16_range_guard_poisons_and_flags():
    """The forward's tensor-core products are exact fp16 x fp16 (all-fp16 3-product scheme, csrc/sage_tc.cuh) while the
    activations stay below 65504.  Beyond that the tile is flagged (status bit 16) and its outputs are NaN -- never a
    silently wrong Q."""
    cfg = CONFIGS["S"]
    o, m = make_pair(cfg["hidden"], cfg["layers"])
    with torch.no_grad():
        m.gnn.convs[0].lin_r.weight.mul_(3e5)          # layer-0 output ~1e5 .. 1e6 for the degree feature
    m = m.to(DEV)
    b = _batch(cfg, "play", 1, 8)
    d = b.to(DEV)
    with torch.no_grad():
        q = m(d.x, d.edge_index, d.batch, d.ptr)
    assert int(m.last_csr.status.item()) & _lib.ST_FP16_RANGE
    assert torch.isnan(q).all()
    m.strict = True
    with pytest.raises(_lib.HexGnnError, match="fp16 range"):
        with torch.no_grad():
            m(d.x, d.edge_index, d.batch, d.ptr)


def test_fused_adam_skips_a_non_finite_step():
    """one NaN gradient (an oversize graph's poisoned record) must not reach the parameters: the fused optimizer
    leaves parameters, moments and step counters untouched and reports the non-finite norm."""
    from gnn_hex_b200 import FusedAdam
    cfg = CONFIGS["S"]
    o, m = make_pair(cfg["hidden"], cfg["layers"])
    m = m.to(DEV)
    b = _batch(cfg, "play", 1, 16)
    d = b.to(DEV)
    action, target, weight = [t.to(DEV) for t in loss_inputs(b)]
    opt = FusedAdam(m, lr=1e-3, max_grad_norm=10.0)
    m.train_step(d.x, d.edge_index, d.batch, d.ptr, action=action, target=target, weight=weight)
    opt.step(d.x)                                      # a regular step: counters (1, 1, 0, taken = 1)
    p0, m0, v0, s0 = m.flat_parameters.clone(), opt.exp_avg.clone(), opt.exp_avg_sq.clone(), opt.state.clone()
    m.zero_grad(set_to_none=True)
    m.train_step(d.x, d.edge_index, d.batch, d.ptr, action=action, target=target, weight=weight)
    g_before = m.flat_gradient.clone()
    m.flat_gradient[123] = float("nan")
    norm = opt.step(d.x)
    torch.cuda.synchronize()
    assert not torch.isfinite(norm)
    assert torch.equal(m.flat_parameters, p0) and torch.equal(opt.exp_avg, m0) and torch.equal(opt.exp_avg_sq, v0)
    assert torch.equal(opt.state, s0)
    g_after = m.flat_gradient.clone()
    g_after[123] = g_before[123]
    assert torch.equal(g_after, g_before)              # not even the in-place clip touched the gradient
    m.flat_gradient[123] = g_before[123]
    assert torch.isfinite(opt.step(d.x))               # the next finite step goes through
    assert int(opt.state[0]) == int(s0[0]) + 1 and not torch.equal(m.flat_parameters, p0)
