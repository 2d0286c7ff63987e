"""GPU: the layer-at-a-time tensor-core path (hexgnn_forward_lw / hexgnn_backward_lw through the module API): norm=True
(PyG LayerNorm in graph mode, GN0/models.py:283-287,481-482,550-551,935,945) and graphs beyond one 128-row tile, training
included (board-size transfer up to 19x19: GN0/RainbowDQN/evaluate_elo.py:228-251, GN0/util/plotting.py:61-80)."""
import numpy as np
import pytest
import torch

import helpers
import positions
from helpers import REL_TOL, loss_inputs, rel_err
from gnn_hex_b200 import FusedAdam, ModernTwoHeaded, _lib, data, get_pre_defined, ops
from oracle import hexgnn_oracle as orc

pytestmark = pytest.mark.gpu
DEV = "cuda"


def make_pair(H, L, Lh=2, norm=False, gain=1.5, seed=3, **kw):
    """Oracle (fp32 + fp64) and product module with the same name-hashed weights (non-degenerate: helpers.set_hash_weights)."""
    o = orc.ModernTwoHeadedOracle(H, L, Lh, norm=norm)
    helpers.set_hash_weights(o, seed=seed, gain=gain)
    o64 = orc.ModernTwoHeadedOracle(H, L, Lh, norm=norm).double()
    o64.load_state_dict({k: v.double() for k, v in o.state_dict().items()})
    m = ModernTwoHeaded(H, L, Lh, norm=norm, **kw)
    m.load_state_dict(o.state_dict())
    return o, o64, m.to(DEV)


def td_loss_of(model, b, action, target, weight, dt):
    q = model(b.x.to(dt), b.edge_index, b.batch, b.ptr)
    return (weight.to(dt) * (q[action + b.ptr[:-1]] - target.to(dt)) ** 2).mean()


def check_grads(m, o, o64):
    """Whole gradient <= 1e-5 of its maximum against the fp64 oracle, or -- per tensor -- inside the band the fp32 oracle's own
    round-off spans (helpers.grad_errs; deep ReLU / LayerNorm stacks amplify ANY fp32 rounding)."""
    glob, worst, name = helpers.grad_errs(m, o, o64)
    assert glob <= REL_TOL or worst <= 1.0, (glob, worst, name)
    return glob, worst


@pytest.mark.parametrize("size,norm", [("S", False), ("S", True), ("L", True)])
@pytest.mark.parametrize("maker", [1, 0])
def test_lw_forward_three_modes(size, norm, maker):
    H, L, fixture, count, gain = {"S": (35, 10, "pos_7x7_seed1234.npz", 96, 1.5),
                                  "L": (110, 15, "pos_11x11_seed1234.npz", 64, 1.7)}[size]
    o, o64, m = make_pair(H, L, norm=norm, gain=gain, path="lw")
    b = data.collate(positions.load_positions(fixture, maker=maker)[:count])
    d = b.to(DEV)
    with torch.no_grad():
        q = m(d.x, d.edge_index, d.batch, d.ptr)
        a = m(d.x, d.edge_index, d.batch, d.ptr, advantages_only=True)
        v, c = m(d.x, d.edge_index, d.batch, d.ptr, seperate=True)
        x64 = b.x.double()
        q64 = o64(x64, b.edge_index, b.batch, b.ptr)
        a64 = o64(x64, b.edge_index, b.batch, b.ptr, advantages_only=True)
        v64, c64 = o64(x64, b.edge_index, b.batch, b.ptr, seperate=True)
    assert int(m.last_csr.status.item()) == 0
    assert q.shape == (b.x.shape[0],) and a.shape == (b.x.shape[0], 1) and v.shape == (b.num_graphs,)
    qs = float(q64.abs().max())
    assert rel_err(q, q64) <= REL_TOL
    assert rel_err(a, a64) <= REL_TOL
    assert float((v.cpu().double() - v64.reshape(-1)).abs().max()) <= REL_TOL * qs
    assert float((c.cpu().double() - c64.reshape(-1)).abs().max()) <= REL_TOL * qs


@pytest.mark.parametrize("size,norm,mode", [("S", False, "q"), ("S", True, "q"), ("S", True, "sep"), ("S", True, "adv"),
                                            ("H60", True, "q")])
def test_lw_autograd_gradients(size, norm, mode):
    """loss.backward() (rainbow.py:316) through the lw path: every parameter gradient -- LayerNorm weights included --
    against the fp64 oracle."""
    H, L, count = {"S": (35, 10, 64), "H60": (60, 4, 16)}[size]
    o, o64, m = make_pair(H, L, norm=norm, path="lw")
    b = data.collate(positions.load_positions("pos_7x7_seed1234.npz", maker=1)[:count])
    d = b.to(DEV)
    action, target, weight = loss_inputs(b)
    idx = action + b.ptr[:-1]

    def loss_of(model, bb, dt, dev):
        x = bb.x.to(dt)
        w, t, i = weight.to(dev).to(dt), target.to(dev).to(dt), idx.to(dev)
        if mode == "q":
            return (w * (model(x, bb.edge_index, bb.batch, bb.ptr)[i] - t) ** 2).mean()
        if mode == "adv":
            a = model(x, bb.edge_index, bb.batch, bb.ptr, advantages_only=True).reshape(-1)
            return (w * (a[i] - t) ** 2).mean()
        v, a = model(x, bb.edge_index, bb.batch, bb.ptr, seperate=True)
        return (w * (a.reshape(-1)[i] - t) ** 2).mean() + (v.reshape(-1) * t).mean()

    loss_of(o, b, torch.float32, "cpu").backward()
    l64 = loss_of(o64, b, torch.float64, "cpu")
    l64.backward()
    m.capture_embed_grads = True
    loss = loss_of(m, d, torch.float32, DEV)
    loss.backward()
    assert rel_err(loss, l64) <= REL_TOL
    check_grads(m, o, o64)
    assert m.final_conv_acts.shape == (b.x.shape[0], H) and m.final_conv_grads.shape == (b.x.shape[0], H)
    assert rel_err(m.final_conv_acts, o64.final_conv_acts) <= REL_TOL
    # a second backward before the optimizer step accumulates (autograd's +=), bit-reproducibly
    g1 = m.flat_gradient.clone()
    n1 = m._norm_grad.clone() if norm else None
    loss_of(m, d, torch.float32, DEV).backward()
    assert torch.equal(m.flat_gradient, 2 * g1)
    if norm:
        assert torch.equal(m._norm_grad, 2 * n1)
        assert m.gnn.norms[0].weight.grad.data_ptr() == m._norm_grad.data_ptr()


def test_norm_golden_vectors_of_the_reference_module():
    """norm=True against tests/golden/model_S_norm.npz (the reference's own DuellingTwoHeaded on the PyG LayerNorm stand-in,
    make_model_fixtures.py): Q, loss and every gradient incl. the LayerNorm affine parameters."""
    fx = helpers.load_model_fixture("S_norm")
    rel = lambda a, r: float(np.abs(np.asarray(a, np.float64) - r).max() / max(np.abs(r).max(), 1e-30))
    for maker in (1, 0):
        b = data.collate(positions.load_positions("pos_7x7_seed1234.npz", maker=maker)[:int(fx["count"])])
        args = type("Args", (), dict(hidden_channels=35, num_layers=10, num_head_layers=2, norm=True, noisy_dqn=False))()
        m = get_pre_defined("modern_two_headed", args)
        helpers.set_hash_weights(m, seed=0, gain=float(fx["gain"]))
        names = [str(n) for n in fx["names"]]
        assert sorted(dict(m.named_parameters())) == names
        m = m.cuda()
        d = b.to(DEV)
        action, target, weight = (torch.from_numpy(fx[k]).cuda() for k in ("action", "target", "weight"))
        q = m(d.x, d.edge_index, d.batch, d.ptr)
        loss = torch.mean(weight * (target - q[action + d.ptr[:-1]]) ** 2)
        loss.backward()
        assert rel(q.detach().cpu().numpy(), fx[f"q64_{maker}"]) <= REL_TOL
        assert rel(loss.detach().cpu().numpy(), fx[f"loss_{maker}"]) <= REL_TOL
        g = helpers.flat_named(m, "grad", names)[:: int(fx["gstride"])]
        gmax = float(fx[f"gradmax_{maker}"])
        assert float(np.abs(g - fx[f"grad64_{maker}"]).max()) / gmax <= REL_TOL


def test_board_size_transfer_up_to_19x19():
    """14x14 .. 19x19 positions (198 .. 363 nodes, up to 2020 directed edges) mixed with 11x11 ones: forward against the
    fp64 oracle; the status word stays clean (no tile limit on this path)."""
    o, o64, m = make_pair(110, 15, gain=1.7, max_graph_nodes=363)
    m.strict = True
    big = positions.load_positions("pos_transfer_14_19_seed1234.npz", maker=0)
    small = positions.load_positions("pos_11x11_seed1234.npz", maker=0)[:20]
    b = data.collate(big[:5] + small + big[5:])
    assert int((b.ptr[1:] - b.ptr[:-1]).max()) == 363
    d = b.to(DEV)
    with torch.no_grad():
        q = m(d.x, d.edge_index, d.batch, d.ptr)
        a = m(d.x, d.edge_index, d.batch, d.ptr, advantages_only=True)
        x64 = b.x.double()
        q64 = o64(x64, b.edge_index, b.batch, b.ptr)
        a64 = o64(x64, b.edge_index, b.batch, b.ptr, advantages_only=True)
    assert rel_err(q, q64) <= REL_TOL and rel_err(a, a64) <= REL_TOL
    # batched == per graph (the path keeps graphs independent)
    g = big[0]
    one = data.collate([g]).to(DEV)
    with torch.no_grad():
        q1 = m(one.x, one.edge_index, one.batch, one.ptr)
    assert rel_err(q1, q64[:g.x.shape[0]]) <= REL_TOL


@pytest.mark.parametrize("norm", [False, True])
def test_train_step_with_oversize_graphs(norm):
    """A 5x5..13x13 mix with 16x16 / 19x19 boards through model.train_step (rainbow.py:301-316): loss, TD errors and
    gradients against the fp64 oracle; the same bits as forward -> td_loss -> backward; torch.optim.Adam steps every
    parameter; the flat-buffer optimizer works without norm and refuses with it."""
    o, o64, m = make_pair(35, 10, norm=norm, max_graph_nodes=363)
    mix = positions.load_positions("pos_mix_5_13_seed1234.npz", maker=1)
    big = positions.load_positions("pos_transfer_14_19_seed1234.npz", maker=1)
    gs = mix[:24] + big[4:9] + [g for g in mix if g.x.shape[0] > 128][:4]
    b = data.collate(gs)
    d = b.to(DEV)
    action, target, weight = loss_inputs(b)
    td_loss_of(o, b, action, target, weight, torch.float32).backward()
    l64 = td_loss_of(o64, b, action, target, weight, torch.float64)
    l64.backward()
    A, T, W = action.to(DEV), target.to(DEV), weight.to(DEV)
    loss, td_err, q = m.train_step(d.x, d.edge_index, d.batch, d.ptr, action=A, target=T, weight=W)
    assert int(m.last_csr.status.item()) == 0
    assert rel_err(loss, l64) <= REL_TOL
    with torch.no_grad():
        q64 = o64(b.x.double(), b.edge_index, b.batch, b.ptr)
    assert rel_err(q, q64) <= REL_TOL
    e64 = q64[action + b.ptr[:-1]] - target.double()
    assert float((td_err.cpu().double() - e64).abs().max()) <= REL_TOL * max(1.0, float(q64.abs().max()))
    check_grads(m, o, o64)
    g_ts = m.flat_gradient.clone()
    m.zero_grad()
    qa = m(d.x, d.edge_index, d.batch, d.ptr)
    la, _ = ops.td_loss(qa, d.ptr, A, T, W)
    la.backward()
    assert torch.equal(qa.detach(), q) and torch.equal(la.detach(), loss) and torch.equal(m.flat_gradient, g_ts)
    if norm:
        with pytest.raises(NotImplementedError):
            FusedAdam(m).step(d.x)
        opt = torch.optim.Adam(m.parameters(), lr=1e-3)
        w0 = m.gnn.norms[3].weight.detach().clone()
        c0 = m.gnn.convs[3].lin_l.weight.detach().clone()
        opt.step()
        assert not torch.equal(m.gnn.norms[3].weight.detach(), w0)
        assert not torch.equal(m.gnn.convs[3].lin_l.weight.detach(), c0)
        with torch.no_grad():   # the kernels see the stepped LayerNorm weights (views of the flat norm block)
            assert not torch.equal(m(d.x, d.edge_index, d.batch, d.ptr), q)
    else:
        FusedAdam(m, lr=1e-3).step(d.x)


def test_lw_matches_fused_path_on_fitting_graphs():
    """Where both paths apply (norm=False, graphs within a tile): each is held to the fp64 oracle on its own, and the two
    fp32 implementations agree with each other to the per-tensor round-off bound of the parity tests."""
    o, o64, m = make_pair(110, 15, gain=1.7)
    _, _, mlw = make_pair(110, 15, gain=1.7, path="lw")
    b = data.collate(positions.load_positions("pos_11x11_seed1234.npz", maker=1)[:48])
    d = b.to(DEV)
    action, target, weight = loss_inputs(b)
    td_loss_of(o, b, action, target, weight, torch.float32).backward()
    td_loss_of(o64, b, action, target, weight, torch.float64).backward()
    A, T, W = action.to(DEV), target.to(DEV), weight.to(DEV)
    l1, _, q1 = m.train_step(d.x, d.edge_index, d.batch, d.ptr, action=A, target=T, weight=W)
    l2, _, q2 = mlw.train_step(d.x, d.edge_index, d.batch, d.ptr, action=A, target=T, weight=W)
    assert rel_err(q2, q1) <= REL_TOL and rel_err(l2, l1) <= REL_TOL
    check_grads(m, o, o64)
    check_grads(mlw, o, o64)
    assert rel_err(mlw.flat_gradient, m.flat_gradient) <= helpers.PER_TENSOR_TOL


def test_lw_argument_checks():
    _, _, m = make_pair(35, 10, norm=True)
    b = data.collate(positions.load_positions("pos_7x7_seed1234.npz", maker=1)[:4]).to(DEV)
    csr = m.build_csr(b.edge_index, b.batch, b.ptr, num_nodes=int(b.x.shape[0]))
    with pytest.raises(_lib.HexGnnError):   # norm block of the wrong size
        ops.lw_forward(m._desc, m.flat_parameters, torch.zeros(7, device=DEV), b.x, csr)
    x = b.x.clone().requires_grad_(True)
    q = m(x, b.edge_index, b.batch, b.ptr)
    with pytest.raises(RuntimeError):        # dL/dx exists on the fused path only: refused, not silently zero
        q.sum().backward()
    with pytest.raises(ValueError):
        ModernTwoHeaded(35, 10, norm=True, path="tc")


def _tiny_graph(n, edges, maker=1):
    """A hand-made graph (undirected edge list -> both directions); rows 0, 1 are the terminals."""
    e = torch.tensor(edges, dtype=torch.long).t().reshape(2, -1)
    ei = torch.cat([e, e.flip(0)], dim=1)
    x = torch.zeros(n, 3)
    x[:, 0] = torch.bincount(ei[1], minlength=n).float() if ei.numel() else 0
    x[:2, 1] = 1
    x[:, 2] = maker
    return data.HexGraph(x=x, edge_index=ei)


@pytest.mark.parametrize("norm", [False, True])
def test_lw_edge_cases_isolated_nodes_single_graph_tiny_batches(norm):
    """Empty neighbourhoods (mean of nothing = 0, PyG / torch_scatter semantics), a single 3-node graph, a batch far smaller
    than one 128-row block, node counts that are not multiples of anything: forward and gradients against the fp64 oracle."""
    o, o64, m = make_pair(35, 4, norm=norm, path="lw")
    graphs = [_tiny_graph(3, [(0, 2)]),                                   # node 1 isolated
              _tiny_graph(7, [(0, 2), (2, 3), (3, 1), (4, 5)]),           # node 6 isolated, two components
              _tiny_graph(5, []),                                          # no edges at all
              positions.load_positions("pos_7x7_seed1234.npz", maker=1)[0]]
    for gs in (graphs[:1], graphs):
        b = data.collate(gs)
        d = b.to(DEV)
        n = b.ptr[1:] - b.ptr[:-1]
        action = (n - 1).clamp(min=0)
        target = torch.linspace(-0.5, 0.5, b.num_graphs)
        weight = torch.ones(b.num_graphs)
        for mod in (o, o64):
            mod.zero_grad()
        td_loss_of(o, b, action, target, weight, torch.float32).backward()
        l64 = td_loss_of(o64, b, action, target, weight, torch.float64)
        l64.backward()
        m.zero_grad()
        loss, td_err, q = m.train_step(d.x, d.edge_index, d.batch, d.ptr, action=action.to(DEV), target=target.to(DEV),
                                       weight=weight.to(DEV))
        with torch.no_grad():
            q64 = o64(b.x.double(), b.edge_index, b.batch, b.ptr)
        assert int(m.last_csr.status.item()) == 0
        assert not torch.isnan(q).any() and rel_err(q, q64) <= REL_TOL and rel_err(loss, l64) <= REL_TOL
        check_grads(m, o, o64)


def test_learner_step_with_norm_model():
    """RainbowLearner over a norm=True pair (rainbow.py:259-338): the target sync carries the LayerNorm weights, a step runs on
    the layer-at-a-time path with torch.optim.Adam over the named parameters and changes conv AND norm weights."""
    from gnn_hex_b200 import RainbowLearner
    o, _, pol = make_pair(35, 4, norm=True)
    tgt = ModernTwoHeaded(35, 4, 2, norm=True).to(DEV)
    lrn = RainbowLearner(pol, tgt, lr=1e-3, adam_eps=1.5e-4, max_grad_norm=10.0, gamma=0.97, n_step=2)
    lrn.sync_Q_target()
    assert torch.equal(tgt.gnn.norms[2].weight, pol.gnn.norms[2].weight) and torch.equal(tgt.after_embed_norm.bias, pol.after_embed_norm.bias)
    gs = positions.load_positions("pos_7x7_seed1234.npz", maker=1)
    hs, hn = data.collate(gs[:32]), data.collate(gs[32:64])
    ds, dn = hs.to(DEV), hn.to(DEV)
    action, _, weight = loss_inputs(hs)
    reward = torch.zeros(32, device=DEV)
    done = torch.zeros(32, device=DEV)
    with torch.no_grad():
        q_before = pol(ds.x, ds.edge_index, ds.batch, ds.ptr).clone()
        t_before = tgt(ds.x, ds.edge_index, ds.batch, ds.ptr).clone()
    assert torch.equal(q_before, t_before)                       # identical replicas after the sync
    n0 = pol.gnn.norms[1].weight.detach().clone()
    out = lrn.train(ds, dn, action.to(DEV), reward, done, weight.to(DEV))
    assert torch.isfinite(out["loss"]) and torch.isfinite(out["grad_norm"])
    assert not torch.equal(pol.gnn.norms[1].weight.detach(), n0)
    with torch.no_grad():
        assert not torch.equal(pol(ds.x, ds.edge_index, ds.batch, ds.ptr), q_before)
        assert torch.equal(tgt(ds.x, ds.edge_index, ds.batch, ds.ptr), t_before)


def test_ragged_transfer_batch_full_size():
    """BASELINE configs[3] at full size: 256 graphs of 5x5 .. 13x13 boards (3 .. 171 nodes, 14,646 rows) through the default
    route for such batches (max_graph_nodes > 128 -> layer-at-a-time path): Q against the fp64 oracle; a permutation of the
    GRAPH order permutes Q (graphs are independent; blocks are cut differently, so equality is to round-off)."""
    o, o64, m = make_pair(110, 15, gain=1.7, max_graph_nodes=256)
    gs = positions.load_positions("pos_mix_5_13_seed1234.npz", maker=1)
    assert len(gs) == 256
    b = data.collate(gs)
    d = b.to(DEV)
    with torch.no_grad():
        q = m(d.x, d.edge_index, d.batch, d.ptr)
        q64 = o64(b.x.double(), b.edge_index, b.batch, b.ptr)
    assert int(m.last_csr.status.item()) == 0 and int((b.ptr[1:] - b.ptr[:-1]).max()) > 128
    assert rel_err(q, q64) <= REL_TOL
    perm = torch.randperm(256, generator=torch.Generator().manual_seed(9)).tolist()
    b2 = data.collate([gs[i] for i in perm])
    d2 = b2.to(DEV)
    with torch.no_grad():
        q2 = m(d2.x, d2.edge_index, d2.batch, d2.ptr).cpu()
    qc = q.cpu()
    for k in (0, 17, 101, 255):
        g = perm[k]
        a = qc[int(b.ptr[g]):int(b.ptr[g + 1])]
        c = q2[int(b2.ptr[k]):int(b2.ptr[k + 1])]
        assert float((a - c).abs().max()) <= REL_TOL * float(qc.abs().max())
