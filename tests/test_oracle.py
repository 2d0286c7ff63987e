"""CPU: pin the oracle (oracle/hexgnn_oracle.py) -- golden fixtures from the reference's C++ game,
a hand-computed known answer, fp64 gradcheck and the structural invariants of the model."""
import numpy as np
import pytest
import torch

from gnn_hex_b200 import data
import positions
from oracle import hexgnn_oracle as orc


def test_empty_board_matches_reference_cpp_game():
    """hex_board_undirected_edges (restating hex_board_game.py:214-232) == the compiled C++ game."""
    z = np.load(positions.GOLDEN_DIR + "/empty_boards_3_13.npz")
    eo = 0
    for g, size in enumerate(z["size"]):
        n, ne = int(z["n"][g]), int(z["ne"][g])
        ref = set(zip(z["src"][eo:eo + ne].tolist(), z["dst"][eo:eo + ne].tolist()))
        eo += ne
        mine = orc.hex_board_undirected_edges(int(size))
        sym = set(mine) | {(b, a) for a, b in mine}
        assert n == size * size + 2
        assert sym == ref, f"size {size}"
    # sizes pinned by construction: 11x11 -> 123/644, 7x7 -> 51/244
    assert 2 * len(orc.hex_board_undirected_edges(11)) == 644
    assert 2 * len(orc.hex_board_undirected_edges(7)) == 244


def test_fixture_graphs_are_valid_and_symmetric():
    for name in ("pos_11x11_seed1234.npz", "pos_7x7_seed1234.npz", "pos_mix_5_13_seed1234.npz"):
        gs = positions.load_positions(name)
        assert len(gs) == 256
        for g in gs[:40]:
            ei = g.edge_index
            n = g.x.shape[0]
            assert ei.min() >= 0 and ei.max() < n
            fwd = set(zip(ei[0].tolist(), ei[1].tolist()))
            assert len(fwd) == ei.shape[1]                       # no duplicates
            assert all((b, a) in fwd for a, b in fwd)            # symmetric
            assert all(a != b for a, b in fwd)                   # no self loops
            assert torch.equal(g.x[:, 0], torch.bincount(ei[1], minlength=n).float())
            assert g.x[0, 1] == 1 and g.x[1, 1] == 1 and g.x[2:, 1].sum() == 0


def test_collate_matches_cpp_collate_batch_semantics():
    gs = positions.load_positions("pos_7x7_seed1234.npz")[:5]
    b = data.collate(gs)
    o = orc.collate(gs)
    for k in ("x", "edge_index", "batch", "ptr"):
        assert torch.equal(getattr(b, k), getattr(o, k)), k
    # util.cpp:22-41: graph i's edges are shifted by the node offset, batch[i-range] = i
    off = 0
    for i, g in enumerate(gs):
        n = g.x.shape[0]
        assert torch.all(b.batch[off:off + n] == i)
        off += n
    assert b.ptr[-1] == off


def test_known_answer_debugging_manager_graph():
    """Hand-computed forward on the reference's 4-node fixture (multi_env_manager.py:170-175):
    x = [[0,0,1]]*3 + [[1,1,1]], one directed edge 2 -> 3, H=2, one body layer, one head layer."""
    m = orc.ModernTwoHeadedOracle(2, 1, 1).double()
    with torch.no_grad():
        for p in m.parameters():
            p.zero_()
        c = m.gnn.convs[0]
        c.lin_l.weight.copy_(torch.tensor([[1., 0.], [0., 1.]]))   # mean of neighbours
        c.lin_r.weight.copy_(torch.tensor([[1., 0.], [0., 2.]]))   # root
        c.lin_l.bias.copy_(torch.tensor([0.5, -0.5]))
        h = m.maker_head
        hc = h.gnn.convs[0]
        hc.lin_l.weight.copy_(torch.eye(2)); hc.lin_r.weight.copy_(torch.eye(2))
        h.linear.weight.copy_(torch.tensor([[1., -1.]])); h.linear.bias.fill_(0.25)
        h.value_head.layers[0].weight.copy_(torch.tensor([[1., 0, 0, 0, 0, 0, 0, 0]]))  # sum of channel 0
        h.value_head.layers[1].weight.fill_(0.5)
    x = torch.tensor([[0, 0, 1], [0, 0, 1], [0, 0, 1], [1, 1, 1]], dtype=torch.float64)
    ei = torch.tensor([[2], [3]])
    q = m(x, ei)
    # body: rows 0..2: x=(0,0), no in-edges -> relu(b) = (0.5, 0); row 3: mean = x_2 = (0,0), root (1,2)+b -> (1.5, 1.5)
    # head layer: rows 0,1,2 -> (0.5,0) (+ nothing); row 3: mean = h_2 = (0.5,0) + own (1.5,1.5) -> (2.0,1.5)
    g = torch.tensor([[0.5, 0.], [0.5, 0.], [0.5, 0.], [2.0, 1.5]], dtype=torch.float64)
    a = 2 * torch.tanh(g[:, 0] - g[:, 1] + 0.25)
    v = torch.tanh(0.5 * torch.relu(g[:, 0].sum()))
    expect = v + a - a.mean()
    assert torch.allclose(q, expect, atol=1e-12), (q, expect)
    assert torch.allclose(m.final_conv_acts, torch.tensor([[0.5, 0], [0.5, 0], [0.5, 0], [1.5, 1.5]], dtype=torch.float64))
    # breaker flag takes the (all-zero) breaker head: Q = 0
    x2 = x.clone(); x2[:, 2] = 0
    assert torch.allclose(m(x2, ei), torch.zeros(4, dtype=torch.float64))


def _small_batch(k=3):
    gs = positions.load_positions("pos_7x7_seed1234.npz", maker=1)[:k]
    return gs, data.collate(gs)


def test_batched_equals_per_graph_and_output_modes():
    torch.manual_seed(0)
    m = orc.ModernTwoHeadedOracle(12, 3, 2).double()
    gs, b = _small_batch(4)
    x = b.x.double()
    q = m(x, b.edge_index, b.batch, b.ptr)
    single = torch.cat([m(g.x.double(), g.edge_index) for g in gs])
    assert torch.allclose(q, single, atol=1e-12)
    adv = m(x, b.edge_index, b.batch, advantages_only=True)
    assert adv.shape == (x.shape[0], 1) and adv.abs().max() <= 2
    v, c = m(x, b.edge_index, b.batch, seperate=True)
    assert v.shape == (4,) and torch.allclose(v[b.batch] + c, q, atol=1e-12)


def test_permutation_equivariance_and_edge_order_invariance():
    torch.manual_seed(0)
    m = orc.ModernTwoHeadedOracle(10, 3, 2).double()
    g = positions.load_positions("pos_7x7_seed1234.npz", maker=0)[7]
    x, ei = g.x.double(), g.edge_index
    q = m(x, ei)
    gen = torch.Generator().manual_seed(3)
    n = x.shape[0]
    perm = torch.cat([torch.arange(2), 2 + torch.randperm(n - 2, generator=gen)])  # terminals stay rows 0,1
    inv = torch.empty_like(perm); inv[perm] = torch.arange(n)
    qp = m(x[perm], inv[ei])
    assert torch.allclose(qp, q[perm], atol=1e-11)
    eperm = torch.randperm(ei.shape[1], generator=gen)
    assert torch.allclose(m(x, ei[:, eperm]), q, atol=1e-11)


def test_gradcheck_fp64():
    torch.manual_seed(0)
    m = orc.ModernTwoHeadedOracle(6, 2, 2).double()
    gs, b = _small_batch(2)
    x = b.x.double()
    params = list(m.maker_head.parameters()) + list(m.gnn.parameters())
    names = [n for n, _ in list(m.maker_head.named_parameters()) + list(m.gnn.named_parameters())]

    def f(*ps):
        with torch.no_grad():
            for p, v in zip(params, ps):
                p.copy_(v)
        sd = {("maker_head." + n if i < len(list(m.maker_head.parameters())) else "gnn." + n): v
              for i, (n, v) in enumerate(zip(names, ps))}
        return torch.func.functional_call(m, sd, (x, b.edge_index, b.batch, b.ptr))

    ins = [p.detach().clone().requires_grad_(True) for p in params]
    assert torch.autograd.gradcheck(f, ins, eps=1e-6, atol=1e-5, nondet_tol=0)


def test_csr_by_destination_is_stable():
    ei = torch.tensor([[3, 1, 2, 0, 2, 1], [0, 0, 1, 1, 0, 2]])
    rowptr, col = orc.csr_by_destination(ei, 4)
    assert rowptr.tolist() == [0, 3, 5, 6, 6]
    assert col.tolist() == [3, 1, 2, 2, 0, 1]


def test_flops_formula_matches_survey():
    assert orc.flops_forward(1, 0, 110, 15, advantages_only=True) == 775500
    assert orc.flops_forward(1, 0, 35, 10, advantages_only=True) == 54250


def test_noisy_linear_known_answer():
    """FactorizedNoisyLinear (GN0/models.py:84-141): init constants, factorised noise, w = mu + sigma * eps."""
    torch.manual_seed(4)
    l = orc.NoisyLinearOracle(9, 1, 0.5)
    scale = 1 / 3.0
    assert float(l.weight_mu.abs().max()) <= scale and torch.all(l.weight_sigma == 0.5 * scale) and torch.all(l.bias_sigma == 0.5 * scale)
    # factorised: weight_epsilon = eps_out (x) eps_in with eps_out = bias_epsilon, and |f(x)| = sqrt|x|
    eps_in = l.weight_epsilon[0] / l.bias_epsilon[0]
    torch.manual_seed(11)
    l.reset_noise()
    torch.manual_seed(11)
    n_in, n_out = torch.randn(9), torch.randn(1)
    f = lambda v: v.sign() * v.abs().sqrt()
    assert torch.equal(l.bias_epsilon, f(n_out)) and torch.equal(l.weight_epsilon, f(n_out).outer(f(n_in)))
    x = torch.arange(18, dtype=torch.float32).reshape(2, 9) / 7
    want = x @ (l.weight_mu + l.weight_sigma * l.weight_epsilon).t() + (l.bias_mu + l.bias_sigma * l.bias_epsilon)
    assert torch.allclose(l(x), want, atol=1e-7)
    l.disable_noise()
    assert torch.allclose(l(x), x @ l.weight_mu.t() + l.bias_mu, atol=1e-7)
    del eps_in


def test_noisy_model_names_and_product_holder():
    """`noisy_dqn=True` swaps only the advantage Linear (models.py:331-334); names interchange with the product module."""
    from gnn_hex_b200 import ModernTwoHeaded, get_pre_defined
    from types import SimpleNamespace
    torch.manual_seed(0)
    o = orc.ModernTwoHeadedOracle(12, 3, 2, noisy_dqn=True, noisy_sigma0=0.5)
    m = get_pre_defined("modern_two_headed", SimpleNamespace(hidden_channels=12, num_layers=3, num_head_layers=2, norm=False,
                                                             noisy_dqn=True, noisy_sigma0=0.5))
    assert isinstance(m, ModernTwoHeaded) and m.noisy_dqn
    assert list(o.state_dict().keys()) == list(m.state_dict().keys())
    assert "maker_head.linear.weight_mu" in m.state_dict() and "breaker_head.linear.bias_epsilon" in m.state_dict()
    assert "maker_head.linear.weight" not in m.state_dict()
    m.load_state_dict(o.state_dict())
    lin = m.maker_head.linear
    torch.manual_seed(5)
    lin.reset_noise()
    torch.manual_seed(5)
    o.maker_head.linear.reset_noise()
    assert torch.equal(lin.weight_epsilon, o.maker_head.linear.weight_epsilon) and torch.equal(lin.bias_epsilon, o.maker_head.linear.bias_epsilon)
    lin.disable_noise()
    assert not lin.weight_epsilon.any() and not lin.bias_epsilon.any()


def test_norm_model_names_and_product_holder():
    """`norm=True` adds PyG LayerNorm parameters (models.py:935,945: one per layer incl. the last, :158-162; plus
    after_embed_norm, :480-483); names and order interchange with the oracle (and so with the reference: test_reference_pin)."""
    from gnn_hex_b200 import ModernTwoHeaded, get_pre_defined
    from types import SimpleNamespace
    o = orc.ModernTwoHeadedOracle(12, 3, 2, norm=True)
    m = get_pre_defined("modern_two_headed", SimpleNamespace(hidden_channels=12, num_layers=3, num_head_layers=2, norm=True,
                                                             noisy_dqn=False))
    assert isinstance(m, ModernTwoHeaded) and m.norm and m._uses_lw()
    assert sorted(o.state_dict().keys()) == sorted(m.state_dict().keys())
    for k in ("gnn.norms.2.weight", "after_embed_norm.bias", "breaker_head.gnn.norms.1.weight"):
        assert k in m.state_dict()
    m.load_state_dict(o.state_dict())
    # the order of the C-ABI's norm block: body layers, after_embed_norm, maker head, breaker head; weight then bias
    ps = m._ordered_norm_params()
    assert len(ps) == 2 * (3 + 1 + 2 + 2) and ps[0] is m.gnn.norms[0].weight and ps[7] is m.after_embed_norm.bias
    assert ps[8] is m.maker_head.gnn.norms[0].weight and ps[-1] is m.breaker_head.gnn.norms[1].bias
    with pytest.raises(NotImplementedError):
        m.grow_depth(1)
    with pytest.raises(ValueError):
        ModernTwoHeaded(12, 3, 2, norm=True, path="simt")
    assert not ModernTwoHeaded(12, 3, 2)._uses_lw() and ModernTwoHeaded(12, 3, 2, max_graph_nodes=363)._uses_lw()


def test_grow_depth_keeps_the_function_and_grow_width_layout():
    """grow_depth appends identity layers (models.py:166-185): same Q-values; grow_width (models.py:187-238,336-358,52-73):
    old weights in the top-left block, zeros right of them, product holder == oracle bit for bit under the same seed."""
    from gnn_hex_b200 import ModernTwoHeaded
    torch.manual_seed(0)
    o = orc.ModernTwoHeadedOracle(10, 3, 2)
    m = ModernTwoHeaded(10, 3, 2)
    m.load_state_dict(o.state_dict())
    gs, b = _small_batch(3)
    with torch.no_grad():
        q0 = o(b.x, b.edge_index, b.batch, b.ptr)
    torch.manual_seed(21)
    o.grow_depth(2)
    torch.manual_seed(21)
    m.grow_depth(2)
    assert o.gnn.num_layers == 5 and len(m.gnn.convs) == 5
    with torch.no_grad():
        assert torch.allclose(o(b.x, b.edge_index, b.batch, b.ptr), q0, atol=1e-6)
    w_old = o.gnn.convs[1].lin_l.weight.clone()
    v_old = o.maker_head.value_head.layers[0].weight.clone()
    torch.manual_seed(22)
    o.grow_width(14)
    o.maker_head.grow_depth(1); o.breaker_head.grow_depth(1)
    torch.manual_seed(22)
    m.grow_width(14)
    m.maker_head.grow_depth(1); m.breaker_head.grow_depth(1)
    sd_o, sd_m = o.state_dict(), m.state_dict()
    assert list(sd_o.keys()) == list(sd_m.keys())
    for k in sd_o:
        assert torch.equal(sd_o[k], sd_m[k]), k
    w = o.gnn.convs[1].lin_l.weight
    assert w.shape == (14, 14) and torch.equal(w[:10, :10], w_old) and not w[:10, 10:].any() and w[10:, :].any()
    v = o.maker_head.value_head.layers[0].weight
    assert v.shape == (7, 56) and torch.equal(v[:5, :40], v_old) and not v[5:, :].any() and not v[:, 40:].any()
    assert o.maker_head.linear.weight.shape == (1, 14) and not o.maker_head.linear.weight[:, 10:].any()
    m._ensure_flat()
    assert (m.hidden_channels, m.num_layers, m.num_head_layers) == (14, 5, 3)
    assert m.flat_parameters.numel() == sum(p.numel() for p in o.parameters())


def test_oracle_against_real_pyg_when_installed():
    """torch_geometric 2.2.0 / torch_scatter 2.1.0 are absent from this image (no network), so the model
    half of the oracle is pinned by known answers only.  Wherever they ARE importable this test pins the two third-party
    semantics the oracle restates: SAGEConv (mean aggregation, lin_l with bias on the neighbours, lin_r without bias on the
    root; message flow edge_index[0] -> edge_index[1]) and scatter(sum / max / min / mean) with empty groups = 0."""
    tg = pytest.importorskip("torch_geometric")
    ts = pytest.importorskip("torch_scatter")
    from torch_geometric.nn import SAGEConv
    torch.manual_seed(0)
    g = positions.load_positions("pos_7x7_seed1234.npz", maker=1)[3]
    x = torch.randn(g.x.shape[0], 6)
    conv = SAGEConv(6, 5)
    want = conv(x, g.edge_index)
    got = orc.sage_conv(x, g.edge_index, conv.lin_l.weight, conv.lin_l.bias, conv.lin_r.weight)
    assert torch.allclose(got, want, atol=1e-6)
    idx = torch.tensor([0, 0, 2, 2, 2, 4])
    src = torch.randn(6, 3)
    for red in ("sum", "max", "min", "mean"):
        assert torch.allclose(orc.scatter(src, idx, 5, red), ts.scatter(src, idx, dim=0, dim_size=5, reduce=red), atol=1e-6), red
    del tg
