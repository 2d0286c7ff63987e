"""CPU: the MODEL half of the oracle pinned against the reference's OWN code.

Two layers of evidence:
  * golden vectors (tests/golden/model_*.npz) produced by `GN0/models.py` itself -- `get_pre_defined("modern_two_headed")`
    -> `DuellingTwoHeaded` / `cachify_gnn(GraphSAGE)` / `HeadNetwork` / `MLP` / `FactorizedNoisyLinear`, unmodified, run on the
    test-only PyG stand-in (oracle/pyg_stub) by tests/golden/make_model_fixtures.py.  These tests run anywhere (GPU box too);
  * the same classes executed LIVE wherever /root/reference is mounted (this container): forward / backward of the whole
    model, `MLP`, `MLP.grow_input_width`, `FactorizedNoisyLinear`, `HeadNetwork.grow_width`, `CachifiedGNN.grow_depth /
    grow_width`, compared with the oracle's restatements bit for bit under one seed.
What stays unpinned: only the third-party primitives the stand-in restates (PyG `SAGEConv.propagate` -- the reference's own
restatement is GN0/torch_script_models.py:52-73 --, `LayerNorm`, torch_scatter `scatter`); see oracle/pyg_stub/README.md."""
from argparse import Namespace

import numpy as np
import pytest
import torch

import helpers
from gnn_hex_b200 import data
import positions
from oracle import hexgnn_oracle as orc

CASES = {
    # tag: (hidden, layers, head layers, norm, noisy, positions) -- must equal tests/golden/make_model_fixtures.py
    "S": (35, 10, 2, False, False, "pos_7x7_seed1234.npz"),
    "L": (110, 15, 2, False, False, "pos_11x11_seed1234.npz"),
    "S_norm": (35, 10, 2, True, False, "pos_7x7_seed1234.npz"),
    "S_noisy": (35, 10, 2, False, True, "pos_7x7_seed1234.npz"),
}


def oracle_of(tag, fx, dtype=torch.float32):
    H, L, Lh, norm, noisy, _ = CASES[tag]
    o = orc.ModernTwoHeadedOracle(H, L, Lh, noisy_dqn=noisy, noisy_sigma0=0.5, norm=norm)
    helpers.set_hash_weights(o, seed=0, gain=float(fx["gain"]))
    return o.to(dtype)


def batch_of(tag, fx, maker):
    return data.collate(positions.load_positions(CASES[tag][5], maker=maker)[:int(fx["count"])])


def max_rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


@pytest.mark.parametrize("tag", list(CASES))
@pytest.mark.parametrize("maker", [1, 0])
def test_oracle_equals_reference_module_code(tag, maker):
    """Oracle forward (three output modes) == the reference module's committed outputs, bit for bit; loss and the whole
    gradient to fp32 round-off (both sides are fp32 autograd over the same formulas; the fp64 reference run arbitrates)."""
    fx = helpers.load_model_fixture(tag)
    b = batch_of(tag, fx, maker)
    action, target, weight = (torch.from_numpy(fx[k]) for k in ("action", "target", "weight"))
    o = oracle_of(tag, fx)
    assert sorted(dict(o.named_parameters())) == [str(n) for n in fx["names"]], "parameter names differ from the reference's"
    q = o(b.x, b.edge_index, b.batch, b.ptr)
    emb = o.final_conv_acts.detach()
    loss = orc.rainbow_loss(q, b.ptr, action, target, weight)
    loss.backward()
    with torch.no_grad():
        adv = o(b.x, b.edge_index, b.batch, b.ptr, advantages_only=True)
        v, c = o(b.x, b.edge_index, b.batch, b.ptr, seperate=True)
    assert np.array_equal(q.detach().numpy(), fx[f"q_{maker}"]), max_rel(q.detach().numpy(), fx[f"q_{maker}"])
    assert np.array_equal(adv.numpy(), fx[f"adv_{maker}"]) and adv.shape == (b.x.shape[0], 1)
    assert np.array_equal(v.numpy(), fx[f"value_{maker}"]) and np.array_equal(c.numpy(), fx[f"centred_{maker}"])
    assert np.array_equal(emb.numpy()[:: max(1, emb.shape[0] // 64)], fx[f"embeds_{maker}"])
    assert max_rel(loss.detach().numpy(), fx[f"loss_{maker}"]) <= 1e-6
    s = int(fx["gstride"])
    g = helpers.flat_named(o, "grad", [str(n) for n in fx["names"]])[::s]
    gmax = float(fx[f"gradmax_{maker}"])
    err32 = float(np.abs(g - fx[f"grad_{maker}"]).max()) / gmax
    err64 = float(np.abs(g - fx[f"grad64_{maker}"]).max()) / gmax
    ref_roundoff = float(np.abs(fx[f"grad_{maker}"].astype(np.float64) - fx[f"grad64_{maker}"]).max()) / gmax
    assert err32 <= 2e-6, (err32, ref_roundoff)
    assert err64 <= max(2e-6, 4 * ref_roundoff), (err64, ref_roundoff)
    # the head that was not taken gets no gradient at all, in the reference (None) and here
    other = "breaker_head." if maker == 1 else "maker_head."
    for n, p in o.named_parameters():
        if n.startswith(other):
            assert p.grad is None or not p.grad.any(), n


def test_fp64_oracle_equals_fp64_reference():
    """The fp64 oracle (the arbiter of every GPU gradient tolerance) == the reference module run in fp64."""
    for tag in ("S", "S_norm"):
        fx = helpers.load_model_fixture(tag)
        b = batch_of(tag, fx, 1)
        o = oracle_of(tag, fx, torch.float64)
        q = o(b.x.double(), b.edge_index, b.batch, b.ptr)
        assert max_rel(q.detach().numpy(), fx["q64_1"]) <= 1e-13
        loss = orc.rainbow_loss(q, b.ptr, torch.from_numpy(fx["action"]), torch.from_numpy(fx["target"]).double(),
                                torch.from_numpy(fx["weight"]).double())
        loss.backward()
        g = helpers.flat_named(o, "grad", [str(n) for n in fx["names"]])[:: int(fx["gstride"])]
        assert float(np.abs(g - fx["grad64_1"]).max()) / float(fx["gradmax_1"]) <= 1e-12


# ---- live: the reference classes themselves (this container only) ----------------------------------------------------
ref = helpers.import_reference_models()
needs_reference = pytest.mark.skipif(ref is None, reason="/root/reference is not mounted (GPU box): golden vectors cover it")


def ref_args(H, L, Lh=2, norm=False, noisy=False):
    return Namespace(hidden_channels=H, num_layers=L, num_head_layers=Lh, norm=norm, noisy_dqn=noisy, noisy_sigma0=0.5)


def assert_same_state(a, b):
    sa, sb = a.state_dict(), b.state_dict()
    assert sorted(sa) == sorted(sb)
    for k in sa:
        assert sa[k].shape == sb[k].shape and torch.equal(sa[k], sb[k]), k


@needs_reference
def test_fixtures_are_what_the_reference_code_produces_now():
    """Regenerating one case in place gives the committed bits (the fixture is not stale, the stand-in has not drifted)."""
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location("make_model_fixtures", os.path.join(helpers.GOLDEN_DIR, "make_model_fixtures.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    out = gen.run_case(ref, "S_noisy")
    fx = helpers.load_model_fixture("S_noisy")
    for k in ("q_1", "q_0", "adv_0", "grad_1", "grad64_0", "value_1"):
        assert np.array_equal(out[k], fx[k]), k


@needs_reference
@pytest.mark.parametrize("norm,noisy", [(False, False), (True, False), (False, True)])
def test_live_reference_model_equals_oracle(norm, noisy):
    """A small model, both heads, single graph without batch vector (models.py:542-543) and a batch: bit-equal outputs."""
    m = ref.get_pre_defined("modern_two_headed", ref_args(12, 3, 2, norm, noisy))
    o = orc.ModernTwoHeadedOracle(12, 3, 2, noisy_dqn=noisy, noisy_sigma0=0.5, norm=norm)
    helpers.set_hash_weights(m, 3)
    helpers.set_hash_weights(o, 3)
    assert_same_state(m, o)
    gs = positions.load_positions("pos_7x7_seed1234.npz", maker=1)[:5]
    b = data.collate(gs)
    for maker in (1.0, 0.0):
        x = b.x.clone()
        x[:, 2] = maker
        assert torch.equal(m(x, b.edge_index, b.batch, b.ptr), o(x, b.edge_index, b.batch, b.ptr))
        assert torch.equal(m(x, b.edge_index, b.batch, advantages_only=True), o(x, b.edge_index, b.batch, advantages_only=True))
        vr, cr = m(x, b.edge_index, b.batch, seperate=True)
        vo, co = o(x, b.edge_index, b.batch, seperate=True)
        assert torch.equal(vr, vo) and torch.equal(cr, co)
        g = gs[2]
        xg = g.x.clone()
        xg[:, 2] = maker
        assert torch.equal(m(xg, g.edge_index), o(xg, g.edge_index))
        assert torch.equal(m.final_conv_acts, o.final_conv_acts)


@needs_reference
def test_live_mlp_and_grow_input_width():
    """`MLP` (models.py:36-50,75-82) and `MLP.grow_input_width` (:52-73) vs the oracle's `_MLP`, one seed, bit for bit."""
    torch.manual_seed(7)
    r = ref.MLP(5, 1, 40, 1)
    torch.manual_seed(7)
    o = orc._MLP(5, 40, 1)
    assert_same_state(r, o)
    x = torch.randn(6, 40)
    assert torch.equal(r(x), o(x))
    torch.manual_seed(8)
    r.grow_input_width(56, 7)
    torch.manual_seed(8)
    o.grow_input_width(56, 7)
    assert_same_state(r, o)
    x = torch.randn(6, 56)
    assert torch.equal(r(x), o(x))


@needs_reference
def test_live_factorized_noisy_linear():
    """`FactorizedNoisyLinear` (models.py:84-141): init, reset_noise, disable_noise and forward vs the oracle AND the
    product module's holder, same seed -> same bits."""
    from gnn_hex_b200.models import FactorizedNoisyLinear as Holder
    mods = []
    for cls in (ref.FactorizedNoisyLinear, orc.NoisyLinearOracle, Holder):
        torch.manual_seed(4)
        mods.append(cls(9, 1, 0.5))
    for other in mods[1:]:
        assert_same_state(mods[0], other)
    for i, l in enumerate(mods):
        torch.manual_seed(11)
        l.reset_noise()
    for other in mods[1:]:
        assert_same_state(mods[0], other)
    x = torch.randn(4, 9)
    assert torch.equal(mods[0](x), mods[1](x))
    for l in mods:
        l.disable_noise()
    assert not mods[2].weight_epsilon.any() and not mods[2].bias_epsilon.any()
    assert torch.equal(mods[0](x), mods[1](x)) and torch.equal(mods[0](x), x @ mods[0].weight_mu.t() + mods[0].bias_mu)


@needs_reference
def test_live_growth_of_the_whole_model():
    """`DuellingTwoHeaded.grow_depth / grow_width` (models.py:494-508) -> `CachifiedGNN.grow_depth / grow_width` (:166-238),
    `HeadNetwork.grow_width` (:336-358): reference vs oracle vs the product module's parameter holder under one seed --
    identical state_dicts, and the reference's function is what the oracle computes afterwards."""
    from gnn_hex_b200 import ModernTwoHeaded
    m = ref.get_pre_defined("modern_two_headed", ref_args(10, 3, 2))
    o = orc.ModernTwoHeadedOracle(10, 3, 2)
    p = ModernTwoHeaded(10, 3, 2)
    for mod in (m, o, p):
        helpers.set_hash_weights(mod, 5)
    b = data.collate(positions.load_positions("pos_7x7_seed1234.npz", maker=1)[:3])
    q0 = m(b.x, b.edge_index, b.batch, b.ptr)
    for mod in (m, o, p):
        torch.manual_seed(21)
        mod.grow_depth(2)
    assert_same_state(m, o)
    assert_same_state(m, p)
    q1 = m(b.x, b.edge_index, b.batch, b.ptr)
    assert torch.allclose(q1, q0, atol=1e-6)              # identity layers: the function does not change
    assert torch.equal(q1, o(b.x, b.edge_index, b.batch, b.ptr))
    for mod in (m, o, p):
        torch.manual_seed(22)
        mod.grow_width(14)
        mod.maker_head.grow_depth(1)
        mod.breaker_head.grow_depth(1)
    assert_same_state(m, o)
    assert_same_state(m, p)
    assert torch.equal(m(b.x, b.edge_index, b.batch, b.ptr), o(b.x, b.edge_index, b.batch, b.ptr))
    assert m.gnn.convs[1].lin_l.weight.shape == (14, 14) and m.maker_head.value_head.layers[0].weight.shape == (7, 56)


@needs_reference
def test_live_head_network_alone():
    """`HeadNetwork.forward` (models.py:368-384): pooling order (sum, max, min, mean) and the MLP input layout."""
    torch.manual_seed(0)
    m = ref.get_pre_defined("modern_two_headed", ref_args(8, 2, 2))
    o = orc.ModernTwoHeadedOracle(8, 2, 2)
    helpers.set_hash_weights(m, 9)
    helpers.set_hash_weights(o, 9)
    b = data.collate(positions.load_positions("pos_7x7_seed1234.npz", maker=1)[:4])
    emb = torch.rand(b.x.shape[0], 8)
    ar, vr = m.maker_head(emb, b.edge_index, b.batch)
    ao, vo = o.maker_head(emb, b.edge_index, b.batch, 4)
    assert torch.equal(ar, ao) and torch.equal(vr, vo)
