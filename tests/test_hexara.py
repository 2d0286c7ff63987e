"""CPU: the HexAra network.  The oracle's restatement of `SAGE_torch_script` (GN0/torch_script_models.py:286-379)
against the golden vectors the reference's OWN module produced (tests/golden/model_ara_*.npz, make_hexara_fixtures.py) and --
where /root/reference is mounted -- against the live reference class; the product holder's names and parameter table."""
import ctypes as C

import numpy as np
import pytest
import torch

import helpers
import positions
from gnn_hex_b200 import SAGETorchScript, _lib, data, get_current_model
from oracle import hexgnn_oracle as orc

CASES = {
    "ara_S": (28, 5, 2, 2, "pos_7x7_seed1234.npz"),
    "ara_L": (60, 15, 2, 2, "pos_11x11_seed1234.npz"),
    "ara_P1": (44, 3, 1, 3, "pos_7x7_seed1234.npz"),
}


@pytest.mark.parametrize("tag", list(CASES))
@pytest.mark.parametrize("swap", [0, 1])
def test_oracle_equals_reference_golden_vectors(tag, swap):
    H, L, Lp, Lv, pos = CASES[tag]
    fx = helpers.load_model_fixture(tag)
    b = data.collate(positions.load_positions_cpp(pos)[:int(fx["count"])])
    o = orc.HexAraOracle(H, L, Lp, Lv, swap_allowed=bool(swap))
    assert list(o.state_dict().keys()) == [str(n) for n in fx["names"]]
    helpers.set_hash_weights(o, seed=0, gain=float(fx["gain"]))
    with torch.no_grad():
        pi, value, gi, ptr = o(b.x, b.edge_index, b.batch, b.ptr)
    assert np.array_equal(gi.numpy(), fx[f"gi_{swap}"]) and np.array_equal(ptr.numpy(), fx[f"ptr_{swap}"])
    assert np.array_equal(pi.numpy(), fx[f"pi_{swap}"]) and np.array_equal(value.numpy(), fx[f"value_{swap}"])   # bit for bit
    # every graph's policy is a distribution; with swap the flagged graphs carry one more entry
    p = torch.zeros(b.num_graphs, dtype=torch.float64).index_add_(0, gi, pi.double().exp())
    assert float((p - 1).abs().max()) <= 1e-5
    n_out = ptr[1:] - ptr[:-1]
    flagged = (b.x[b.ptr[1:] - 1, 2] != 0).long() * swap
    assert torch.equal(n_out, (b.ptr[1:] - b.ptr[:-1]) - 2 + flagged) and int(flagged.sum()) == (len(flagged) + 2) // 3 * swap


@pytest.mark.parametrize("swap", [False, True])
def test_oracle_equals_live_reference_class(swap):
    ref = helpers.import_reference_torch_script_models()
    if ref is None:
        pytest.skip("/root/reference not mounted (GPU box): the committed golden vectors stand in")
    m = ref.get_current_model("SAGE", hidden_channels=20, hidden_layers=4, policy_layers=3, value_layers=1, swap_allowed=swap)
    o = orc.HexAraOracle(20, 4, 3, 1, swap_allowed=swap)
    assert list(m.state_dict().keys()) == list(o.state_dict().keys())
    helpers.set_hash_weights(m, seed=4, gain=1.5)
    o.load_state_dict(m.state_dict())
    b = data.collate(positions.load_positions_cpp("pos_7x7_seed1234.npz")[5:19])
    r = m(b.x, b.edge_index, b.batch, b.ptr)
    with torch.no_grad():
        q = o(b.x, b.edge_index, b.batch, b.ptr)
    for a, c in zip(r, q):
        assert torch.equal(a.detach(), c)


def test_product_holder_names_and_parameter_table(lib_built):
    m = get_current_model("SAGE", hidden_channels=28, hidden_layers=5, policy_layers=2, value_layers=2, swap_allowed=True)
    assert isinstance(m, SAGETorchScript)
    fx = helpers.load_model_fixture("ara_S")
    assert list(m.state_dict().keys()) == [str(n) for n in fx["names"]]
    assert tuple(m.my_modules["policy_head"].convs[1].lin_l.weight.shape) == (1, 28)
    assert tuple(m.gnn.convs[0].lin_r.weight.shape) == (28, 3)
    l = _lib.lib()
    assert l.hexgnn_hexara_param_count(C.byref(m._desc)) == sum(p.numel() for p in m.parameters())
    n = l.hexgnn_hexara_param_offsets(C.byref(m._desc), None, 0)
    offs = (C.c_int64 * n)()
    l.hexgnn_hexara_param_offsets(C.byref(m._desc), offs, n)
    ps = m._ordered_params()
    assert n == len(ps) and list(offs) == list(np.cumsum([0] + [p.numel() for p in ps[:-1]]))
    assert [id(p) for p in ps] == [id(p) for p in m.parameters()]          # C-ABI order == state_dict order
    bad = _lib.HexAraDesc(28, 5, 2, 2, 2, 0)
    assert l.hexgnn_hexara_param_count(C.byref(bad)) == -1
    with pytest.raises(NotImplementedError):
        get_current_model("PNA")
    with pytest.raises(_lib.HexGnnError):      # no CPU fallback: CPU tensors are refused
        m(torch.zeros(4, 3), torch.zeros(2, 0, dtype=torch.long), torch.zeros(4, dtype=torch.long), torch.tensor([0, 4]))
