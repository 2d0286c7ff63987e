"""GPU: hexgnn_hexara_forward (HexAra `SAGE_torch_script`, GN0/torch_script_models.py:286-379) through the module API against
the reference module's golden vectors and, at larger sizes, against the fp64 oracle."""
import numpy as np
import pytest
import torch

import helpers
import positions
from helpers import REL_TOL, rel_err
from gnn_hex_b200 import _lib, data, get_current_model
from oracle import hexgnn_oracle as orc

pytestmark = pytest.mark.gpu
DEV = "cuda"
CASES = {
    "ara_S": (28, 5, 2, 2, "pos_7x7_seed1234.npz"),
    "ara_L": (60, 15, 2, 2, "pos_11x11_seed1234.npz"),
    "ara_P1": (44, 3, 1, 3, "pos_7x7_seed1234.npz"),
}


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


@pytest.mark.parametrize("tag", list(CASES))
@pytest.mark.parametrize("swap", [0, 1])
def test_cuda_path_equals_reference_golden_vectors(tag, swap):
    H, L, Lp, Lv, pos = CASES[tag]
    fx = helpers.load_model_fixture(tag)
    b = data.collate(positions.load_positions_cpp(pos)[:int(fx["count"])])
    m = get_current_model("SAGE", hidden_channels=H, hidden_layers=L, policy_layers=Lp, value_layers=Lv, swap_allowed=bool(swap))
    helpers.set_hash_weights(m, seed=0, gain=float(fx["gain"]))
    m = m.to(DEV)
    m.capture_embeds = True
    d = b.to(DEV)
    with torch.no_grad():
        pi, value, gi, ptr = m(d.x, d.edge_index, d.batch, d.ptr)
    torch.cuda.synchronize()
    assert int(m.last_csr.status.item()) == 0
    assert np.array_equal(gi.cpu().numpy(), fx[f"gi_{swap}"]) and np.array_equal(ptr.cpu().numpy(), fx[f"ptr_{swap}"])
    assert pi.shape == fx[f"pi_{swap}"].shape and value.shape == (b.num_graphs,)
    assert rel(pi.cpu().numpy(), fx[f"pi64_{swap}"]) <= REL_TOL and rel(pi.cpu().numpy(), fx[f"pi_{swap}"]) <= REL_TOL
    assert float(np.abs(value.cpu().numpy() - fx[f"value64_{swap}"]).max()) <= REL_TOL
    emb = m.final_conv_acts.cpu().numpy()
    assert rel(emb[:: max(1, emb.shape[0] // 64)], fx[f"embeds_{swap}"]) <= REL_TOL


@pytest.mark.parametrize("swap", [False, True])
def test_default_model_full_batch_and_large_boards(swap):
    """get_current_model's defaults (hidden 60, 15 + 2 + 2 convs) on 256 11x11 positions plus 14x14 .. 19x19 boards (the
    kernels have no graph-size limit) against the fp64 oracle; repeated calls give the same bits."""
    o = orc.HexAraOracle(60, 15, 2, 2, swap_allowed=swap)
    helpers.set_hash_weights(o, seed=2, gain=1.7)
    o64 = orc.HexAraOracle(60, 15, 2, 2, swap_allowed=swap).double()
    o64.load_state_dict({k: v.double() for k, v in o.state_dict().items()})
    m = get_current_model("SAGE", swap_allowed=swap)
    m.load_state_dict(o.state_dict())
    m = m.to(DEV)
    gs = positions.load_positions_cpp("pos_11x11_seed1234.npz") + positions.load_positions_cpp("pos_transfer_14_19_seed1234.npz")
    b = data.collate(gs)
    d = b.to(DEV)
    with torch.no_grad():
        pi, value, gi, ptr = m(d.x, d.edge_index, d.batch, d.ptr)
        pi2, value2, _, _ = m(d.x, d.edge_index, d.batch, d.ptr)
        r = o64(b.x.double(), b.edge_index, b.batch, b.ptr)
    assert torch.equal(pi, pi2) and torch.equal(value, value2)
    assert torch.equal(gi.cpu(), r[2]) and torch.equal(ptr.cpu(), r[3])
    assert rel_err(pi, r[0]) <= REL_TOL
    assert float((value.cpu().double() - r[1]).abs().max()) <= REL_TOL
    # arg-max move per graph agrees with the oracle wherever the oracle's top two are not tied to round-off
    pc, rc = pi.cpu().double(), r[0]
    for g in range(0, b.num_graphs, 17):
        s, e = int(r[3][g]), int(r[3][g + 1])
        top = torch.topk(rc[s:e], 2).values
        if float(top[0] - top[1]) > 1e-4:
            assert int(pc[s:e].argmax()) == int(rc[s:e].argmax())


def test_hexara_argument_checks():
    m = get_current_model("SAGE", hidden_channels=28, hidden_layers=3).to(DEV)
    b = data.collate(positions.load_positions_cpp("pos_7x7_seed1234.npz")[:4]).to(DEV)
    with pytest.raises(_lib.HexGnnError):          # grad mode: inference only
        m(b.x, b.edge_index, b.batch, b.ptr)
    with pytest.raises(NotImplementedError):
        get_current_model("SAGE", norm=torch.nn.LayerNorm)
    with torch.no_grad(), pytest.raises(_lib.HexGnnError):
        m(b.x[:, :2].contiguous(), b.edge_index, b.batch, b.ptr)
