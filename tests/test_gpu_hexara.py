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
    with pytest.raises(NotImplementedError):
        get_current_model("SAGE", norm=torch.nn.LayerNorm)
    with torch.no_grad(), pytest.raises(_lib.HexGnnError):
        m(b.x[:, :2].contiguous(), b.edge_index, b.batch, b.ptr)


@pytest.mark.parametrize("shape,swap", [((28, 5, 2, 2), True), ((60, 15, 2, 2), False), ((44, 3, 1, 3), True), ((32, 1, 2, 1), True)])
def test_training_gradients(shape, swap):
    """`combined_loss.backward()` of the HexAra trainer (rl_loop/trainer_agent_pytorch.py:245-275: value MSE * batch size +
    policy cross-entropy) through hexgnn_hexara_backward: every parameter gradient against the fp64 oracle; a second backward
    accumulates; the result is bit-reproducible."""
    H, L, Lp, Lv = shape
    o = orc.HexAraOracle(H, L, Lp, Lv, swap_allowed=swap)
    helpers.set_hash_weights(o, seed=5, gain=1.5)
    o64 = orc.HexAraOracle(H, L, Lp, Lv, swap_allowed=swap).double()
    o64.load_state_dict({k: v.double() for k, v in o.state_dict().items()})
    m = get_current_model("SAGE", hidden_channels=H, hidden_layers=L, policy_layers=Lp, value_layers=Lv, swap_allowed=swap)
    m.load_state_dict(o.state_dict())
    m = m.to(DEV)
    b = data.collate(positions.load_positions_cpp("pos_7x7_seed1234.npz")[:48])
    d = b.to(DEV)
    g = torch.Generator().manual_seed(3)
    y = torch.rand(b.num_graphs, generator=g) * 2 - 1

    def loss_of(model, bb, dt, dev):
        pi, value, gi, bp = model(bb.x.to(dt), bb.edge_index, bb.batch, bb.ptr)
        # a seeded target distribution per graph over the policy entries
        gt = torch.Generator().manual_seed(4)
        t = torch.rand(pi.shape[0], generator=gt).double()
        gic = gi.cpu()
        tsum = torch.zeros(bb.num_graphs, dtype=torch.float64).index_add_(0, gic, t)       # on the CPU: CUDA index_add_ is
        t = (t / tsum[gic]).to(dt).to(dev)                                                  # an atomic, order-dependent sum
        policy_loss = -(t * pi).sum() / bb.num_graphs
        value_loss = torch.nn.functional.mse_loss(value.reshape(-1), y.to(dev).to(dt)) * bb.num_graphs
        return 0.01 * value_loss + 0.99 * policy_loss

    loss_of(o, b, torch.float32, "cpu").backward()
    l64 = loss_of(o64, b, torch.float64, "cpu")
    l64.backward()
    loss = loss_of(m, d, torch.float32, DEV)
    loss.backward()
    assert rel_err(loss, l64) <= REL_TOL
    glob, worst, name = helpers.grad_errs(m, o, o64)
    assert glob <= REL_TOL or worst <= 1.0, (glob, worst, name)
    if not swap:    # without a swap entry the swap MLP is not on the path: the reference leaves its .grad at None, here zeros
        assert not m.my_modules["swap_linear"].layers[0].weight.grad.any()
    g1 = m.flat_gradient.clone()
    loss_of(m, d, torch.float32, DEV).backward()
    assert torch.equal(m.flat_gradient, g1)                        # the same bits again ...
    assert torch.equal(m.gnn.convs[0].lin_l.weight.grad, 2 * g1[:m.gnn.convs[0].lin_l.weight.numel()].view_as(m.gnn.convs[0].lin_l.weight))  # ... accumulated
    opt = torch.optim.Adam(m.parameters(), lr=1e-3)
    w0 = m.my_modules["policy_head"].convs[-1].lin_l.weight.detach().clone()
    opt.step()
    assert not torch.equal(m.my_modules["policy_head"].convs[-1].lin_l.weight.detach(), w0)
    with torch.no_grad():
        m(d.x, d.edge_index, d.batch, d.ptr)                       # the kernels see the stepped weights (views of the flat buffer)


def test_hexara_single_graph_and_isolated_nodes():
    """One graph per call (the C++ engine's batch of one, nn_api.cpp:14-19), isolated nodes, a graph of three nodes (the
    smallest the reference accepts, torch_script_models.py:311): outputs against the fp64 oracle."""
    o = orc.HexAraOracle(28, 3, 2, 2, swap_allowed=True)
    helpers.set_hash_weights(o, seed=6, gain=1.5)
    o64 = orc.HexAraOracle(28, 3, 2, 2, swap_allowed=True).double()
    o64.load_state_dict({k: v.double() for k, v in o.state_dict().items()})
    m = get_current_model("SAGE", hidden_channels=28, hidden_layers=3, swap_allowed=True)
    m.load_state_dict(o.state_dict())
    m = m.to(DEV)
    full = positions.load_positions_cpp("pos_7x7_seed1234.npz")
    e = torch.tensor([[0, 2], [2, 0]], dtype=torch.long)
    x3 = torch.tensor([[1., 0, 1], [1, 0, 1], [0, 0, 1]])
    tiny = data.HexGraph(x=x3, edge_index=e)                      # node 1 isolated, swap flag set
    for gs in ([full[0]], [tiny], [tiny, full[3], tiny]):
        b = data.collate(gs)
        d = b.to(DEV)
        with torch.no_grad():
            pi, value, gi, bp = m(d.x, d.edge_index, d.batch, d.ptr)
            r = o64(b.x.double(), b.edge_index, b.batch, b.ptr)
        assert torch.equal(gi.cpu(), r[2]) and torch.equal(bp.cpu(), r[3])
        assert float((pi.cpu().double() - r[0]).abs().max()) <= REL_TOL * max(1.0, float(r[0].abs().max()))
        assert float((value.cpu().double() - r[1]).abs().max()) <= REL_TOL


def test_hexara_custom_op_matches_the_module():
    """torch.ops.hexgnn.hexara_forward (the functional face a TorchScript / C++ front end binds) == the module, bit for bit."""
    m = get_current_model("SAGE", hidden_channels=28, hidden_layers=4, swap_allowed=True).to(DEV)
    b = data.collate(positions.load_positions_cpp("pos_7x7_seed1234.npz")[:12]).to(DEV)
    with torch.no_grad():
        pi, value, gi, bp = m(b.x, b.edge_index, b.batch, b.ptr)
        pi2, value2, gi2, bp2 = torch.ops.hexgnn.hexara_forward(m.flat_parameters, b.x, b.edge_index, b.batch, b.ptr, 28, 4, 2, 2, True)
    n = int(bp[-1])
    assert torch.equal(bp, bp2) and torch.equal(value, value2) and torch.equal(pi, pi2[:n]) and torch.equal(gi, gi2[:n])
