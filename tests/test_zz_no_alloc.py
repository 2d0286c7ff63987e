"""GPU: hexgnn_forward_static -- the inference forward whose scratch (weight images) is a static device array inside
libhexgnn.so, for callers that cannot allocate (no reference counterpart).
Called through the C-ABI (ctypes) and compared with hexgnn_forward on two calls in a row; the second call uses other
weights, another head and another batch, so stale images of the first call in the shared scratch would show.  Collected
last on purpose; passes alone in a fresh process and after the other tests."""
import ctypes as C

import pytest
import torch

import positions
from helpers import REL_TOL, make_pair, rel_err
from gnn_hex_b200 import _lib, data, ops

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _both(m, batch, flags):
    d = batch.to(DEV)
    csr = m.build_csr(d.edge_index, d.batch, d.ptr, num_nodes=int(d.x.shape[0]))
    flat = m.flat_parameters
    ref_nodes, ref_value, _, _ = ops.fused_forward(m._desc, flat, d.x, csr, flags)
    out_nodes = torch.full((csr.N,), float("nan"), device=DEV)
    out_value = torch.full((csr.B,), float("nan"), device=DEV)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    ops.fused_forward_static(m._desc, flat, d.x, csr, out_nodes, out_value, flags)
    assert torch.cuda.memory_allocated() == before        # nothing came from torch's allocator either
    torch.cuda.synchronize()
    assert int(csr.status.item()) == 0
    return (out_nodes, out_value), (ref_nodes, ref_value)


def test_forward_static_matches_forward_twice_in_a_row():
    _, m1 = make_pair(110, 15, seed=0)
    _, m2 = make_pair(110, 15, seed=7)
    m1, m2 = m1.to(DEV), m2.to(DEV)
    b1 = data.collate(positions.load_positions("pos_11x11_seed1234.npz", maker=1)[:96])
    b2 = data.collate(positions.load_positions("pos_11x11_seed1234.npz", maker=0)[100:140])
    (q, _), (q_ref, _) = _both(m1, b1, 0)
    assert rel_err(q, q_ref) <= REL_TOL and torch.equal(q, q_ref)
    (a, v), (a_ref, v_ref) = _both(m2, b2, ops.HEXGNN_SEPARATE)      # second call: other weights / head / batch / mode
    assert rel_err(a, a_ref) <= REL_TOL and rel_err(v, v_ref) <= REL_TOL
    assert torch.equal(a, a_ref) and torch.equal(v, v_ref)
    (q3, _), (q3_ref, _) = _both(m1, b1, 0)                          # and back: the first result again, bit for bit
    assert torch.equal(q3, q) and torch.equal(q3_ref, q_ref)


def test_forward_static_small_model_and_size_limit():
    _, m = make_pair(35, 10)
    m = m.to(DEV)
    b = data.collate(positions.load_positions("pos_7x7_seed1234.npz", maker=1)[:32])
    (q, _), (q_ref, _) = _both(m, b, 0)
    assert rel_err(q, q_ref) <= REL_TOL
    # a model whose weight images exceed HEXGNN_STATIC_WS_BYTES is refused with a message naming both sizes
    d = b.to(DEV)
    csr = m.build_csr(d.edge_index, d.batch, d.ptr, num_nodes=int(d.x.shape[0]))
    big = _lib.make_desc(110, 40, 2)
    out = torch.empty(csr.N, device=DEV)
    val = torch.empty(csr.B, device=DEV)
    params = torch.zeros(int(_lib.lib().hexgnn_param_count(C.byref(big))), device=DEV)
    with pytest.raises(_lib.HexGnnError, match="HEXGNN_STATIC_WS_BYTES"):
        ops.fused_forward_static(big, params, d.x, csr, out, val, 0)
    with pytest.raises(_lib.HexGnnError):                            # CUDA-core path: no static variant
        ops.fused_forward_static(m._desc, m.flat_parameters, d.x, csr, out, val, ops.HEXGNN_PATH_SIMT)
