"""CPU: the C-ABI library builds, loads and exports every symbol include/hexgnn.h declares."""
import ctypes
import os
import re

from gnn_hex_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_exported(lib_built):
    hdr = open(os.path.join(ROOT, "include", "hexgnn.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(hexgnn_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    so = ctypes.CDLL(lib_built)
    for name in declared:
        assert hasattr(so, name), name


def test_abi_version_and_param_tables(lib_built):
    l = _lib.lib()
    assert l.hexgnn_abi_version() == _lib.ABI_VERSION
    # parameter counts from SURVEY.md §7.0
    assert l.hexgnn_param_count(_lib.make_desc(110, 15, 2)) == 486974
    assert l.hexgnn_param_count(_lib.make_desc(35, 10, 2)) == 37382
    offs = _lib.param_offsets(_lib.make_desc(35, 10, 2))
    assert len(offs) == 3 * 10 + 2 * (3 * 2 + 6)
    assert offs == sorted(offs) and offs[0] == 0


def test_error_convention(lib_built):
    l = _lib.lib()
    bad = _lib.make_desc(35, 10, 2)
    bad.in_channels = 3
    assert l.hexgnn_param_count(bad) == -1
    s = _lib.HexSizes()
    rc = l.hexgnn_sizes(ctypes.byref(_lib.make_desc(35, 10, 2)), 10, 10, 1, 77, ctypes.byref(s))
    assert rc == -2 and b"tile_rows" in l.hexgnn_last_error()
    # null pointers are rejected before any CUDA call
    rc = l.hexgnn_forward(ctypes.byref(_lib.make_desc(35, 10, 2)), None, None, 1, 0, 1, None, 128, 0, None, None, None,
                          None, 0, None)
    assert rc == -1


def test_state_dict_names_match_reference_convention(lib_built):
    from gnn_hex_b200 import ModernTwoHeaded
    m = ModernTwoHeaded(35, 10)
    keys = list(m.state_dict().keys())
    assert "gnn.convs.0.lin_l.weight" in keys and "gnn.convs.0.lin_l.bias" in keys
    assert "gnn.convs.9.lin_r.weight" in keys and "gnn.convs.0.lin_r.bias" not in keys
    assert "maker_head.gnn.convs.1.lin_l.weight" in keys and "breaker_head.linear.bias" in keys
    assert "maker_head.value_head.layers.0.weight" in keys and "maker_head.value_head.layers.1.bias" in keys
    assert tuple(m.maker_head.value_head.layers[0].weight.shape) == (17, 140)
    assert m.supports_cache and not m.gnn.has_cache and m.gnn.hidden_channels == 35


def test_custom_ops_registered_with_fake_kernels(lib_built):
    """torch.ops.hexgnn.{build_csr, fused_forward, fused_backward}: schemas and shape inference, no compute."""
    import torch
    import gnn_hex_b200  # noqa: F401
    from torch._subclasses.fake_tensor import FakeTensorMode
    for name in ("build_csr", "fused_forward", "fused_backward"):
        assert hasattr(torch.ops.hexgnn, name)
    desc = _lib.make_desc(35, 10, 2)
    P = _lib.lib().hexgnn_param_count(desc)
    with FakeTensorMode():
        ei = torch.empty(2, 500, dtype=torch.int64, device="cuda")
        x = torch.empty(100, 3, device="cuda")
        ptr = torch.empty(5, dtype=torch.int64, device="cuda")
        params = torch.empty(P, device="cuda")
        csr = torch.ops.hexgnn.build_csr(ei, None, ptr, 100, 4, 128)
        assert csr.dtype == torch.int32 and csr.numel() == _lib.csr_layout(100, 500, 4).total_words
        qn, qv, st = torch.ops.hexgnn.fused_forward(params, x, csr, 500, 4, 35, 10, 2, 128, 0, True)
        assert qn.shape == (100,) and qv.shape == (4,)
        assert st.numel() * 4 == _lib.sizes(desc, 100, 500, 4, 128).stash_bytes
        _, _, st0 = torch.ops.hexgnn.fused_forward(params, x, csr, 500, 4, 35, 10, 2, 128, 1, False)
        assert st0.numel() == 0
        g = torch.ops.hexgnn.fused_backward(params, x, csr, st, qn, None, 500, 4, 35, 10, 2, 128, 0)
        assert g.shape == (P,)
        # HexAra network: capacity shapes of the 4-tuple (N - 2B entries, + B with swap)
        gi = torch.empty(100, dtype=torch.int64, device="cuda")
        pa = torch.empty(1000, device="cuda")
        pi, v, ogi, obp = torch.ops.hexgnn.hexara_forward(pa, x, ei, gi, ptr, 60, 15, 2, 2, True)
        assert pi.shape == (100 - 8 + 4,) and v.shape == (4,) and ogi.dtype == torch.int64 and obp.shape == (5,)
        assert torch.ops.hexgnn.hexara_forward(pa, x, ei, gi, ptr, 60, 15, 2, 2, False)[0].shape == (92,)


def test_deepcopy_and_pickle_reflatten(lib_built):
    """`copy.deepcopy(model)` / pickling clone every Parameter, so the copy's names no longer alias its flat
    buffer; the module must notice and re-flatten instead of feeding stale weights to the kernels."""
    import copy
    import io
    import torch
    from gnn_hex_b200 import ModernTwoHeaded
    torch.manual_seed(0)
    m = ModernTwoHeaded(12, 3, 2)
    m._ensure_flat()
    assert m._aliased()
    for clone in (copy.deepcopy(m), None):
        if clone is None:
            buf = io.BytesIO()
            torch.save(m, buf)
            buf.seek(0)
            clone = torch.load(buf, weights_only=False)
        clone.load_state_dict({k: v + 1 for k, v in m.state_dict().items()})
        clone._ensure_flat()
        assert clone._aliased() and clone._flat.data_ptr() != m._flat.data_ptr()
        w = clone.gnn.convs[1].lin_r.weight
        off = clone._slices[5][0]
        assert torch.equal(clone._flat[off:off + w.numel()], w.reshape(-1))
        assert torch.equal(w, m.gnn.convs[1].lin_r.weight + 1)
        with torch.no_grad():
            w.add_(1)                                   # an optimizer step on the names moves the kernels' buffer too
        assert torch.equal(clone._flat[off:off + w.numel()], ((m.gnn.convs[1].lin_r.weight + 1) + 1).reshape(-1))
    assert m._aliased()
