"""Tensor-level wrappers over the C-ABI (include/hexgnn.h).  PyTorch only provides device memory
and the current stream here; all arithmetic happens in libhexgnn.so."""
import ctypes as C
import os

import torch

from . import _lib
from ._lib import HEXGNN_ADV_ONLY, HEXGNN_SEPARATE, HEXGNN_PATH_SIMT, HEXGNN_PATH_TC, HEXGNN_TF32X1  # noqa: F401


# kernels launched by each C-ABI call (csr: init, hist, scan, fill, sort; forward: pack, fused;
# backward: pack, fused, reduce, unpack) -- bench.py's `gpu_launches` is read off this counter.
LAUNCHES = {"build_csr": 1, "forward": 2, "backward": 5, "td_loss": 1, "train_step": 6, "adam_step": 2, "mask_argmax": 1, "td_target": 1, "forward_big": 0, "allreduce_p2p": 1, "noisy": 1, "forward_lw": 0, "backward_lw": 0, "hexara_forward": 0, "hexara_backward": 0}
launch_counter = 0
NVTX = os.environ.get("HEXGNN_NVTX", "0") == "1"


def _count(what):
    global launch_counter
    launch_counter += LAUNCHES[what]


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _stream(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _on_tensor_device(fn):
    """Run a C-ABI wrapper with the CUDA device of its first CUDA argument current: the library launches on the current
    device and passes `torch.cuda.current_stream(tensor.device)`, so a model on cuda:1 called while cuda:0 is current
    would otherwise hand a stream of one device to a launch on another.  Costs nothing when the device already is
    current (the normal one-process-per-GPU case)."""
    import functools

    @functools.wraps(fn)
    def wrapped(*args, **kwargs):
        if NVTX:                                 # HEXGNN_NVTX=1: one range per C-ABI call (tracing)
            torch.cuda.nvtx.range_push("hexgnn::" + fn.__name__)
            try:
                return guarded(*args, **kwargs)
            finally:
                torch.cuda.nvtx.range_pop()
        return guarded(*args, **kwargs)

    def guarded(*args, **kwargs):
        idx = -1
        for a in args:
            t = a.blob if isinstance(a, CsrBlob) else a
            if isinstance(t, torch.Tensor) and t.is_cuda:
                idx = t.device.index
                break
        if idx < 0 or idx == torch.cuda.current_device():
            return fn(*args, **kwargs)
        with torch.cuda.device(idx):
            return fn(*args, **kwargs)
    return wrapped


def _require_cuda(t, name):
    if not t.is_cuda:
        raise _lib.HexGnnError(f"{name} must be a CUDA tensor: the fused path has no CPU fallback")


class CsrBlob:
    """Destination-sorted CSR + tile table of one batch, living in one int32 device blob."""

    def __init__(self, blob, N, E, B, tile_rows):
        self.blob, self.N, self.E, self.B, self.tile_rows = blob, N, E, B, tile_rows
        self.layout = _lib.csr_layout(N, E, B)

    def _view(self, off, n):
        return self.blob[off:off + n]

    @property
    def status(self):
        return self._view(self.layout.status, 1)

    @property
    def ntiles(self):
        return self._view(self.layout.ntiles, 1)

    @property
    def gptr(self):
        return self._view(self.layout.gptr, self.B + 1)

    @property
    def rowptr(self):
        return self._view(self.layout.rowptr, self.N + 1)

    @property
    def col(self):
        return self._view(self.layout.col, self.E)

    @property
    def tile_g0(self):
        return self._view(self.layout.tile_g0, self.B + 1)

    def raise_on_error(self):
        """Host sync: read the device status word and raise if any kernel flagged the batch."""
        st = int(self.status.item())
        if st:
            msgs = [m for bit, m in _lib.STATUS_NAMES.items() if st & bit]
            raise _lib.HexGnnError("hexgnn device status %d: %s" % (st, "; ".join(msgs)))


@_on_tensor_device
def build_csr(edge_index, batch=None, ptr=None, num_nodes=None, num_graphs=None, tile_rows=128):
    """Device CSR builder (hexgnn_build_csr).  edge_index [2,E] int64 CUDA; batch [N] int64 and/or
    ptr [B+1] int64; neither => single graph.  No host sync."""
    _require_cuda(edge_index, "edge_index")
    if edge_index.dtype != torch.int64 or edge_index.dim() != 2 or edge_index.shape[0] != 2:
        raise _lib.HexGnnError("edge_index must be int64 [2,E]")
    edge_index = edge_index.contiguous()
    E = int(edge_index.shape[1])
    if batch is not None:
        _require_cuda(batch, "batch")
        batch = batch.contiguous()
        if batch.dtype != torch.int64:
            raise _lib.HexGnnError("batch must be int64")
    if ptr is not None:
        if not ptr.is_cuda:
            ptr = ptr.to(edge_index.device, non_blocking=True)
        ptr = ptr.contiguous()
        if ptr.dtype != torch.int64:
            raise _lib.HexGnnError("ptr must be int64")
    if num_nodes is None:
        if batch is None:
            raise _lib.HexGnnError("num_nodes is required when batch is not given")
        num_nodes = int(batch.shape[0])
    if num_graphs is None:
        if ptr is not None:
            num_graphs = int(ptr.shape[0]) - 1
        elif batch is None:
            num_graphs = 1
        else:
            raise _lib.HexGnnError("num_graphs is required when only `batch` is given (avoids a host sync)")
    N, B = int(num_nodes), int(num_graphs)
    lay = _lib.csr_layout(N, E, B)
    blob = torch.empty(lay.total_words, dtype=torch.int32, device=edge_index.device)
    rc = _lib.lib().hexgnn_build_csr(_ptr(edge_index), E, _ptr(batch), _ptr(ptr), N, B, tile_rows, _ptr(blob),
                                     blob.numel() * 4, _stream(edge_index.device))
    _lib.check(rc, "hexgnn_build_csr")
    _count("build_csr")
    return CsrBlob(blob, N, E, B, tile_rows)


@_on_tensor_device
def fused_forward(desc, flat_params, x, csr, flags=0, want_stash=False, ws=None):
    """hexgnn_forward.  Returns (out_nodes [N], out_value [B], stash or None, hidden_padded)."""
    _require_cuda(x, "x")
    _require_cuda(flat_params, "parameters")
    if x.dtype != torch.float32 or x.dim() != 2 or x.shape[1] != 3:
        raise _lib.HexGnnError("x must be float32 [N,3]")
    x = x.contiguous()
    N, E, B = csr.N, csr.E, csr.B
    if x.shape[0] != N:
        raise _lib.HexGnnError("x rows != csr nodes")
    sz = _lib.sizes(desc, N, E, B, csr.tile_rows)
    dev = x.device
    if ws is None or ws.numel() * 4 < sz.fwd_ws_bytes:
        ws = torch.empty(sz.fwd_ws_bytes // 4, dtype=torch.float32, device=dev)
    out_nodes = torch.empty(N, dtype=torch.float32, device=dev)
    out_value = torch.empty(B, dtype=torch.float32, device=dev)
    stash = torch.empty(sz.stash_bytes // 4, dtype=torch.float32, device=dev) if want_stash else None
    rc = _lib.lib().hexgnn_forward(C.byref(desc), _ptr(flat_params), _ptr(x), N, E, B, _ptr(csr.blob), csr.tile_rows,
                                   flags, _ptr(out_nodes), _ptr(out_value), _ptr(stash), _ptr(ws), ws.numel() * 4,
                                   _stream(dev))
    _lib.check(rc, "hexgnn_forward")
    _count("forward")
    return out_nodes, out_value, stash, int(sz.hidden_padded)


STATIC_WS_BYTES = 2 * 1024 * 1024   # HEXGNN_STATIC_WS_BYTES


@_on_tensor_device
def fused_forward_static(desc, flat_params, x, csr, out_nodes, out_value, flags=0):
    """hexgnn_forward_static: inference forward whose scratch is a static device array inside libhexgnn.so -- the call
    allocates nothing (the caller passes the outputs: out_nodes [N] f32, out_value [B] f32).  Models beyond
    HEXGNN_STATIC_WS_BYTES are refused; calls that overlap on the GPU race on the shared scratch."""
    _require_cuda(x, "x")
    _require_cuda(flat_params, "parameters")
    N, E, B = csr.N, csr.E, csr.B
    if x.dtype != torch.float32 or not x.is_contiguous() or tuple(x.shape) != (N, 3):
        raise _lib.HexGnnError("x must be a contiguous float32 [N,3] tensor")
    for t, n, name in ((out_nodes, N, "out_nodes"), (out_value, B, "out_value")):
        if t is None or not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous() or t.numel() != n:
            raise _lib.HexGnnError(f"fused_forward_static: {name} must be a contiguous float32 CUDA tensor of {n} elements")
    rc = _lib.lib().hexgnn_forward_static(C.byref(desc), _ptr(flat_params), _ptr(x), N, E, B, _ptr(csr.blob), csr.tile_rows,
                                          flags, _ptr(out_nodes), _ptr(out_value), _stream(x.device))
    _lib.check(rc, "hexgnn_forward_static")
    _count("forward")
    return out_nodes, out_value


@_on_tensor_device
def fused_backward(desc, flat_params, x, csr, flags, d_nodes, d_value, stash, want_d_embeds=False, ws=None,
                   grad_out=None, d_x=None):
    """hexgnn_backward.  Returns (flat gradient [P], d_embeds [N,Hp] or None).  `d_x` ([N,3] f32 CUDA, optional): also
    receives dL/dx (hexgnn_backward_dx; zero-initialised by the caller for rows of graphs the kernels refuse)."""
    N, E, B = csr.N, csr.E, csr.B
    sz = _lib.sizes(desc, N, E, B, csr.tile_rows)
    dev = x.device
    x = x.contiguous()
    if ws is None or ws.numel() * 4 < sz.bwd_ws_bytes:
        ws = torch.empty(sz.bwd_ws_bytes // 4, dtype=torch.float32, device=dev)
    if grad_out is None:
        grad_out = torch.empty(flat_params.numel(), dtype=torch.float32, device=dev)
    d_nodes = d_nodes.contiguous().float()
    if d_value is not None:
        d_value = d_value.contiguous().float()
    d_embeds = torch.empty((N, int(sz.hidden_padded)), dtype=torch.float32, device=dev) if want_d_embeds else None
    if d_x is not None:
        if not d_x.is_cuda or d_x.dtype != torch.float32 or not d_x.is_contiguous() or tuple(d_x.shape) != (N, 3):
            raise _lib.HexGnnError("fused_backward: d_x must be a contiguous float32 CUDA tensor [N,3]")
        rc = _lib.lib().hexgnn_backward_dx(C.byref(desc), _ptr(flat_params), _ptr(x), N, E, B, _ptr(csr.blob), csr.tile_rows,
                                           flags, _ptr(d_nodes), _ptr(d_value), _ptr(stash), _ptr(grad_out), _ptr(d_embeds),
                                           _ptr(d_x), _ptr(ws), ws.numel() * 4, _stream(dev))
        _lib.check(rc, "hexgnn_backward_dx")
        _count("backward")
        return grad_out, d_embeds
    rc = _lib.lib().hexgnn_backward(C.byref(desc), _ptr(flat_params), _ptr(x), N, E, B, _ptr(csr.blob), csr.tile_rows,
                                    flags, _ptr(d_nodes), _ptr(d_value), _ptr(stash), _ptr(grad_out), _ptr(d_embeds),
                                    _ptr(ws), ws.numel() * 4, _stream(dev))
    _lib.check(rc, "hexgnn_backward")
    _count("backward")
    return grad_out, d_embeds


@_on_tensor_device
def lw_forward(desc, flat_params, norm_params, x, csr, flags=0, want_stash=False, ws=None):
    """hexgnn_forward_lw (layer-at-a-time tensor-core path: any graph size, optional graph-mode LayerNorm).
    Returns (out_nodes [N], out_value [B], stash or None, HexLwSizes)."""
    _require_cuda(x, "x")
    _require_cuda(flat_params, "parameters")
    if x.dtype != torch.float32 or x.dim() != 2 or x.shape[1] != 3:
        raise _lib.HexGnnError("x must be float32 [N,3]")
    x = x.contiguous()
    N, E, B = csr.N, csr.E, csr.B
    if x.shape[0] != N:
        raise _lib.HexGnnError("x rows != csr nodes")
    sz = _lib.lw_sizes(desc, N, E, B, norm_params is not None)
    if norm_params is not None and (not norm_params.is_cuda or norm_params.numel() != sz.norm_floats):
        raise _lib.HexGnnError("lw_forward: norm_params must be a CUDA tensor of hexgnn_lw_sizes().norm_floats")
    dev = x.device
    if ws is None or ws.numel() * 4 < sz.fwd_ws_bytes:
        ws = torch.empty((sz.fwd_ws_bytes + 3) // 4, dtype=torch.float32, device=dev)
    out_nodes = torch.empty(N, dtype=torch.float32, device=dev)
    out_value = torch.empty(B, dtype=torch.float32, device=dev)
    stash = torch.empty(sz.stash_bytes // 4, dtype=torch.float32, device=dev) if want_stash else None
    rc = _lib.lib().hexgnn_forward_lw(C.byref(desc), _ptr(flat_params), _ptr(norm_params), _ptr(x), N, E, B, _ptr(csr.blob),
                                      flags, _ptr(out_nodes), _ptr(out_value), _ptr(stash), _ptr(ws), ws.numel() * 4,
                                      _stream(dev))
    _lib.check(rc, "hexgnn_forward_lw")
    _count("forward_lw")
    return out_nodes, out_value, stash, sz


@_on_tensor_device
def lw_backward(desc, flat_params, norm_params, x, csr, flags, d_nodes, d_value, stash, want_d_embeds=False, ws=None,
                grad_out=None, grad_norm=None):
    """hexgnn_backward_lw.  Returns (flat gradient [P], norm gradient or None, d_embeds [N,Hp] or None)."""
    N, E, B = csr.N, csr.E, csr.B
    sz = _lib.lw_sizes(desc, N, E, B, norm_params is not None)
    dev = x.device
    x = x.contiguous()
    if ws is None or ws.numel() * 4 < sz.bwd_ws_bytes:
        ws = torch.empty((sz.bwd_ws_bytes + 3) // 4, dtype=torch.float32, device=dev)
    if grad_out is None:
        grad_out = torch.empty(flat_params.numel(), dtype=torch.float32, device=dev)
    if norm_params is not None and grad_norm is None:
        grad_norm = torch.empty_like(norm_params)
    d_nodes = d_nodes.contiguous().float()
    if d_value is not None:
        d_value = d_value.contiguous().float()
    d_embeds = torch.empty((N, int(sz.hidden_padded)), dtype=torch.float32, device=dev) if want_d_embeds else None
    rc = _lib.lib().hexgnn_backward_lw(C.byref(desc), _ptr(flat_params), _ptr(norm_params), _ptr(x), N, E, B, _ptr(csr.blob),
                                       flags, _ptr(d_nodes), _ptr(d_value), _ptr(stash), _ptr(grad_out), _ptr(grad_norm),
                                       _ptr(d_embeds), _ptr(ws), ws.numel() * 4, _stream(dev))
    _lib.check(rc, "hexgnn_backward_lw")
    _count("backward_lw")
    return grad_out, grad_norm, d_embeds


@torch.no_grad()
@_on_tensor_device
def hexara_forward(desc, flat_params, x, csr, want_embeds=False, ws=None, want_stash=False):
    """hexgnn_hexara_forward (HexAra `SAGE_torch_script`, GN0/torch_script_models.py:286-379).  Returns
    (pi [max entries], value [B], out_graph_indices [max entries] i64, out_batch_ptr [B+1] i64, embeds [N,Hp] or None, ws,
    stash or None); out_batch_ptr[B] entries of pi / out_graph_indices are valid."""
    _require_cuda(x, "x")
    _require_cuda(flat_params, "parameters")
    if x.dtype != torch.float32 or x.dim() != 2 or x.shape[1] != 3:
        raise _lib.HexGnnError("x must be float32 [N,3]")
    x = x.contiguous()
    N, E, B = csr.N, csr.E, csr.B
    if x.shape[0] != N:
        raise _lib.HexGnnError("x rows != csr nodes")
    sz = _lib.HexAraSizes()
    _lib.check(_lib.lib().hexgnn_hexara_sizes(C.byref(desc), N, E, B, C.byref(sz)), "hexgnn_hexara_sizes")
    dev = x.device
    if ws is None or ws.numel() * 4 < sz.ws_bytes or ws.device != dev:
        ws = torch.empty((sz.ws_bytes + 3) // 4, dtype=torch.float32, device=dev)
    cap = int(sz.max_policy_entries)
    pi = torch.empty(cap, dtype=torch.float32, device=dev)
    value = torch.empty(B, dtype=torch.float32, device=dev)
    gi = torch.empty(cap, dtype=torch.int64, device=dev)
    bp = torch.empty(B + 1, dtype=torch.int64, device=dev)
    emb = torch.empty((N, int(sz.hidden_padded)), dtype=torch.float32, device=dev) if want_embeds else None
    stash = torch.empty(int(sz.stash_bytes) // 4, dtype=torch.float32, device=dev) if want_stash else None
    rc = _lib.lib().hexgnn_hexara_forward(C.byref(desc), _ptr(flat_params), _ptr(x), N, E, B, _ptr(csr.blob), _ptr(pi),
                                          _ptr(value), _ptr(gi), _ptr(bp), _ptr(emb), _ptr(stash), _ptr(ws), ws.numel() * 4,
                                          _stream(dev))
    _lib.check(rc, "hexgnn_hexara_forward")
    _count("hexara_forward")
    return pi, value, gi, bp, emb, ws, stash


@torch.no_grad()
@_on_tensor_device
def hexara_backward(desc, flat_params, x, csr, pi, out_batch_ptr, d_pi, d_value, stash, grad_out=None, ws=None):
    """hexgnn_hexara_backward.  Returns (flat gradient [P], ws)."""
    N, E, B = csr.N, csr.E, csr.B
    sz = _lib.HexAraSizes()
    _lib.check(_lib.lib().hexgnn_hexara_sizes(C.byref(desc), N, E, B, C.byref(sz)), "hexgnn_hexara_sizes")
    dev = x.device
    x = x.contiguous()
    if ws is None or ws.numel() * 4 < sz.bwd_ws_bytes or ws.device != dev:
        ws = torch.empty((sz.bwd_ws_bytes + 3) // 4, dtype=torch.float32, device=dev)
    if grad_out is None:
        grad_out = torch.empty(flat_params.numel(), dtype=torch.float32, device=dev)
    d_pi = d_pi.contiguous().float() if d_pi is not None else None
    d_value = d_value.contiguous().float() if d_value is not None else None
    rc = _lib.lib().hexgnn_hexara_backward(C.byref(desc), _ptr(flat_params), _ptr(x), N, E, B, _ptr(csr.blob), _ptr(pi),
                                           _ptr(out_batch_ptr), _ptr(d_pi), _ptr(d_value), _ptr(stash), _ptr(grad_out),
                                           _ptr(ws), ws.numel() * 4, _stream(dev))
    _lib.check(rc, "hexgnn_hexara_backward")
    _count("hexara_backward")
    return grad_out, ws


LOSS_KINDS = {"mse": 0, "smooth_l1": 1, "huber": 1}


@_on_tensor_device
def fused_train_step(desc, flat_params, x, csr, ptr, action, target, weight=None, kind="mse", scale=1.0, grad_out=None,
                     ws=None):
    """hexgnn_train_step: forward + TD loss + backward + gradient reduction in one call (rainbow.py:301-316).
    Returns (loss [], td_err [B], q [N], flat gradient [P]).  Launches pack, forward, chain, dW, reduce (+ loss block), unpack."""
    _require_cuda(x, "x")
    _require_cuda(flat_params, "parameters")
    if kind not in LOSS_KINDS:
        raise _lib.HexGnnError(f"train_step: unknown loss kind {kind!r}")
    if x.dtype != torch.float32 or x.dim() != 2 or x.shape[1] != 3:
        raise _lib.HexGnnError("x must be float32 [N,3]")
    x = x.contiguous()
    N, E, B = csr.N, csr.E, csr.B
    dev = x.device
    ptr = ptr.to(dev, non_blocking=True).contiguous()
    action = action.to(dev, non_blocking=True).contiguous()
    target = target.to(dev, non_blocking=True).contiguous().float()
    if weight is not None:
        weight = weight.to(dev, non_blocking=True).contiguous().float()
    if ptr.dtype != torch.int64 or action.dtype != torch.int64 or ptr.numel() != B + 1 or target.numel() != B or \
            action.numel() != B or x.shape[0] != N:
        raise _lib.HexGnnError("train_step: x [N,3], ptr [B+1] int64, action [B] int64, target [B] expected")
    sz = _lib.sizes(desc, N, E, B, csr.tile_rows)
    if ws is None or ws.numel() * 4 < sz.bwd_ws_bytes:
        ws = torch.empty(sz.bwd_ws_bytes // 4, dtype=torch.float32, device=dev)
    if grad_out is None:
        grad_out = torch.empty(flat_params.numel(), dtype=torch.float32, device=dev)
    q = torch.empty(N, dtype=torch.float32, device=dev)
    # loss and the B TD errors share one buffer: what the learner reads back every step (loss + PER priorities,
    # rainbow.py:306-309) is ONE device-to-host copy of `loss_and_td_err(loss)`
    out = torch.empty(1 + B, dtype=torch.float32, device=dev)
    loss, td_err = out[:1], out[1:]
    stash = torch.empty(sz.stash_bytes // 4, dtype=torch.float32, device=dev)
    rc = _lib.lib().hexgnn_train_step(C.byref(desc), _ptr(flat_params), _ptr(x), N, E, B, _ptr(csr.blob), csr.tile_rows, 0,
                                      _ptr(ptr), _ptr(action), _ptr(target), _ptr(weight), LOSS_KINDS[kind], float(scale),
                                      _ptr(q), _ptr(loss), _ptr(td_err), _ptr(stash), _ptr(grad_out),
                                      _ptr(ws), ws.numel() * 4, _stream(dev))
    _lib.check(rc, "hexgnn_train_step")
    _count("train_step")
    return loss.reshape(()), td_err, q, grad_out


def loss_and_td_err(loss):
    """The contiguous [1 + B] device tensor behind the (loss, td_err) pair `fused_train_step` / `model.train_step` return
    (loss first): one device-to-host copy fetches both."""
    base = loss._base if loss._base is not None else loss
    return base


class _TDLoss(torch.autograd.Function):
    """hexgnn_td_loss behind autograd: forward computes loss, td errors AND d loss / d q in one launch;
    backward only scales the stored gradient by the incoming scalar."""

    @staticmethod
    def forward(ctx, q, ptr, action, target, weight, kind, scale):
        _require_cuda(q, "q")
        dev = q.device
        q = q.contiguous().float()
        N, B = int(q.shape[0]), int(action.shape[0])
        ptr = ptr.to(dev, non_blocking=True).contiguous()
        action = action.to(dev, non_blocking=True).contiguous()
        target = target.to(dev, non_blocking=True).contiguous().float()
        if weight is not None:
            weight = weight.to(dev, non_blocking=True).contiguous().float()
        if ptr.dtype != torch.int64 or action.dtype != torch.int64 or ptr.numel() != B + 1 or target.numel() != B:
            raise _lib.HexGnnError("td_loss: ptr [B+1] int64, action [B] int64, target [B] expected")
        loss = torch.empty(1, dtype=torch.float32, device=dev)
        td_err = torch.empty(B, dtype=torch.float32, device=dev)
        need_grad = ctx.needs_input_grad[0]
        d_q = torch.empty(N, dtype=torch.float32, device=dev) if need_grad else None
        with torch.cuda.device(dev):
            rc = _lib.lib().hexgnn_td_loss(_ptr(q), _ptr(ptr), _ptr(action), _ptr(target), _ptr(weight), N, B, int(kind),
                                           float(scale), _ptr(loss), _ptr(td_err), _ptr(d_q), _stream(dev))
        _lib.check(rc, "hexgnn_td_loss")
        _count("td_loss")
        ctx.d_q = d_q
        ctx.mark_non_differentiable(td_err)
        return loss.reshape(()), td_err

    @staticmethod
    def backward(ctx, g_loss, _g_err):
        d_q = ctx.d_q
        ctx.d_q = None
        return (d_q * g_loss if d_q is not None else None), None, None, None, None, None, None


def td_loss(q, ptr, action, target, weight=None, kind="mse", scale=1.0):
    """Rainbow learner loss on device (rainbow.py:301-312): returns (loss, td_err) with
    ``td_est = q[action + ptr[:-1]]``, ``td_err = td_est - target`` (PER priorities = |td_err| + 1e-6) and
    ``loss = scale * mean(weight * l(target, td_est))``, l = MSE ("mse") or SmoothL1 ("smooth_l1").
    Differentiable w.r.t. ``q``; one kernel launch forward, one elementwise scale backward."""
    if kind not in LOSS_KINDS:
        raise _lib.HexGnnError(f"td_loss: unknown loss kind {kind!r}")
    return _TDLoss.apply(q, ptr, action, target, weight, LOSS_KINDS[kind], scale)


@torch.no_grad()
@_on_tensor_device
def mask_argmax(q, ptr, mask_terminals=True, mask_value=-5.0):
    """rainbow.py:222-225 / 263-266 on device, one launch: overwrite the two terminal rows of every graph in ``q``
    (in place) with ``mask_value`` and return (max [B], argmax [B] as GLOBAL row ids; ties -> lowest index).
    Local actions are ``argmax - ptr[:-1]``."""
    _require_cuda(q, "q")
    if q.dtype != torch.float32 or not q.is_contiguous() or q.dim() != 1:
        raise _lib.HexGnnError("mask_argmax: q must be a contiguous float32 [N] tensor (it is masked in place)")
    dev = q.device
    ptr = ptr.to(dev, non_blocking=True).contiguous()
    N, B = int(q.shape[0]), int(ptr.shape[0]) - 1
    mx = torch.empty(B, dtype=torch.float32, device=dev)
    arg = torch.empty(B, dtype=torch.int64, device=dev)
    rc = _lib.lib().hexgnn_mask_argmax(_ptr(q), _ptr(ptr), N, B, 1 if mask_terminals else 0, float(mask_value),
                                       _ptr(mx), _ptr(arg), _stream(dev))
    _lib.check(rc, "hexgnn_mask_argmax")
    _count("mask_argmax")
    return mx, arg


@torch.no_grad()
@_on_tensor_device
def td_target(q_target, best, reward, done, gamma_n):
    """rainbow.py:268-269 on device: ``reward + gamma_n * q_target[best] * (1 - done)``."""
    _require_cuda(q_target, "q_target")
    dev = q_target.device
    q_target = q_target.contiguous().float()
    best = best.to(dev).contiguous()
    reward = reward.to(dev).contiguous().float()
    done = done.to(dev).contiguous().float()
    B = int(best.shape[0])
    out = torch.empty(B, dtype=torch.float32, device=dev)
    rc = _lib.lib().hexgnn_td_target(_ptr(q_target), _ptr(best), _ptr(reward), _ptr(done), float(gamma_n),
                                     int(q_target.shape[0]), B, _ptr(out), _stream(dev))
    _lib.check(rc, "hexgnn_td_target")
    _count("td_target")
    return out


@_on_tensor_device
def forward_big(desc, flat_params, x, csr, flags, max_rows, out_nodes, out_value, ws=None):
    """hexgnn_forward_big: after fused_forward, evaluate the graphs that do not fit one tile (inference only) and
    overwrite their rows of ``out_nodes`` / ``out_value`` in place.  Launches 2 + num_layers + num_head_layers kernels."""
    N, E, B = csr.N, csr.E, csr.B
    dev = x.device
    hp = (int(desc.hidden) + 15) // 16 * 16
    lt = int(desc.num_layers) + int(desc.num_head_layers)
    need = 2 * N * hp + (lt - 1) * hp * 2 * hp * 2 + lt * hp      # activations ping-pong + packed weights (pack.cu)
    if ws is None or ws.numel() < need:
        ws = torch.empty(need, dtype=torch.float32, device=dev)
    rc = _lib.lib().hexgnn_forward_big(C.byref(desc), _ptr(flat_params), _ptr(x.contiguous()), N, E, B, _ptr(csr.blob),
                                       csr.tile_rows, flags, int(max_rows), _ptr(out_nodes), _ptr(out_value), _ptr(ws),
                                       ws.numel() * 4, _stream(dev))
    _lib.check(rc, "hexgnn_forward_big")
    global launch_counter
    launch_counter += 3 + int(desc.num_layers) + int(desc.num_head_layers)
    return ws


@_on_tensor_device
def allreduce_p2p(peer_ptrs_dev, world, n, out):
    """hexgnn_allreduce_p2p: out[n] = sum over ranks (in rank order) of the peer-mapped buffers whose device pointers
    are stored in the DEVICE array at address `peer_ptrs_dev`.  The caller provides the cross-rank barriers."""
    _require_cuda(out, "out")
    rc = _lib.lib().hexgnn_allreduce_p2p(C.c_void_p(int(peer_ptrs_dev)), int(world), int(n), _ptr(out), _stream(out.device))
    _lib.check(rc, "hexgnn_allreduce_p2p")
    _count("allreduce_p2p")
    return out


def allreduce_mc(mc_ptr, peer_flags_dev, rank, world, n, epoch, counter):
    """hexgnn_allreduce_mc: two-shot in-place all-reduce of the symmetric buffers behind the multicast address `mc_ptr`
    (multimem.ld_reduce / multimem.st through the NVSwitch), cross-rank barriers included."""
    _require_cuda(counter, "counter")
    with torch.cuda.device(counter.device):
        rc = _lib.lib().hexgnn_allreduce_mc(C.c_void_p(int(mc_ptr)), C.c_void_p(int(peer_flags_dev)), int(rank), int(world),
                                            int(n), int(epoch), _ptr(counter), _stream(counter.device))
    _lib.check(rc, "hexgnn_allreduce_mc")
    _count("allreduce_p2p")


@torch.no_grad()
@_on_tensor_device
def adam_step(desc, flat_params, flat_grad, exp_avg, exp_avg_sq, state, x, lr, betas=(0.9, 0.999), eps=1e-8,
              max_grad_norm=0.0, skip_untaken=False, grad_norm_out=None, ws=None):
    """hexgnn_adam_step over the flat buffers (rainbow.py:320-326): clip_grad_norm_ + Adam.step in two launches.
    `state` is a zero-initialised int32 [4] device tensor; returns the pre-clip gradient norm as a device scalar."""
    for t, name in ((flat_params, "parameters"), (flat_grad, "gradient"), (exp_avg, "exp_avg"), (exp_avg_sq, "exp_avg_sq")):
        _require_cuda(t, name)
        if t.dtype != torch.float32 or not t.is_contiguous() or t.numel() != flat_params.numel():
            raise _lib.HexGnnError(f"adam_step: {name} must be a contiguous float32 buffer of the flat parameter size")
    if state.dtype != torch.int32 or state.numel() < 4 or not state.is_cuda:
        raise _lib.HexGnnError("adam_step: state must be an int32 [4] CUDA tensor")
    dev = flat_params.device
    need = int(_lib.lib().hexgnn_adam_ws_bytes()) // 4
    if ws is None or ws.numel() < need:
        ws = torch.empty(need, dtype=torch.float32, device=dev)
    if grad_norm_out is None:
        grad_norm_out = torch.empty(1, dtype=torch.float32, device=dev)
    rc = _lib.lib().hexgnn_adam_step(C.byref(desc), _ptr(flat_params), _ptr(flat_grad), _ptr(exp_avg), _ptr(exp_avg_sq),
                                     _ptr(state), _ptr(x), float(lr), float(betas[0]), float(betas[1]), float(eps),
                                     float(max_grad_norm), 1 if skip_untaken else 0, _ptr(grad_norm_out), _ptr(ws),
                                     ws.numel() * 4, _stream(dev))
    _lib.check(rc, "hexgnn_adam_step")
    _count("adam_step")
    return grad_norm_out.reshape(())


@_on_tensor_device
def noisy_compose(desc, flat_params, noisy):
    """hexgnn_noisy_compose: write `mu + sigma * epsilon` of both heads' FactorizedNoisyLinear (GN0/models.py:134-141) into
    the advantage-Linear slots of the flat parameter buffer.  noisy: [2, 3, H+1] f32 CUDA."""
    _require_cuda(flat_params, "parameters")
    _require_cuda(noisy, "noisy")
    if noisy.dtype != torch.float32 or not noisy.is_contiguous() or noisy.numel() != 6 * (desc.hidden + 1):
        raise _lib.HexGnnError("noisy_compose: noisy must be a contiguous float32 [2, 3, hidden + 1] tensor")
    _lib.check(_lib.lib().hexgnn_noisy_compose(C.byref(desc), _ptr(flat_params), _ptr(noisy), _stream(flat_params.device)),
               "hexgnn_noisy_compose")
    _count("noisy")


@_on_tensor_device
def noisy_grads(desc, flat_grad, noisy, d_noisy, accumulate):
    """hexgnn_noisy_grads: d_mu = g, d_sigma = g * epsilon from the composed weights' slice of the flat gradient.
    d_noisy: [2, 2, H+1] f32 CUDA, overwritten or (accumulate) added to."""
    for t, name in ((flat_grad, "gradient"), (noisy, "noisy"), (d_noisy, "d_noisy")):
        _require_cuda(t, name)
        if t.dtype != torch.float32 or not t.is_contiguous():
            raise _lib.HexGnnError(f"noisy_grads: {name} must be contiguous float32")
    if noisy.numel() != 6 * (desc.hidden + 1) or d_noisy.numel() != 4 * (desc.hidden + 1):
        raise _lib.HexGnnError("noisy_grads: noisy is [2, 3, hidden + 1], d_noisy is [2, 2, hidden + 1]")
    _lib.check(_lib.lib().hexgnn_noisy_grads(C.byref(desc), _ptr(flat_grad), _ptr(noisy), _ptr(d_noisy),
                                             1 if accumulate else 0, _stream(flat_grad.device)), "hexgnn_noisy_grads")
    _count("noisy")
