"""ctypes binding of libhexgnn.so (include/hexgnn.h).  Fails loudly when the library is missing:
there is no CPU or stock-PyTorch fallback for the product path."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HEXGNN_LIB") or os.path.join(_HERE, "libhexgnn.so")

HEXGNN_ADV_ONLY = 1
HEXGNN_SEPARATE = 2
HEXGNN_PATH_SIMT = 0x100
HEXGNN_PATH_TC = 0x200
HEXGNN_TF32X1 = 0x400

ST_GRAPH_TOO_LARGE = 1
ST_EDGE_CROSSES = 2
ST_BATCH_UNSORTED = 4
ST_FLAG_MIXED = 8
ST_FP16_RANGE = 16
STATUS_NAMES = {
    ST_GRAPH_TOO_LARGE: "a graph has more nodes/edges than one tile holds (raise tile_rows)",
    ST_EDGE_CROSSES: "an edge joins two different graphs or is out of range",
    ST_BATCH_UNSORTED: "`batch` is not sorted / out of range",
    ST_FLAG_MIXED: "x[:,2] is not constant over the batch (reference asserts this, GN0/models.py:538)",
    ST_FP16_RANGE: "an activation left the fp16 range of the forward kernel's mixed product scheme (|h| >= 60000): use path='simt'",
}

ABI_VERSION = 11


class HexModelDesc(C.Structure):
    _fields_ = [("hidden", C.c_int32), ("num_layers", C.c_int32), ("num_head_layers", C.c_int32),
                ("in_channels", C.c_int32)]


class HexSizes(C.Structure):
    _fields_ = [("csr_bytes", C.c_int64), ("fwd_ws_bytes", C.c_int64), ("stash_bytes", C.c_int64),
                ("bwd_ws_bytes", C.c_int64), ("hidden_padded", C.c_int64)]


class HexLwSizes(C.Structure):
    _fields_ = [("fwd_ws_bytes", C.c_int64), ("bwd_ws_bytes", C.c_int64), ("stash_bytes", C.c_int64),
                ("norm_floats", C.c_int64), ("embed_offset_floats", C.c_int64), ("hidden_padded", C.c_int64)]


class HexAraDesc(C.Structure):
    _fields_ = [("hidden", C.c_int32), ("hidden_layers", C.c_int32), ("policy_layers", C.c_int32),
                ("value_layers", C.c_int32), ("in_channels", C.c_int32), ("swap_allowed", C.c_int32)]


class HexAraSizes(C.Structure):
    _fields_ = [("ws_bytes", C.c_int64), ("max_policy_entries", C.c_int64), ("hidden_padded", C.c_int64),
                ("stash_bytes", C.c_int64), ("bwd_ws_bytes", C.c_int64)]


class HexCsrLayout(C.Structure):
    _fields_ = [("status", C.c_int64), ("ntiles", C.c_int64), ("gptr", C.c_int64), ("rowptr", C.c_int64),
                ("col", C.c_int64), ("tile_g0", C.c_int64), ("scratch", C.c_int64), ("total_words", C.c_int64)]


EXPORTS = ["hexgnn_abi_version", "hexgnn_last_error", "hexgnn_param_count", "hexgnn_param_offsets",
           "hexgnn_sizes", "hexgnn_csr_layout", "hexgnn_build_csr", "hexgnn_forward", "hexgnn_backward",
           "hexgnn_profile_enable", "hexgnn_profile_read", "hexgnn_selftest_mma", "hexgnn_debug_phases", "hexgnn_stash_embed", "hexgnn_td_loss", "hexgnn_train_step", "hexgnn_adam_step", "hexgnn_adam_ws_bytes", "hexgnn_collate", "hexgnn_collate_ws_bytes",
           "hexgnn_mask_argmax", "hexgnn_td_target", "hexgnn_forward_big", "hexgnn_allreduce_p2p",
           "hexgnn_noisy_compose", "hexgnn_noisy_grads", "hexgnn_backward_dx",
           "hexgnn_lw_sizes", "hexgnn_forward_lw", "hexgnn_backward_lw", "hexgnn_forward_static", "hexgnn_allreduce_mc",
           "hexgnn_hexara_param_count", "hexgnn_hexara_param_offsets", "hexgnn_hexara_sizes", "hexgnn_hexara_forward", "hexgnn_hexara_backward"]

_lib = None


class HexGnnError(RuntimeError):
    pass


def lib():
    """Load libhexgnn.so once.  Raises if it has not been built (python -m gnn_hex_b200.build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise HexGnnError(f"{LIB_PATH} is missing: build it with `python -m gnn_hex_b200.build` "
                          "(nvcc, sm_100a).  There is no fallback path.")
    l = C.CDLL(LIB_PATH)
    p, i32, i64, sz = C.c_void_p, C.c_int32, C.c_int64, C.c_size_t
    D = C.POINTER(HexModelDesc)
    l.hexgnn_abi_version.restype = C.c_int
    l.hexgnn_abi_version.argtypes = []
    l.hexgnn_last_error.restype = C.c_char_p
    l.hexgnn_last_error.argtypes = []
    l.hexgnn_param_count.restype = i64
    l.hexgnn_param_count.argtypes = [D]
    l.hexgnn_param_offsets.restype = i32
    l.hexgnn_param_offsets.argtypes = [D, C.POINTER(i64), i32]
    l.hexgnn_sizes.restype = C.c_int
    l.hexgnn_sizes.argtypes = [D, i64, i64, i64, i32, C.POINTER(HexSizes)]
    l.hexgnn_csr_layout.restype = C.c_int
    l.hexgnn_csr_layout.argtypes = [i64, i64, i64, C.POINTER(HexCsrLayout)]
    l.hexgnn_build_csr.restype = C.c_int
    l.hexgnn_build_csr.argtypes = [p, i64, p, p, i64, i64, i32, p, sz, p]
    l.hexgnn_forward.restype = C.c_int
    l.hexgnn_forward.argtypes = [D, p, p, i64, i64, i64, p, i32, i32, p, p, p, p, sz, p]
    l.hexgnn_backward.restype = C.c_int
    l.hexgnn_backward.argtypes = [D, p, p, i64, i64, i64, p, i32, i32, p, p, p, p, p, p, sz, p]
    l.hexgnn_backward_dx.restype = C.c_int
    l.hexgnn_backward_dx.argtypes = [D, p, p, i64, i64, i64, p, i32, i32, p, p, p, p, p, p, p, sz, p]
    l.hexgnn_selftest_mma.restype = C.c_int
    l.hexgnn_selftest_mma.argtypes = [p, p, p, i32, i32, i32, p]
    l.hexgnn_stash_embed.restype = C.c_int
    l.hexgnn_stash_embed.argtypes = [D, i64, i64, i32, C.POINTER(i64), C.POINTER(i64)]
    l.hexgnn_train_step.restype = C.c_int
    l.hexgnn_train_step.argtypes = [p, p, p, i64, i64, i64, p, i32, i32, p, p, p, p, i32, C.c_float, p, p, p, p, p,
                                    p, C.c_size_t, p]
    l.hexgnn_collate_ws_bytes.restype = i64
    l.hexgnn_collate_ws_bytes.argtypes = [i64]
    l.hexgnn_collate.restype = C.c_int
    l.hexgnn_collate.argtypes = [p, p, p, p, i64, p, p, p, p, p, C.c_size_t, p, C.c_size_t, p]
    l.hexgnn_adam_ws_bytes.restype = i64
    l.hexgnn_adam_ws_bytes.argtypes = []
    l.hexgnn_adam_step.restype = C.c_int
    l.hexgnn_adam_step.argtypes = [p, p, p, p, p, p, p, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, i32, p, p,
                                   C.c_size_t, p]
    l.hexgnn_noisy_compose.restype = C.c_int
    l.hexgnn_noisy_compose.argtypes = [D, p, p, p]
    l.hexgnn_noisy_grads.restype = C.c_int
    l.hexgnn_noisy_grads.argtypes = [D, p, p, p, i32, p]
    l.hexgnn_td_loss.restype = C.c_int
    l.hexgnn_td_loss.argtypes = [p, p, p, p, p, i64, i64, i32, C.c_float, p, p, p, p]
    l.hexgnn_forward_big.restype = C.c_int
    l.hexgnn_forward_big.argtypes = [D, p, p, i64, i64, i64, p, i32, i32, i32, p, p, p, sz, p]
    l.hexgnn_allreduce_p2p.restype = C.c_int
    l.hexgnn_allreduce_p2p.argtypes = [p, i32, i64, p, p]
    l.hexgnn_allreduce_mc.restype = C.c_int
    l.hexgnn_allreduce_mc.argtypes = [p, p, i32, i32, i64, i32, p, p]
    l.hexgnn_mask_argmax.restype = C.c_int
    l.hexgnn_mask_argmax.argtypes = [p, p, i64, i64, i32, C.c_float, p, p, p]
    l.hexgnn_td_target.restype = C.c_int
    l.hexgnn_td_target.argtypes = [p, p, p, p, C.c_float, i64, i64, p, p]
    l.hexgnn_lw_sizes.restype = C.c_int
    l.hexgnn_lw_sizes.argtypes = [D, i64, i64, i64, i32, C.POINTER(HexLwSizes)]
    l.hexgnn_forward_lw.restype = C.c_int
    l.hexgnn_forward_lw.argtypes = [D, p, p, p, i64, i64, i64, p, i32, p, p, p, p, sz, p]
    l.hexgnn_backward_lw.restype = C.c_int
    l.hexgnn_backward_lw.argtypes = [D, p, p, p, i64, i64, i64, p, i32, p, p, p, p, p, p, p, sz, p]
    A = C.POINTER(HexAraDesc)
    l.hexgnn_hexara_param_count.restype = C.c_int64
    l.hexgnn_hexara_param_count.argtypes = [A]
    l.hexgnn_hexara_param_offsets.restype = C.c_int32
    l.hexgnn_hexara_param_offsets.argtypes = [A, C.POINTER(i64), i32]
    l.hexgnn_hexara_sizes.restype = C.c_int
    l.hexgnn_hexara_sizes.argtypes = [A, i64, i64, i64, C.POINTER(HexAraSizes)]
    l.hexgnn_hexara_forward.restype = C.c_int
    l.hexgnn_hexara_forward.argtypes = [A, p, p, i64, i64, i64, p, p, p, p, p, p, p, p, sz, p]
    l.hexgnn_hexara_backward.restype = C.c_int
    l.hexgnn_hexara_backward.argtypes = [A, p, p, i64, i64, i64, p, p, p, p, p, p, p, p, sz, p]
    l.hexgnn_forward_static.restype = C.c_int
    l.hexgnn_forward_static.argtypes = [D, p, p, i64, i64, i64, p, i32, i32, p, p, p]
    l.hexgnn_debug_phases.restype = C.c_int
    l.hexgnn_debug_phases.argtypes = [i32, C.POINTER(i64), i32]
    l.hexgnn_profile_enable.restype = C.c_int
    l.hexgnn_profile_enable.argtypes = [C.c_int]
    l.hexgnn_profile_read.restype = C.c_int
    l.hexgnn_profile_read.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(i64)]
    v = l.hexgnn_abi_version()
    if v != ABI_VERSION:
        raise HexGnnError(f"libhexgnn.so ABI {v} != binding ABI {ABI_VERSION}: rebuild")
    _lib = l
    return l


def check(rc, what):
    if rc != 0:
        msg = lib().hexgnn_last_error().decode(errors="replace")
        raise HexGnnError(f"{what} failed (code {rc}): {msg}")


def make_desc(hidden, num_layers, num_head_layers):
    return HexModelDesc(int(hidden), int(num_layers), int(num_head_layers), 2)


def sizes(desc, N, E, B, tile_rows):
    s = HexSizes()
    check(lib().hexgnn_sizes(C.byref(desc), N, E, B, tile_rows, C.byref(s)), "hexgnn_sizes")
    return s


def lw_sizes(desc, N, E, B, norm):
    s = HexLwSizes()
    check(lib().hexgnn_lw_sizes(C.byref(desc), N, E, B, 1 if norm else 0, C.byref(s)), "hexgnn_lw_sizes")
    return s


def csr_layout(N, E, B):
    l = HexCsrLayout()
    check(lib().hexgnn_csr_layout(N, E, B, C.byref(l)), "hexgnn_csr_layout")
    return l


def param_offsets(desc):
    n = lib().hexgnn_param_offsets(C.byref(desc), None, 0)
    if n < 0:
        check(n, "hexgnn_param_offsets")
    arr = (C.c_int64 * n)()
    lib().hexgnn_param_offsets(C.byref(desc), arr, n)
    return list(arr)


def profile_enable(on=True):
    check(lib().hexgnn_profile_enable(1 if on else 0), "hexgnn_profile_enable")


def profile_read(which):
    """(total ms, launches) of the dominant kernel (0 forward, 1 backward) since the last read."""
    ms, n = C.c_double(0.0), C.c_int64(0)
    check(lib().hexgnn_profile_read(which, C.byref(ms), C.byref(n)), "hexgnn_profile_read")
    return ms.value, n.value


def debug_phases(which):
    """[148][16] cycle totals of the last tcgen05 launch (0 forward, 1 backward); see csrc/sage_*_tc.cu."""
    buf = (C.c_int64 * (148 * 16))()
    check(lib().hexgnn_debug_phases(which, buf, 148 * 16), "hexgnn_debug_phases")
    return [list(buf[i * 16:(i + 1) * 16]) for i in range(148)]


def stash_embed(desc, N, B, flags):
    """(float offset, row pitch) of the row-major body embedding inside the forward stash."""
    off, pitch = C.c_int64(0), C.c_int64(0)
    check(lib().hexgnn_stash_embed(C.byref(desc), N, B, flags, C.byref(off), C.byref(pitch)), "hexgnn_stash_embed")
    return off.value, pitch.value
