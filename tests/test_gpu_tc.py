"""GPU (B200): the tcgen05 building blocks (TMEM-resident A, shared-memory B in both canonical layouts,
3xTF32 split) reproduce an fp64 GEMM to fp32 round-off.  Called through the C-ABI."""
import ctypes as C

import pytest
import torch

from gnn_hex_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _gemm(A, B, layout):
    N, K = B.shape[1], B.shape[0]
    D = torch.full((128, N), float("nan"), device=DEV)
    rc = _lib.lib().hexgnn_selftest_mma(C.c_void_p(A.data_ptr()), C.c_void_p(B.data_ptr()), C.c_void_p(D.data_ptr()),
                                        N, K, layout, C.c_void_p(torch.cuda.current_stream().cuda_stream))
    _lib.check(rc, "hexgnn_selftest_mma")
    torch.cuda.synchronize()
    return D


# MN-major tf32 operands exist only in the SWIZZLE_128B_BASE32B form (mode 3); MN-major with SWIZZLE_NONE or
# plain SWIZZLE_128B silently produced zeros on B200 and is not used by any kernel.
MODES = {"A tmem, B K-major": 0x00, "A tmem, B MN-major sw128/32B": 0x03, "A smem K-major, B K-major": 0x10,
         "A smem MN-major, B MN-major (sw128/32B)": 0x33, "A smem K-major, B MN-major sw128/32B": 0x13}


@pytest.mark.parametrize("N,K", [(224, 112), (112, 128), (48, 48), (96, 48), (16, 8), (256, 96), (128, 64)])
@pytest.mark.parametrize("mode", list(MODES))
def test_tcgen05_3xtf32_gemm(N, K, mode):
    layout = MODES[mode]
    if layout >= 0x10 and K * (N + 31) // 32 * 32 * 8 + 128 * K * 8 > 200 * 1024:
        pytest.skip("operand images exceed shared memory")
    g = torch.Generator().manual_seed(N * 1000 + K + layout)
    A = torch.randn(128, K, generator=g).to(DEV)
    B = torch.randn(K, N, generator=g).to(DEV)
    D = _gemm(A, B, layout)
    ref = A.double() @ B.double()
    err = float((D.double() - ref).abs().max() / ref.abs().max())
    assert err < 2e-6, err
    # structured operands: exact small integers, any layout permutation shows up as a wrong integer
    A2 = torch.zeros(128, K, device=DEV); A2[:, 0] = torch.arange(128, device=DEV)
    A2[:, K - 1] = 1.0
    B2 = torch.zeros(K, N, device=DEV); B2[0] = torch.arange(N, device=DEV); B2[K - 1] = 1000.0
    D2 = _gemm(A2, B2, layout)
    assert torch.equal(D2, A2 @ B2)
