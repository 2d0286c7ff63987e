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


# layout 1 (MN-major SWIZZLE_NONE B) returned zeros on B200 and is not used by any kernel: K-major only.
@pytest.mark.parametrize("N,K,layout", [(224, 112, 0), (112, 112, 0), (112, 128, 0), (48, 48, 0), (96, 48, 0),
                                        (48, 128, 0), (16, 8, 0), (256, 96, 0)])
def test_tcgen05_3xtf32_gemm(N, K, layout):
    g = torch.Generator().manual_seed(N * 1000 + K + layout)
    A = torch.randn(128, K, generator=g).to(DEV)
    B = torch.randn(K, N, generator=g).to(DEV)
    D = _gemm(A, B, layout)
    ref = A.double() @ B.double()
    err = float((D.double() - ref).abs().max() / ref.abs().max())
    # structured operands catch layout permutations that random data would also catch, but readably
    assert err < 2e-6, err
    A2 = torch.zeros(128, K, device=DEV); A2[:, 0] = torch.arange(128, device=DEV)
    A2[:, K - 1] = 1.0
    B2 = torch.zeros(K, N, device=DEV); B2[0] = torch.arange(N, device=DEV); B2[K - 1] = 1000.0
    D2 = _gemm(A2, B2, layout)
    assert torch.equal(D2, A2 @ B2)
