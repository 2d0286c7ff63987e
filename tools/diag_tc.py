import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_hex_b200 import _lib
DEV = "cuda:0"
def gemm(A, B, layout):
    N, K = B.shape[1], B.shape[0]
    D = torch.full((128, N), float("nan"), device=DEV)
    rc = _lib.lib().hexgnn_selftest_mma(C.c_void_p(A.data_ptr()), C.c_void_p(B.data_ptr()), C.c_void_p(D.data_ptr()), N, K, layout, C.c_void_p(torch.cuda.current_stream().cuda_stream))
    _lib.check(rc, "selftest"); torch.cuda.synchronize(); return D
torch.set_printoptions(linewidth=250, sci_mode=False)
for (N, K) in [(32, 8), (64, 16)]:
    A = torch.zeros(128, K, device=DEV)
    for i in range(128): A[i, i % K] = 1.0
    B = (torch.arange(K, device=DEV).float()[:, None] * 100 + torch.arange(N, device=DEV).float()[None, :]).contiguous()
    for layout in (0x03, 0x33):
        D = gemm(A, B, layout)
        print(f"N={N} K={K} modes={layout:#x}: expected row k = k*100 + n")
        print(D[:K].int())
