"""Tiny driver for `ncu --set full`: one large plain NT GEMM, one LAUUM-shaped (both-upper, lower-only) GEMM
and one SYRK-shaped accumulate GEMM of the DMMA kernel at N=8192."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unmanned_aerial_vehicles_b200 import _lib  # noqa: E402

lib = _lib.load()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
P = lambda t: C.c_void_p(t.data_ptr())
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
a = torch.randn(n, n, dtype=torch.float64, device="cuda")
b = torch.randn(n, n, dtype=torch.float64, device="cuda")
c = torch.zeros(n, n, dtype=torch.float64, device="cuda")
assert lib.gpk_gemm_nt(P(a), n, P(b), n, P(c), n, n, n, n, 1.0, 0, 0, 0, 0, st) == 0
u = torch.triu(a)
assert lib.gpk_gemm_nt(P(u), n, P(u), n, P(c), n, n, n, n, 1.0, 0, 2, 2, 1, st) == 0
assert lib.gpk_gemm_nt(P(a), n, P(a), n, P(c), n, n, n, n // 2, -1.0, 1, 0, 0, 1, st) == 0
torch.cuda.synchronize()
print("ok")
