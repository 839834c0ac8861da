"""Tiny driver for `ncu --set full`: three Cholesky-leaf launches on a fresh SPD 128x128 block (restored before
each launch: refactoring the factor in place drifts into denormals)."""
import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unmanned_aerial_vehicles_b200 import _lib
lib = _lib.load()
g = torch.Generator(device="cuda"); g.manual_seed(0)
B = torch.randn(128, 128, dtype=torch.float64, device="cuda", generator=g)
A0 = B @ B.T / 128 + torch.eye(128, dtype=torch.float64, device="cuda")
A = A0.clone(); Ml = torch.empty_like(A); Mu = torch.empty_like(A); info = torch.zeros(1, dtype=torch.int32, device="cuda")
P = lambda t: C.c_void_p(t.data_ptr()); st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for _ in range(3):
    A.copy_(A0)
    assert lib.gpk_potrf_inv(P(A), P(Ml), P(Mu), 128, P(info), st) == 0
torch.cuda.synchronize()
L = torch.tril(A)
print("ok", int(info[0]), float((L @ L.T - A0).abs().max()), float((Ml @ L - torch.eye(128, dtype=torch.float64, device="cuda")).abs().max()))
