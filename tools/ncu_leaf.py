import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unmanned_aerial_vehicles_b200 import _lib
lib = _lib.load()
A = torch.eye(128, dtype=torch.float64, device="cuda") * 2 + 0.01
Ml = torch.empty_like(A); Mu = torch.empty_like(A); info = torch.zeros(1, dtype=torch.int32, device="cuda")
P = lambda t: C.c_void_p(t.data_ptr()); st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for _ in range(3):
    A.copy_(torch.eye(128, dtype=torch.float64, device="cuda") * 2 + 0.01)
    assert lib.gpk_potrf_inv(P(A), P(Ml), P(Mu), 128, P(info), st) == 0
torch.cuda.synchronize(); print("ok", int(info[0]))
