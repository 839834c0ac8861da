import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unmanned_aerial_vehicles_b200 import _lib
lib = _lib.load(); n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
P = lambda t: C.c_void_p(t.data_ptr()); st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
a = torch.randn(n, n, dtype=torch.float64, device="cuda"); b = torch.randn(n, n, dtype=torch.float64, device="cuda"); c = torch.zeros(n, n, dtype=torch.float64, device="cuda")
best = 1e9
for i in range(6):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); lib.gpk_gemm_nt(P(a), n, P(b), n, P(c), n, n, n, n, 1.0, 0, 0, 0, 0, st); e1.record(); torch.cuda.synchronize()
    if i: best = min(best, e0.elapsed_time(e1))
print(os.environ.get("GPK_LIB", "default").split("/")[-1], n, f"{2*n**3/best/1e9:.2f} TF/s")
