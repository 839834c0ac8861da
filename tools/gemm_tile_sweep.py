"""Time the NT GEMM at the mid-size shapes of the Cholesky recursion for one forced tile size (GPK_FORCE_TS env)."""
import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unmanned_aerial_vehicles_b200 import _lib
lib = _lib.load()
P = lambda t: C.c_void_p(t.data_ptr()); st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
out = []
for n in (256, 384, 640, 1280, 2560, 5120):
    a = torch.randn(n, n, dtype=torch.float64, device="cuda"); b = torch.tril(torch.randn(n, n, dtype=torch.float64, device="cuda")); c = torch.zeros(n, n, dtype=torch.float64, device="cuda")
    for name, args, flops in (("plain", (1.0, 0, 0, 0, 0), 2.0 * n ** 3), ("b_lower", (1.0, 0, 0, 1, 0), 1.0 * n ** 3), ("syrk_lower_only", (-1.0, 1, 0, 0, 1), 1.0 * n ** 3)):
        best = 1e9
        for i in range(8):
            e0.record(); lib.gpk_gemm_nt(P(a), n, P(b), n, P(c), n, n, n, n, *args, st); e1.record(); torch.cuda.synchronize()
            if i: best = min(best, e0.elapsed_time(e1))
        out.append(f"{n}:{name}:{best*1e3:.1f}us")
print("TS=" + os.environ.get("GPK_FORCE_TS", "auto"), " ".join(out))
