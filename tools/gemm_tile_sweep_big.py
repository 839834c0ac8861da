"""Large triangular shapes of the recursion (n=10112 tri / lower-only, n=20096 LAUUM) for one forced tile size."""
import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unmanned_aerial_vehicles_b200 import _lib
lib = _lib.load()
P = lambda t: C.c_void_p(t.data_ptr()); st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
out = []
for n in (10112, 20096):
    a = torch.randn(n, n, dtype=torch.float64, device="cuda"); c = torch.zeros(n, n, dtype=torch.float64, device="cuda")
    u = torch.triu(a); l = torch.tril(a)
    cases = (("b_lower", (a, l, 1.0, 0, 0, 1, 0)), ("syrk_lower_only", (a, a, -1.0, 1, 0, 0, 1)), ("a_upper", (u, a, 1.0, 0, 2, 0, 0)),
             ("lauum", (u, u, 1.0, 0, 2, 2, 1))) if n < 20000 else (("lauum", (u, u, 1.0, 0, 2, 2, 1)),)
    for name, (A, B, alpha, acc, at, bt, lo) in cases:
        best = 1e9
        for i in range(3):
            e0.record(); lib.gpk_gemm_nt(P(A), n, P(B), n, P(c), n, n, n, n, alpha, acc, at, bt, lo, st); e1.record(); torch.cuda.synchronize()
            if i: best = min(best, e0.elapsed_time(e1))
        out.append(f"{n}:{name}:{best:.2f}ms")
    del a, c, u, l
print("TS=" + os.environ.get("GPK_FORCE_TS", "auto"), " ".join(out))
