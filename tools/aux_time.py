"""CUDA-event timings of the non-GEMM kernels at N=20000 through the C ABI (Gram, cross-Gram, predict mean)."""
import ctypes as C, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unmanned_aerial_vehicles_b200 import _lib
from unmanned_aerial_vehicles_b200.engine import DeviceGP
lib = _lib.load()
rng = np.random.default_rng(20251129)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
X = rng.standard_normal((N, 9)); Y = np.sin(X @ (rng.standard_normal((9, 3)) / 3)) + 0.1 * rng.standard_normal((N, 3))
eng = DeviceGP(X, Y, 1e-6); eng._bufs()
P = lambda t: C.c_void_p(t.data_ptr()); st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
inv = _lib.darr(np.ones(9))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
def t(fn, reps=5):
    fn(); best = 1e9
    for _ in range(reps):
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1))
    return best
print("gram lower N=%d: %.3f ms" % (N, t(lambda: lib.gpk_gram(P(eng.X_d), N, 9, inv, 1.0, 0.01, 1e-6, P(eng.A_d), eng.NP, 1, st))))
Xq = torch.as_tensor(X[:8192] + 0.05, device="cuda")
Ks = torch.empty((8192, eng.NP), dtype=torch.float64, device="cuda")
print("cross_gram 8192 x N: %.3f ms" % t(lambda: lib.gpk_cross_gram(P(Xq), 8192, P(eng.X_d), N, 9, inv, 1.0, P(Ks), 8192, eng.NP, st)))
ym = torch.zeros(3, dtype=torch.float64, device="cuda"); ys = torch.ones(3, dtype=torch.float64, device="cuda")
mean = torch.empty((8192, 3), dtype=torch.float64, device="cuda")
for m in (1, 25, 1024, 8192):
    print("predict_mean %d rows: %.3f ms" % (m, t(lambda: lib.gpk_predict_mean(P(Ks), m, N, eng.NP, P(eng.alpha_d), 3, P(ym), P(ys), P(mean), st))))
assert eng.factor(np.ones(9), 1.0, 0.01) == 0
for m in (1, 25, 8192):
    x = X[:m] + 0.05
    eng.predict(x, 0.0, 1.0, True)
    import time; t0 = time.perf_counter()
    for _ in range(5): eng.predict(x, 0.0, 1.0, True)
    print("engine.predict(mean+std) %d rows: %.3f ms" % (m, (time.perf_counter() - t0) / 5 * 1e3))
