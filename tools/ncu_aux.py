"""Driver for `ncu --set full` on the non-GEMM kernels of the path at BASELINE size (N=20000, D=9, T=3):
one LML+gradient evaluation (Gram, trimv, gradient, finalize), one 8192-row and one 1-row predict."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unmanned_aerial_vehicles_b200.engine import DeviceGP
rng = np.random.default_rng(20251129)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
X = rng.standard_normal((N, 9)); Y = np.sin(X @ (rng.standard_normal((9, 3)) / 3)) + 0.1 * rng.standard_normal((N, 3))
eng = DeviceGP(X, Y, 1e-6)
print("lml", eng.lml(np.ones(9), 1.0, 0.01, True)[0])
assert eng.factor(np.ones(9), 1.0, 0.01) == 0
m, s, _, _ = eng.predict(X[:8192] + 0.05, 0.0, 1.0, True)
m1, s1, _, _ = eng.predict(X[:1] + 0.05, 0.0, 1.0, True)
print("ok", float(m.sum()), float(s.sum()), float(s1.sum()))
