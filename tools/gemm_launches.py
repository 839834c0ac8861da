"""Per-launch list (M, N, K, us, executed TF/s) of the GEMMs of one potrf_inv+lauum at N, grouped by shape."""
import collections, ctypes as C, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unmanned_aerial_vehicles_b200 import _lib
from unmanned_aerial_vehicles_b200.engine import DeviceGP
lib = _lib.load()
N = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
rng = np.random.default_rng(0)
X = rng.standard_normal((N, 9)); Y = rng.standard_normal((N, 3))
eng = DeviceGP(X, Y, 1e-6)
eng.lml(np.ones(9), 1.0, 0.01, True)
lib.gpk_profile_begin()
eng.lml(np.ones(9), 1.0, 0.01, True)
sec, fl, nl = C.c_double(0), C.c_double(0), C.c_int(0)
lib.gpk_profile_end(C.byref(sec), C.byref(fl), C.byref(nl))
agg = collections.OrderedDict()
mnk = (C.c_int * 3)()
for i in range(nl.value):
    s, f = C.c_double(0), C.c_double(0)
    lib.gpk_profile_launch(i, C.byref(s), C.byref(f), mnk)
    key = (mnk[0], mnk[1], mnk[2], round(f.value / (2.0 * mnk[0] * mnk[1] * mnk[2]), 2))
    a = agg.setdefault(key, [0, 0.0, 0.0]); a[0] += 1; a[1] += s.value; a[2] += f.value
print("M,N,K,exec_frac,launches,total_ms,us_each,TF/s")
for (m, n, k, fr), (c, s, f) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    if max(m, n) >= 256:
        print(f"{m},{n},{k},{fr},{c},{s*1e3:.3f},{s/c*1e6:.1f},{f/s/1e12:.2f}")
