"""Per-launch GEMM breakdown of one potrf_inv+lauum at N (CUDA events around every GEMM launch)."""
import collections, ctypes as C, json, os, sys
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
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); eng.lml(np.ones(9), 1.0, 0.01, True); e1.record()
sec, fl, nl = C.c_double(0), C.c_double(0), C.c_int(0)
lib.gpk_profile_end(C.byref(sec), C.byref(fl), C.byref(nl))
print("step", e0.elapsed_time(e1), "ms; gemm", sec.value * 1e3, "ms; launches", nl.value)
agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
mnk = (C.c_int * 3)()
for i in range(nl.value):
    s, f = C.c_double(0), C.c_double(0)
    lib.gpk_profile_launch(i, C.byref(s), C.byref(f), mnk)
    key = max(mnk[0], mnk[1])
    key = 1 << int(np.ceil(np.log2(key)))
    a = agg[key]; a[0] += 1; a[1] += s.value; a[2] += f.value
for k in sorted(agg):
    n, s, f = agg[k]
    print(f"dim<={k:6d} launches {n:4d} time {s*1e3:8.2f} ms  flops {f:.3e}  {f/s/1e12:6.2f} TF/s")
cyc = (C.c_longlong * 8)()
lib.gpk_leaf_phase_cycles(cyc)
print("last leaf of the step, phase cycles [load,P1,P2,P3,I1,I2,store]:", list(cyc)[:7], "total", sum(cyc))
# leaf latency: potrf_inv on a single 128x128 block, averaged over 200 back-to-back launches
A = torch.eye(128, dtype=torch.float64, device="cuda") * 2 + 0.01
Ml = torch.empty_like(A); Mu = torch.empty_like(A); info = torch.zeros(1, dtype=torch.int32, device="cuda")
P = lambda t: C.c_void_p(t.data_ptr()); st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
A0 = A.clone()
# (A is restored before every launch: refactoring the factor in place drifts into denormals and a slow rsqrt path)
for rep in range(2):
    tot = 0.0
    for _ in range(100):
        A.copy_(A0)
        e0.record(); lib.gpk_potrf_inv(P(A), P(Ml), P(Mu), 128, P(info), st); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
print("leaf potrf_inv(128) us per launch (incl. memset, event overhead):", tot * 1e3 / 100)
cyc = (C.c_longlong * 8)()
lib.gpk_leaf_phase_cycles(cyc)
names = ["load", "P1 diag potrf", "P2 panel", "P3 trailing", "I1 diag inverse", "I2 off-diag inverse", "store", "-"]
print("leaf phase cycles:", {n: int(v) for n, v in zip(names, cyc)}, "total", sum(cyc))
