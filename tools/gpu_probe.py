"""GPU-box probe: FP64 peaks (DMMA register loop, cuBLAS dgemm via torch) and phase timings of the hot path.
Writes gpurun_out/probe.json.  Not part of the product path."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unmanned_aerial_vehicles_b200 import _lib  # noqa: E402
from unmanned_aerial_vehicles_b200.engine import DeviceGP  # noqa: E402

lib = _lib.load()
P = lambda t: C.c_void_p(t.data_ptr())
ST = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)


def timeit(fn, reps=3, warm=1):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e-3)
    return min(ts), float(np.median(ts))


out = {"gpu": torch.cuda.get_device_name(0)}
sink = torch.zeros(8, dtype=torch.float64, device="cuda")
fl = C.c_double(0)
for blocks in (148, 296, 592):
    t, _ = timeit(lambda: lib.gpk_dmma_probe(blocks, 20000, P(sink), C.byref(fl), ST()))
    out[f"dmma_probe_tflops_b{blocks}"] = fl.value / t / 1e12
n = 8192
a = torch.randn(n, n, dtype=torch.float64, device="cuda"); b = torch.randn(n, n, dtype=torch.float64, device="cuda")
t, _ = timeit(lambda: torch.matmul(a, b), reps=5)
out["cublas_dgemm_8192_tflops"] = 2 * n ** 3 / t / 1e12
c = torch.empty(n, n, dtype=torch.float64, device="cuda")
t, _ = timeit(lambda: lib.gpk_gemm_nt(P(a), n, P(b), n, P(c), n, n, n, n, 1.0, 0, 0, 0, 0, ST()), reps=5)
out["gpk_gemm_nt_8192_tflops"] = 2 * n ** 3 / t / 1e12
del a, b, c
print(json.dumps(out), flush=True)

sizes = [int(s) for s in (sys.argv[1:] or ["4096", "20000"])]
for N in sizes:
    rng = np.random.default_rng(20251129)
    X = rng.standard_normal((N, 9)); w = rng.standard_normal((9, 3)) / 3
    Y = np.sin(X @ w) + 0.1 * rng.standard_normal((N, 3))
    eng = DeviceGP(X, Y, 1e-6)
    inv = np.ones(9); eng._bufs()
    r = {}
    r["gram_s"] = timeit(lambda: lib.gpk_gram(P(eng.X_d), N, 9, _lib.darr(inv), 1.0, 0.01, 1e-6, P(eng.A_d), eng.NP, 1, ST()))[0]

    def fac():
        lib.gpk_gram(P(eng.X_d), N, 9, _lib.darr(inv), 1.0, 0.01, 1e-6, P(eng.A_d), eng.NP, 1, ST())
        lib.gpk_potrf_inv(P(eng.A_d), P(eng.Mlow_d), P(eng.Mup_d), eng.NP, P(eng.info_d), ST())
    r["gram_potrf_inv_s"] = timeit(fac)[0]
    r["lauum_s"] = timeit(lambda: lib.gpk_lauum(P(eng.Mup_d), P(eng.A_d), eng.NP, ST()))[0]
    t0 = time.time(); v = eng.lml(inv, 1.0, 0.01, True); r["lml_first_wall_s"] = time.time() - t0
    r["lml_grad_eval_s"] = timeit(lambda: eng.lml(inv, 1.0, 0.01, True))[0]
    r["lml_only_eval_s"] = timeit(lambda: eng.lml(inv, 1.0, 0.01, False))[0]
    r["lml"] = v[0]
    eng.factor(inv, 1.0, 0.01)
    Xq = rng.standard_normal((16384, 9))
    t0 = time.time(); eng.predict(Xq, 0.0, 1.0, True); torch.cuda.synchronize(); t1 = time.time()
    eng.predict(Xq, 0.0, 1.0, True); torch.cuda.synchronize(); r["predict_std_pts_per_s"] = 16384 / (time.time() - t1)
    t1 = time.time(); eng.predict(Xq, 0.0, 1.0, False); torch.cuda.synchronize(); r["predict_mean_pts_per_s"] = 16384 / (time.time() - t1)
    r["eval_flops_N3"] = float(N) ** 3
    r["eval_tflops_vs_N3"] = float(N) ** 3 / r["lml_grad_eval_s"] / 1e12
    out[f"N{N}"] = r
    print(json.dumps({f"N{N}": r}), flush=True)
    del eng
    torch.cuda.empty_cache()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/probe.json", "w"), indent=1)
