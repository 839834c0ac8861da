"""Mean-only predict (return_std=False: evaluate_gp_offline.py:198, gp_enhanced_mpc.py:165,175) at BASELINE size:
the fused kernel (no cross-Gram in HBM) against the two-pass route it replaces, and its own roofline -- FP64 issue:
per (query, training row) pair 2*DP distance operations + ~20 for exp + TT accumulations = 44 FP64 lane-operations
at D=9 (DP=10), T=3 (TT=4), counted in the SASS of gpk_predict_mean_fused_kernel<10,4>.

    python tools/mean_predict_run.py [--n 20000] [--rows 81920]        -> JSON line, appended to gpurun_out/mean_predict.jsonl
"""
import argparse, ctypes as C, json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unmanned_aerial_vehicles_b200 import _lib
from unmanned_aerial_vehicles_b200.engine import DeviceGP

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=20000)
ap.add_argument("--rows", type=int, default=81920)
ap.add_argument("--d", type=int, default=9)
ap.add_argument("--t", type=int, default=3)
a = ap.parse_args()
lib = _lib.load()
rng = np.random.default_rng(20251129)
N, D, T, M = a.n, a.d, a.t, a.rows
X = rng.standard_normal((N, D)); Y = np.sin(X @ (rng.standard_normal((D, T)) / 3)) + 0.1 * rng.standard_normal((N, T))
eng = DeviceGP(X, Y, 1e-6)
assert eng.factor(np.ones(D), 1.0, 0.01) == 0
rq = np.random.default_rng(7)
Xq = (rq.standard_normal((M // 20, 1, D)) + 0.05 * np.cumsum(rq.standard_normal((M // 20, 20, D)), 1)).reshape(-1, D)
M = Xq.shape[0]                                          # whole 20-step trajectories
P = lambda t: C.c_void_p(t.data_ptr()); st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
inv = _lib.darr(np.ones(D))
ym = torch.zeros(T, dtype=torch.float64, device="cuda"); ys = torch.ones(T, dtype=torch.float64, device="cuda")


def best_ms(fn, reps=5):
    fn(); b = 1e9
    for _ in range(reps):
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); b = min(b, e0.elapsed_time(e1))
    return b


out = {"what": "mean_only_predict", "N": N, "D": D, "T": T, "rows": M}
# device-resident: the fused kernel alone
MP = -(-M // 128) * 128
xq_d = torch.zeros((MP, D), dtype=torch.float64, device="cuda"); xq_d[:M] = torch.as_tensor(Xq, device="cuda")
nws = int(lib.gpk_predict_mean_workspace_doubles(MP, N, D, T))
ws = torch.empty(nws, dtype=torch.float64, device="cuda")
mean_d = torch.empty((MP, T), dtype=torch.float64, device="cuda")
fused = lambda: _lib.check(lib.gpk_predict_mean_fused(P(xq_d), M, MP, P(eng.X_d), N, D, inv, 1.0, P(eng.alpha_d), T,
                                                      P(ym), P(ys), P(ws), P(mean_d), st), "fused")
ms = best_ms(fused)
pairs = float(M) * N
DP = 4 if D <= 4 else 8 if D <= 8 else 10 if D <= 10 else 12 if D <= 12 else 16 if D <= 16 else 32
TT = 4 if T <= 4 else 8
ops_per_pair = 2 * DP + 20 + TT
clk = torch.cuda.clock_rate() * 1e3 if hasattr(torch.cuda, "clock_rate") else 1.965e9
sms = torch.cuda.get_device_properties(0).multi_processor_count
out.update(fused_kernel_ms=ms, fused_pts_per_s=M / (ms * 1e-3), fused_pairs_per_s=pairs / (ms * 1e-3),
           fp64_lane_ops_per_pair=ops_per_pair, fp64_lane_ops_per_s=pairs * ops_per_pair / (ms * 1e-3),
           fp64_issue_peak_lane_ops_per_s=sms * 64 * 1.965e9,
           frac_fp64_issue_peak=pairs * ops_per_pair / (ms * 1e-3) / (sms * 64 * 1.965e9),
           algorithmic_hbm_bytes=8.0 * M * (D + T), algorithmic_hbm_GBps=8.0 * M * (D + T) / (ms * 1e-3) / 1e9)
# the two-pass route it replaces (cross-Gram to HBM, then K* alpha), 8192-row chunks as predict used them
ch = min(8192, MP)
Ks = torch.empty((ch, eng.NP), dtype=torch.float64, device="cuda")
m2 = torch.empty((ch, T), dtype=torch.float64, device="cuda")
two = lambda: (_lib.check(lib.gpk_cross_gram(P(xq_d), ch, P(eng.X_d), N, D, inv, 1.0, P(Ks), ch, eng.NP, st), "cg"),
               _lib.check(lib.gpk_predict_mean(P(Ks), ch, N, eng.NP, P(eng.alpha_d), T, P(ym), P(ys), P(m2), st), "pm"))
ms2 = best_ms(two)
out.update(two_pass_ms_per_8192_rows=ms2, two_pass_pts_per_s=ch / (ms2 * 1e-3),
           two_pass_hbm_bytes_per_chunk=2 * 8.0 * ch * eng.NP)
fused()
torch.cuda.synchronize()
out["max_abs_diff_fused_vs_two_pass"] = float((mean_d[:ch] - m2).abs().max())
# through the engine: host arrays in, host arrays out (the user's call)
eng.predict(Xq[:4096], 0.0, 1.0, False)
t0 = time.perf_counter(); eng.predict(Xq, 0.0, 1.0, False); torch.cuda.synchronize()
out["engine_host_to_host_pts_per_s"] = M / (time.perf_counter() - t0)
for m in (1, 25):
    x = np.ascontiguousarray(Xq[:m])
    eng.predict(x, 0.0, 1.0, False)
    t0 = time.perf_counter()
    for _ in range(50):
        eng.predict(x, 0.0, 1.0, False)
    out[f"engine_mean_only_{m}_rows_ms"] = (time.perf_counter() - t0) / 50 * 1e3
print(json.dumps(out))
os.makedirs("gpurun_out", exist_ok=True)
open("gpurun_out/mean_predict.jsonl", "a").write(json.dumps(out) + "\n")
