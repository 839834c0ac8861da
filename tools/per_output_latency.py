"""Per-tick latency of the reference's PER-OUTPUT loaders (pretrained_gp.py:52-98, gp_enhanced_mpc.py:142-183): six
single-output ARD members with their own StandardScalers, one row per call.  Same fitted members
on every arm (N=800 rows of the reference CSV, D=10; hyper-parameters fixed):
  A. the reference loop on stock scikit-learn (host cores): scaler.transform + predict(return_std) + inverse, x6
  B. the same loop on this repo's estimator: six C-ABI calls, six stream syncs
  C. PerOutputGP.predict_batch: gpk_predict_stacked, ONE call and ONE sync for the six members
  D. C on a 25-row horizon
Prints one JSON line and appends it to gpurun_out/per_output_latency.jsonl."""
import json, os, sys, time, warnings
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sklearn.gaussian_process import GaussianProcessRegressor as SkGPR
from sklearn.gaussian_process.kernels import RBF, ConstantKernel as CK, WhiteKernel
from sklearn.preprocessing import StandardScaler
from unmanned_aerial_vehicles_b200 import GaussianProcessRegressor
from unmanned_aerial_vehicles_b200.callers import OUTPUT_NAMES, PerOutputGP

g = np.load(os.path.join(ROOT, "tests", "golden", "simple_gp_csv.npz"))
X, Y = g["X"][:800], g["Y"][:800]
sk, sX, sY = {}, {}, {}
with warnings.catch_warnings():
    warnings.simplefilter("ignore")
    for i, n in enumerate(OUTPUT_NAMES):
        sX[n], sY[n] = StandardScaler().fit(X), StandardScaler().fit(Y[:, i:i + 1])
        k = CK(1.0, constant_value_bounds="fixed") * RBF(np.linspace(0.5, 3.0, 10), (0.1, 10)) + WhiteKernel(0.05, (1e-5, 10))
        sk[n] = SkGPR(kernel=k, alpha=1e-6, optimizer=None).fit(sX[n].transform(X), sY[n].transform(Y[:, i:i + 1]).ravel())
ours = {n: GaussianProcessRegressor.from_sklearn(m) for n, m in sk.items()}
members = PerOutputGP(ours, sX, sY, names=OUTPUT_NAMES)
row = X[5:6] + 0.01
horizon = X[10:35] + 0.01


def loop(models, x):
    mean, std = np.zeros(6), np.zeros(6)
    for i, n in enumerate(OUTPUT_NAMES):
        m, s = models[n].predict(sX[n].transform(x), return_std=True)
        mean[i] = sY[n].inverse_transform(m.reshape(-1, 1)).flatten()[0]
        std[i] = abs(s[0] * sY[n].scale_[0])
    return mean, std


def timeit(fn, reps):
    fn(); fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); r = fn(); ts.append(time.perf_counter() - t0)
    return float(np.median(ts)), r


tA, (mA, sA) = timeit(lambda: loop(sk, row), 50)
tB, (mB, sB) = timeit(lambda: loop(ours, row), 200)
tC, (mC, sC) = timeit(lambda: members.predict_batch(row), 200)
tD, _ = timeit(lambda: members.predict_batch(horizon), 200)
out = {"what": "per_output_latency", "config": "6 single-output ARD members, N=800 D=10, one row per call", "cores": os.cpu_count(),
       "A_sklearn_loop_ms": tA * 1e3, "B_b200_member_loop_ms": tB * 1e3, "C_b200_stacked_one_call_ms": tC * 1e3,
       "D_b200_stacked_25rows_ms": tD * 1e3, "max_abs_diff_mean_vs_sklearn": float(np.max(np.abs(mC[0] - mA))),
       "max_abs_diff_std_vs_sklearn": float(np.max(np.abs(sC[0] - sA))),
       "stacked_equals_member_loop": bool(np.array_equal(mC[0], mB) and np.array_equal(sC[0], sB))}
print(json.dumps(out), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
open(os.path.join(ROOT, "gpurun_out", "per_output_latency.jsonl"), "a").write(json.dumps(out) + "\n")
