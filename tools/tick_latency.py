"""Control-tick latency of the GP residual lookup at the reference's production shape (BASELINE config 1:
N=1000 rows of gp_mpc_data_20251129_221039.csv, D=10, T=6, RBF+White, normalize_y), SURVEY 8(f) row 1.

The reference's QuadrotorMPC._build_gp_residuals (src/px4/mpc.py:1490-1509) makes 25 one-row predict_residual
calls (mean+std each) per 20 ms tick.  Timed here, same fitted hyper-parameters on both sides:
  A. that loop on stock scikit-learn (host cores)            -- the reference path
  B. that loop on this repo's estimator (one device round trip per row)
  C. build_gp_residuals: ONE 25-row device call              -- the batched caller this repo adds
  D. build_gp_residuals_scenarios: 4096 scenarios x 20 steps -- one call
Data come from tests/golden/simple_gp_csv.npz (the CSV rows after the reference loader's filter).
Prints one JSON line and appends it to gpurun_out/tick_latency.jsonl."""
import json, os, sys, time, warnings
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sklearn.gaussian_process import GaussianProcessRegressor as SkGPR
from sklearn.gaussian_process.kernels import RBF, WhiteKernel
from unmanned_aerial_vehicles_b200 import GaussianProcessRegressor
from unmanned_aerial_vehicles_b200.callers import SimpleQuadrotorGP, build_gp_residuals, build_gp_residuals_scenarios

g = np.load(os.path.join(ROOT, "tests", "golden", "simple_gp_csv.npz"))
X, Y, theta = g["X"], g["Y"], g["theta"]
kern = RBF(length_scale=float(np.exp(theta[0]))) + WhiteKernel(noise_level=float(np.exp(theta[1])))
with warnings.catch_warnings():
    warnings.simplefilter("ignore")
    sk = SkGPR(kernel=kern, alpha=1e-4, normalize_y=True, optimizer=None).fit(X, Y)
    ours = GaussianProcessRegressor(kernel=kern, alpha=1e-4, normalize_y=True, optimizer=None).fit(X, Y)
gp = SimpleQuadrotorGP(); gp.gp_model = ours; gp.is_trained = True
rng = np.random.default_rng(3)
base = X[rng.integers(0, len(X), 26)]
Xg = base[:, :6].T.copy(); Ug = base[:25, 6:].T.copy()          # warm-start (6,26), (4,25)


def loop(model):
    out = np.zeros((6, 25))
    for k in range(25):
        m, s = model.predict(np.concatenate([Xg[:, k], Ug[:, k]]).reshape(1, -1), return_std=True)
        out[3:6, k] = 0.1 * (m.flatten()[3:6] / 0.02)
    return out


def timeit(fn, reps):
    fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); r = fn(); ts.append(time.perf_counter() - t0)
    return float(np.median(ts)), r


tA, DA = timeit(lambda: loop(sk), 20)
tB, DB = timeit(lambda: loop(ours), 20)
tC, DC = timeit(lambda: build_gp_residuals(gp, Xg, Ug, 0.02, 0.1), 50)
S = 4096
Xs = np.repeat(Xg[None, :, :21], S, 0) + 0.01 * rng.standard_normal((S, 6, 21)); Us = np.repeat(Ug[None, :, :20], S, 0)
tD, DD = timeit(lambda: build_gp_residuals_scenarios(gp, Xs, Us, 0.02, 0.1), 5)
out = {"what": "tick_latency", "config": "config1 shape: N=1000 D=10 T=6 iso RBF+White normalize_y", "cores": os.cpu_count(),
       "A_sklearn_25x1row_mean_std_ms": tA * 1e3, "B_b200_25x1row_mean_std_ms": tB * 1e3,
       "C_b200_batched_25rows_mean_ms": tC * 1e3, "D_b200_scenarios_81920rows_mean_ms": tD * 1e3,
       "D_rows_per_s": S * 20 / tD, "max_abs_diff_loop_vs_sklearn": float(np.max(np.abs(DA - DB))),
       "max_abs_diff_batched_vs_sklearn": float(np.max(np.abs(DA - DC)))}
print(json.dumps(out), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
open(os.path.join(ROOT, "gpurun_out", "tick_latency.jsonl"), "a").write(json.dumps(out) + "\n")
