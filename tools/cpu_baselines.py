"""Reference CPU path on the box's host cores at BASELINE size:
scipy.linalg.cholesky alone at N=20000, and stock scikit-learn GaussianProcessRegressor (optimizer=None) fit +
predict(return_std=True) on one query chunk.  No GPU is touched.  Appends one JSON line to gpurun_out/cpu_baselines.jsonl."""
import json, os, sys, time, warnings
import numpy as np
from scipy.linalg import cholesky
from sklearn.gaussian_process import GaussianProcessRegressor
from sklearn.gaussian_process.kernels import RBF, ConstantKernel, WhiteKernel
from threadpoolctl import threadpool_info
N = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
MQ = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
D, T = 9, 3
rng = np.random.default_rng(20251129)
X = rng.standard_normal((N, D)); w = rng.standard_normal((D, T)) / 3
Y = np.sin(X @ w) + 0.1 * rng.standard_normal((N, T))
ls = np.array([2.91, 2.21, 4.09, 4.53, 3.17, 4.25, 9.06, 3.86, 3.95])
k = ConstantKernel(1.0, constant_value_bounds="fixed") * RBF(ls) + WhiteKernel(0.00996)
t0 = time.perf_counter(); K = k(X); K[np.diag_indices_from(K)] += 1e-6; t_gram = time.perf_counter() - t0
t0 = time.perf_counter(); L = cholesky(K, lower=True, check_finite=False); t_chol = time.perf_counter() - t0
del K, L
with warnings.catch_warnings():
    warnings.simplefilter("ignore")
    t0 = time.perf_counter(); g = GaussianProcessRegressor(kernel=k, alpha=1e-6, optimizer=None).fit(X, Y); t_fit = time.perf_counter() - t0
    r = np.random.default_rng(7)
    Xq = (r.standard_normal((MQ // 20, 1, D)) + 0.05 * np.cumsum(r.standard_normal((MQ // 20, 20, D)), 1)).reshape(-1, D)
    t0 = time.perf_counter(); m, s = g.predict(Xq, return_std=True); t_pred = time.perf_counter() - t0
out = {"what": "cpu_baselines", "N": N, "cores": os.cpu_count(),
       "blas": [(d.get("internal_api"), d.get("num_threads")) for d in threadpool_info()],
       "kernel_matrix_s": t_gram, "scipy_cholesky_s": t_chol, "scipy_cholesky_tflops": N ** 3 / 3 / t_chol / 1e12,
       "sklearn_fit_optimizer_none_s": t_fit, "sklearn_predict_mean_std_rows": int(len(Xq)), "sklearn_predict_s": t_pred,
       "sklearn_predict_pts_per_s": len(Xq) / t_pred, "extrapolated_1e7_points_s": 1e7 / (len(Xq) / t_pred)}
print(json.dumps(out), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
open("gpurun_out/cpu_baselines.jsonl", "a").write(json.dumps(out) + "\n")
