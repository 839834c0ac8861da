"""BASELINE config 1: the reference trainer's fit (SimpleQuadrotorGP.train_gp, src/px4/simple_gp.py:156-185) on the
rows of gp_datasets/gp_mpc_data_20251129_221039.csv (tests/golden/simple_gp_csv.npz holds them after the reference
loader's filter: N=1000, D=10, T=6), RBF(0.5)+WhiteKernel(0.1), alpha=1e-4, normalize_y=True, 1 restart,
random_state=0 -- stock scikit-learn on the host cores beside this repo's estimator on the GPU, same arrays.
Prints one JSON line and appends it to gpurun_out/config1.jsonl."""
import json, os, sys, time, warnings
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from sklearn.gaussian_process import GaussianProcessRegressor as SkGPR
from sklearn.gaussian_process.kernels import RBF, WhiteKernel
from unmanned_aerial_vehicles_b200 import GaussianProcessRegressor
g = np.load(os.path.join(ROOT, "tests", "golden", "simple_gp_csv.npz"))
X, Y = g["X"], g["Y"]


def run(cls):
    k = RBF(length_scale=0.5) + WhiteKernel(noise_level=0.1)
    m = cls(kernel=k, alpha=1e-4, normalize_y=True, n_restarts_optimizer=1, random_state=0)
    t0 = time.perf_counter()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        m.fit(X, Y)
    t_fit = time.perf_counter() - t0
    t0 = time.perf_counter()
    mean, std = m.predict(X, return_std=True)
    return m, t_fit, time.perf_counter() - t0, mean, std


GaussianProcessRegressor(kernel=RBF(0.5) + WhiteKernel(0.1), alpha=1e-4, optimizer=None).fit(X[:256], Y[:256])  # CUDA context
ours, tf_o, tp_o, m_o, s_o = run(GaussianProcessRegressor)
sk, tf_s, tp_s, m_s, s_s = run(SkGPR)
out = {"what": "config1", "N": int(X.shape[0]), "D": int(X.shape[1]), "T": int(Y.shape[1]), "cores": os.cpu_count(),
       "fit_s_sklearn_cpu": tf_s, "fit_s_b200": tf_o, "predict_1000_mean_std_s_sklearn_cpu": tp_s, "predict_1000_mean_std_s_b200": tp_o,
       "theta_sklearn": sk.kernel_.theta.tolist(), "theta_b200": ours.kernel_.theta.tolist(),
       "theta_reference_trainer_golden": g["theta"].tolist(),
       "lml_sklearn": float(sk.log_marginal_likelihood_value_), "lml_b200": float(ours.log_marginal_likelihood_value_),
       "lml_reference_trainer_golden": float(g["lml"]),
       "max_rel_mean_diff": float(np.max(np.abs(m_o - m_s) / np.maximum(np.abs(m_s), 1e-12))),
       "max_abs_std_diff": float(np.max(np.abs(s_o - s_s)))}
print(json.dumps(out), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
open(os.path.join(ROOT, "gpurun_out", "config1.jsonl"), "a").write(json.dumps(out) + "\n")
