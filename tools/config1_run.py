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


# (ii) the GPTrainer loop (src/px4/gp_trainer.py:121-205): 80/20 split (seed 42), per output: StandardScaler on X and y,
# C(1,fixed)*RBF(ARD,(0.1,10))+White(0.01,(1e-5,10)), alpha=1e-6, 3 restarts (seed injected), test RMSE / R^2, LML.
from sklearn.gaussian_process.kernels import ConstantKernel
from sklearn.metrics import mean_squared_error, r2_score
from sklearn.model_selection import train_test_split
from sklearn.preprocessing import StandardScaler
Xa, Ya = X[:, :9], Y[:, 3:6]                                   # [x..az] -> res_dvx, res_dvy, res_dvz
Xtr, Xte, Ytr, Yte = train_test_split(Xa, Ya, test_size=0.2, random_state=42)


def trainer(cls):
    res, t0 = [], time.perf_counter()
    for i in range(Ya.shape[1]):
        sx, sy = StandardScaler(), StandardScaler()
        xs, ys = sx.fit_transform(Xtr), sy.fit_transform(Ytr[:, i].reshape(-1, 1)).flatten()
        k = (ConstantKernel(1.0, constant_value_bounds="fixed") * RBF([1.0] * Xa.shape[1], (0.1, 10.0))
             + WhiteKernel(0.01, (1e-5, 1e1)))
        gp = cls(kernel=k, n_restarts_optimizer=3, alpha=1e-6, normalize_y=False, random_state=0)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            gp.fit(xs, ys)
        pred = sy.inverse_transform(gp.predict(sx.transform(Xte)).reshape(-1, 1)).flatten()
        res.append({"rmse": float(np.sqrt(mean_squared_error(Yte[:, i], pred))), "r2": float(r2_score(Yte[:, i], pred)),
                    "lml": float(gp.log_marginal_likelihood()), "theta": gp.kernel_.theta.tolist()})
    return time.perf_counter() - t0, res


t_o, r_o = trainer(GaussianProcessRegressor)
t_s, r_s = trainer(SkGPR)
out["ard_trainer"] = {"N_train": int(len(Xtr)), "D": 9, "outputs": 3, "runs_per_output": 4,
                      "fit_eval_s_sklearn_cpu": t_s, "fit_eval_s_b200": t_o,
                      "max_rel_lml_diff": max(abs(a["lml"] - b["lml"]) / abs(b["lml"]) for a, b in zip(r_o, r_s)),
                      "max_abs_theta_diff": max(float(np.max(np.abs(np.array(a["theta"]) - np.array(b["theta"])))) for a, b in zip(r_o, r_s)),
                      "max_abs_rmse_diff": max(abs(a["rmse"] - b["rmse"]) for a, b in zip(r_o, r_s)),
                      "rmse_b200": [a["rmse"] for a in r_o], "r2_b200": [a["r2"] for a in r_o]}
print(json.dumps(out), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
open(os.path.join(ROOT, "gpurun_out", "config1.jsonl"), "a").write(json.dumps(out) + "\n")
