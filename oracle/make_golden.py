"""Generate tests/golden/*.npz by executing the REFERENCE's own code (read-only, from /root/reference) and the
sklearn 1.9.0 it depends on.  Run in the build container only (the reference does not travel to the GPU box):

    python oracle/make_golden.py

Vectors:
  simple_gp_csv.npz   -- /root/reference/src/px4/simple_gp.py SimpleQuadrotorGP.train_gp() (:156-185) on
                         gp_datasets/gp_mpc_data_20251129_221039.csv loaded with the filter of
                         src/px4/train_gp_offline.py:43-65; the only change is random_state=0 injected into the
                         GaussianProcessRegressor constructor so the restart draw is reproducible.
  ard_trainer.npz     -- the src/px4/gp_trainer.py:152-176 configuration (StandardScaler, C*RBF(ARD)+White,
                         alpha=1e-6) on the first 400 rows, D=9 live features -> res_dvx/dvy/dvz, fixed seed.
  lml_grid_*.npz      -- sklearn log_marginal_likelihood(theta, eval_gradient=True) at fixed theta grids.
"""
import os
import sys
import warnings

import numpy as np
import pandas as pd

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
CSV = os.path.join(REF, "gp_datasets", "gp_mpc_data_20251129_221039.csv")
FEAT = ["x", "y", "z", "vx", "vy", "vz", "ax", "ay", "az", "yaw_rate"]
RES = ["res_dx", "res_dy", "res_dz", "res_dvx", "res_dvy", "res_dvz"]


def load_reference_arrays():
    df = pd.read_csv(CSV)
    X = df[FEAT].to_numpy(dtype=np.float64)
    Y = df[RES].to_numpy(dtype=np.float64)
    ok = np.isfinite(X).all(1) & np.isfinite(Y).all(1) & (np.linalg.norm(Y, axis=1) < 5.0)
    return X[ok], Y[ok]


def theta_grid(kernel, n_random, seed):
    rng = np.random.default_rng(seed)
    b = kernel.bounds
    lo = np.maximum(b[:, 0], -6.0)
    hi = np.minimum(b[:, 1], 3.0)
    grid = [kernel.theta.copy()]
    grid += [rng.uniform(lo, hi) for _ in range(n_random)]
    return np.array(grid)


def main():
    warnings.simplefilter("ignore")
    os.makedirs(OUT, exist_ok=True)
    sys.path.insert(0, os.path.join(REF, "src", "px4"))
    import simple_gp  # the reference module, unmodified
    from sklearn.gaussian_process import GaussianProcessRegressor as SkGPR
    from sklearn.gaussian_process.kernels import RBF, WhiteKernel
    from sklearn.gaussian_process.kernels import ConstantKernel as C
    from sklearn.preprocessing import StandardScaler

    X, Y = load_reference_arrays()
    print("csv rows kept", X.shape, Y.shape)

    # ---- G5: the reference trainer, seed injected ------------------------------------------------
    simple_gp.GaussianProcessRegressor = lambda **kw: SkGPR(random_state=0, **kw)
    gp = simple_gp.SimpleQuadrotorGP(max_data_points=10000)
    for xi, yi in zip(X, Y):
        gp.X_train.append(xi)
        gp.Y_train.append(yi)
    gp.train_gp()
    m = gp.gp_model
    rng = np.random.default_rng(42)
    Xq = np.vstack([
        X[::40],                                          # training rows: variance ~ sigma_n^2 (worst cancellation)
        X[5::97] + 1e-3 * rng.standard_normal((len(X[5::97]), 10)),
        X.mean(0) + 3.0 * X.std(0) * rng.standard_normal((16, 10)),   # far field
    ])
    mean, std = m.predict(Xq, return_std=True)
    state = np.array([0.3, -0.2, 3.05, 0.5, 0.1, 0.0])
    control = np.array([0.1, 0.05, -2.66, 0.0])
    pr_mean, pr_var = gp.predict_residual(state, control)
    grid = theta_grid(m.kernel_, 8, 1)
    lmls, grads = zip(*[m.log_marginal_likelihood(t, eval_gradient=True) for t in grid])
    np.savez_compressed(
        os.path.join(OUT, "simple_gp_csv.npz"), X=X, Y=Y, theta=m.kernel_.theta, lml=m.log_marginal_likelihood_value_,
        y_mean=m._y_train_mean, y_std=m._y_train_std, alpha_=m.alpha_, L_diag=np.diag(m.L_), L_rowsum=m.L_.sum(1),
        Xq=Xq, mean=mean, std=std, state=state, control=control, pr_mean=pr_mean, pr_var=pr_var,
        theta_grid=grid, lml_grid=np.array(lmls), grad_grid=np.array(grads))
    print("simple_gp: kernel_", m.kernel_, "lml", m.log_marginal_likelihood_value_)

    # ---- ARD per-output trainer configuration (gp_trainer.py:152-176) -------------------------------
    Xa, Ya = X[:400, :9], Y[:400, 3:6]
    sx = StandardScaler().fit(Xa)
    Xs = sx.transform(Xa)
    out = {}
    for t in range(3):
        sy = StandardScaler().fit(Ya[:, t:t + 1])
        ys = sy.transform(Ya[:, t:t + 1]).ravel()
        k = C(1.0, constant_value_bounds="fixed") * RBF([1.0] * 9, (0.1, 10.0)) + WhiteKernel(0.01, (1e-5, 10.0))
        g = SkGPR(kernel=k, n_restarts_optimizer=3, alpha=1e-6, normalize_y=False, random_state=42 + t).fit(Xs, ys)
        mq, sq = g.predict(Xs[::7] + 0.01, return_std=True)
        grid = theta_grid(g.kernel_, 6, 10 + t)
        lmls, grads = zip(*[g.log_marginal_likelihood(th, eval_gradient=True) for th in grid])
        out.update({f"theta{t}": g.kernel_.theta, f"lml{t}": g.log_marginal_likelihood_value_, f"ys{t}": ys,
                    f"mean{t}": mq, f"std{t}": sq, f"alpha{t}": g.alpha_, f"theta_grid{t}": grid,
                    f"lml_grid{t}": np.array(lmls), f"grad_grid{t}": np.array(grads)})
        print("ard", t, g.kernel_, g.log_marginal_likelihood_value_)
    np.savez_compressed(os.path.join(OUT, "ard_trainer.npz"), Xs=Xs, Xq=Xs[::7] + 0.01, **out)

    # ---- free-constant kernel, multi-output, LML grid (exercises d/dlog c and T>1) ------------------
    rng = np.random.default_rng(3)
    Xr = rng.standard_normal((300, 5))
    Yr = np.sin(Xr @ rng.standard_normal((5, 3))) + 0.1 * rng.standard_normal((300, 3))
    k = C(2.0, (1e-2, 1e2)) * RBF([1.0, 2.0, 0.5, 1.5, 1.0], (1e-2, 1e2)) + WhiteKernel(0.05, (1e-5, 1e1))
    g = SkGPR(kernel=k, alpha=1e-8, optimizer=None, normalize_y=True).fit(Xr, Yr)
    grid = theta_grid(g.kernel_, 10, 5)
    lmls, grads = zip(*[g.log_marginal_likelihood(th, eval_gradient=True) for th in grid])
    mq, sq = g.predict(Xr[:50] * 1.1, return_std=True)
    np.savez_compressed(os.path.join(OUT, "lml_grid_const_ard.npz"), X=Xr, Y=Yr, theta_grid=grid,
                        lml_grid=np.array(lmls), grad_grid=np.array(grads), Xq=Xr[:50] * 1.1, mean=mq, std=sq,
                        alpha_=g.alpha_)
    # ---- the one metrics artefact the reference committed (evaluate_gp_offline.py:277-312) ---------------
    ds = os.path.join(REF, "gp_datasets", "gp_mpc_data_20251124_225535.csv")
    mt = pd.read_csv(os.path.join(REF, "gp_datasets", "gp_mpc_data_20251124_225535_metrics.csv"))
    R = pd.read_csv(ds)[RES].to_numpy(dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, "metrics_nominal.npz"), R_true=R, mse_nom=mt["mse_nom"].to_numpy(),
                        rmse_nom=mt["rmse_nom"].to_numpy(), r2_nom=mt["r2_nom"].to_numpy())
    print("done")


if __name__ == "__main__":
    main()
