"""Generate tests/golden/gp_trainer_loop.npz by executing the REFERENCE's own offline trainer
(/root/reference/src/px4/gp_trainer.py, unmodified, read-only) -- test infrastructure, build container only:

    python oracle/make_golden_trainer.py

GPTrainer.train_gp_models (:121-205) runs on the first 250 filtered rows of gp_datasets/gp_mpc_data_20251129_221039.csv
(X = the ten state/control columns, y = the six residual columns): 80/20 split with seed 42, per component a
StandardScaler pair and C(1, fixed)*RBF([1]*10, (0.1, 10)) + White(0.01, (1e-5, 10)) with 3 restarts, scored on the
held-out rows.  The only change is random_state=0 injected into the GaussianProcessRegressor constructor, so the
restart draws are reproducible (the reference leaves them on NumPy's global stream).  Also written: a synthetic
flight_data npz pushed through GPTrainer.load_training_data (:48-103), for the vectorised ingest.
"""
import os
import sys
import tempfile
import warnings

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden import OUT, REF, load_reference_arrays  # noqa: E402


def main():
    warnings.simplefilter("ignore")
    sys.path.insert(0, os.path.join(REF, "src", "px4"))
    import gp_trainer  # the reference module, unmodified
    from sklearn.gaussian_process import GaussianProcessRegressor as SkGPR

    gp_trainer.GaussianProcessRegressor = lambda **kw: SkGPR(random_state=0, **kw)
    X, Y = load_reference_arrays()
    X, Y = X[:250], Y[:250]
    with tempfile.TemporaryDirectory() as tmp:
        tr = gp_trainer.GPTrainer(data_dir=os.path.join(tmp, "data"), model_dir=os.path.join(tmp, "models"))
        stats = tr.train_gp_models(X, Y)
        names = list(stats.keys())
        out = {"X": X, "Y": Y, "names": np.array(names)}
        for n in names:
            g = tr.gp_models[n]
            out[f"theta_{n}"] = g.kernel_.theta
            for k in ("mse", "rmse", "r2", "log_marginal_likelihood"):
                out[f"{k}_{n}"] = np.float64(stats[n][k])
            out[f"kernel_{n}"] = np.array(stats[n]["kernel"])
        # ingest: two synthetic flight files through the reference loader
        rng = np.random.default_rng(5)
        os.makedirs(tr.data_dir)
        for j, n in enumerate((37, 21)):
            prev = rng.standard_normal((n, 6))
            ctrl = rng.standard_normal((n, 4))
            dt = 0.02 + 0.001 * rng.random(n)
            nxt = prev + 0.02 * rng.standard_normal((n, 6))
            np.savez(os.path.join(tr.data_dir, f"flight_data_{j}.npz"), states_prev=prev, controls=ctrl,
                     states_next=nxt, dt_values=dt)
            out[f"flight{j}_prev"], out[f"flight{j}_ctrl"], out[f"flight{j}_next"], out[f"flight{j}_dt"] = prev, ctrl, nxt, dt
        Xl, yl = tr.load_training_data()
        order = np.lexsort(Xl.T[::-1])                      # glob order is not defined: compare as sorted row sets
        out["ingest_X"], out["ingest_y"] = Xl[order], yl[order]
    np.savez_compressed(os.path.join(OUT, "gp_trainer_loop.npz"), **out)
    for n in names:
        print(n, stats[n]["kernel"], stats[n]["log_marginal_likelihood"], stats[n]["rmse"], stats[n]["r2"])


if __name__ == "__main__":
    main()
