"""Shared body of the GPTrainer-mirror check: the CPU suite runs it through the oracle-backed engine double
(test_estimator_host_logic.py), the GPU suite through libgpk.so (test_callers.py)."""
import pickle
import warnings

import numpy as np


def check_trainer_against_reference(golden_dir, tmp_path):
    """callers.GPTrainer on the rows the reference's own GPTrainer.train_gp_models was run on
    (oracle/make_golden_trainer.py, seed injected on both sides): same optimum, same held-out scores, schema-B pickle."""
    import os

    from unmanned_aerial_vehicles_b200 import callers

    z = np.load(os.path.join(golden_dir, "gp_trainer_loop.npz"))
    tr = callers.GPTrainer(data_dir=str(tmp_path / "data"), model_dir=str(tmp_path / "models"), random_state=0)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        stats = tr.train_gp_models(z["X"], z["Y"])
    assert list(stats) == [str(n) for n in z["names"]] == callers.OUTPUT_NAMES
    for n in stats:
        lml = float(z[f"log_marginal_likelihood_{n}"])
        assert abs(stats[n]["log_marginal_likelihood"] - lml) <= 1e-6 * abs(lml), n      # north star: LML 1e-6 rel
        np.testing.assert_allclose(tr.gp_models[n].kernel_.theta, z[f"theta_{n}"], rtol=0, atol=2e-3, err_msg=n)
        np.testing.assert_allclose(stats[n]["rmse"], float(z[f"rmse_{n}"]), rtol=1e-4, err_msg=n)
        np.testing.assert_allclose(stats[n]["mse"], float(z[f"mse_{n}"]), rtol=2e-4, err_msg=n)
        np.testing.assert_allclose(stats[n]["r2"], float(z[f"r2_{n}"]), rtol=0, atol=2e-4, err_msg=n)
    path = tr.save_models("unit")
    blob = pickle.load(open(path, "rb"))
    assert set(blob) == {"gp_models", "scalers_X", "scalers_y", "training_stats", "model_name", "creation_time"}
    pg = callers.adopt_pickle(blob)
    mean, std = pg.predict_batch(z["X"][:5])
    assert mean.shape == (5, 6) and std.shape == (5, 6) and np.isfinite(mean).all() and (std > 0).all()
    tr2 = callers.GPTrainer(data_dir=str(tmp_path / "data"), model_dir=str(tmp_path / "models"))
    assert tr2.load_models(path).keys() == stats.keys()
    return z, tr
