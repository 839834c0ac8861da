"""GPU tests written after the round's last GPU session (they first run with the round-end suite).  Kept in the file
that sorts last so that, under `pytest -x`, a surprise here cannot hide the parity tests that were already green on
the device."""
import os
import pickle

import numpy as np
import pytest

from unmanned_aerial_vehicles_b200 import callers


@pytest.mark.gpu
@pytest.mark.parametrize("rows", [1, 25, 300])
def test_stacked_predict_is_bit_identical_to_the_member_loop(rows):
    """gpk_predict_stacked (one call, one sync for all per-output members) runs the same kernels as gpk_predict:
    mean and std equal the member-by-member loop bit for bit, for a control-tick row, a horizon and a batch large
    enough to take the DMMA contraction instead of the small-batch kernel."""
    from sklearn.gaussian_process.kernels import RBF, ConstantKernel as CK, WhiteKernel
    from unmanned_aerial_vehicles_b200 import GaussianProcessRegressor, gpr
    rng = np.random.default_rng(3)
    models, Xs = [], []
    for t, (n, d) in enumerate([(400, 10), (700, 10), (400, 9)]):             # members differ in N and D
        X = rng.standard_normal((n, d))
        y = np.sin(X @ rng.standard_normal(d) / 2) + 0.05 * rng.standard_normal(n)
        k = CK(1.0, constant_value_bounds="fixed") * RBF(np.linspace(0.7, 1.5, d), (0.1, 10)) + WhiteKernel(0.01, (1e-5, 10))
        models.append(GaussianProcessRegressor(kernel=k, alpha=1e-6, optimizer=None, normalize_y=bool(t % 2)).fit(X, y))
        Xs.append(np.ascontiguousarray(X[rng.integers(0, n, rows)] + 0.02 * rng.standard_normal((rows, d))))
    want = [m.predict(x, return_std=True) for m, x in zip(models, Xs)]
    from unmanned_aerial_vehicles_b200 import engine
    res = engine.predict_stacked([m._engine for m in models], Xs, [m._yscale_b for m in models], True)
    assert res is not None                                                    # this batch qualifies: no silent loop
    got = gpr.predict_stacked(models, Xs, return_std=True)
    for (gm, gs), (wm, ws) in zip(got, want):
        np.testing.assert_array_equal(gm, wm)
        np.testing.assert_array_equal(gs, ws)
    got_m = gpr.predict_stacked(models, Xs, return_std=False)
    for gm, (wm, _) in zip(got_m, want):
        np.testing.assert_array_equal(gm, wm)


@pytest.mark.gpu
def test_trainer_mirror_matches_the_reference_trainer_on_device(golden_dir, tmp_path):
    """callers.GPTrainer through libgpk.so against the golden made by the reference's own GPTrainer.train_gp_models
    (gp_trainer.py:121-205; oracle/make_golden_trainer.py): six ARD fits with 3 restarts each, held-out scores through
    the stacked predict, schema-B pickle round trip."""
    from _trainer_check import check_trainer_against_reference
    check_trainer_against_reference(golden_dir, tmp_path)


@pytest.mark.gpu
def test_offline_flow_on_device_reaches_the_reference_optimum(golden_dir, tmp_path):
    """train_gp_offline.py flow -> schema-A pickle -> evaluate_gp_offline.py flow through libgpk.so on the reference's
    CSV rows: theta* and LML of the golden made by the reference's own SimpleQuadrotorGP.train_gp (seed injected)."""
    import warnings
    from unmanned_aerial_vehicles_b200 import offline
    z = np.load(os.path.join(golden_dir, "simple_gp_csv.npz"))
    gp = callers.SimpleQuadrotorGP(max_data_points=10000)
    gp.extend(z["X"], z["Y"])
    (tmp_path / "data").mkdir()
    gp.save_dataset(str(tmp_path / "data" / "gp_mpc_data_a.csv"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        path = offline.train_offline(str(tmp_path / "data"), str(tmp_path / "models"), random_state=0)
    model = offline.load_gp_model(path)
    np.testing.assert_allclose(model.kernel_.theta, z["theta"], rtol=0, atol=1e-4)
    assert abs(model.log_marginal_likelihood_value_ - float(z["lml"])) <= 1e-6 * abs(float(z["lml"]))
    df, X_feat, R_true, X_state, U_ctrl = offline.load_dataset(tmp_path / "data" / "gp_mpc_data_a.csv")
    out = offline.evaluate_gp(model, X_feat, R_true, X_state, U_ctrl, save_prefix=tmp_path / "run")
    assert out["global"]["mse_gp"] < out["global"]["mse_nom"] and os.path.exists(tmp_path / "run_metrics.csv")
    np.testing.assert_allclose(out["per_component"]["mse_nom"], np.mean(R_true ** 2, axis=0), rtol=1e-9)
