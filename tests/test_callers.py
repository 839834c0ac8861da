"""Callers either side of the hot path: host logic on CPU, device paths under -m gpu."""
import os
import pickle

import numpy as np
import pytest

from unmanned_aerial_vehicles_b200 import callers


def test_metrics_reproduce_reference_artifact(golden_dir):
    """*_nom columns of gp_datasets/gp_mpc_data_20251124_225535_metrics.csv (the only numbers the reference
    committed for this path) from the CSV's residual columns."""
    z = np.load(os.path.join(golden_dir, "metrics_nominal.npz"))
    rows = callers.evaluate_residual_metrics(z["R_true"], np.zeros_like(z["R_true"]))
    np.testing.assert_allclose([r["mse_nom"] for r in rows], z["mse_nom"], rtol=1e-12)
    np.testing.assert_allclose([r["rmse_nom"] for r in rows], z["rmse_nom"], rtol=1e-12)
    np.testing.assert_allclose([r["r2_nom"] for r in rows], z["r2_nom"], rtol=1e-9)
    assert [r["component"] for r in rows] == ["dx", "dy", "dz", "dvx", "dvy", "dvz"]
    assert all(r["improvement_%"] == 0.0 and r["frac_better"] == 0.0 for r in rows)


def test_wrapper_filters_and_csv_roundtrip(tmp_path):
    gp = callers.SimpleQuadrotorGP(max_data_points=5)
    s = np.array([0, 0, 3.0, 0.5, 0, 0]); u = np.array([0.1, 0, 0, 0.0])
    nxt = gp._nominal_dynamics(s, u, 0.02) + 0.01
    gp.add_training_data(s, u, nxt)
    assert len(gp.X_train) == 1 and np.allclose(gp.Y_train[0], 0.01)
    gp.add_training_data(np.array([0, 0, 3, 6.0, 0, 0]), u, nxt)           # |v| > 5 dropped
    gp.add_training_data(s, np.array([4.0, 0, 0, 0]), nxt)                 # |a| > 3 dropped
    gp.add_training_data(s, u, nxt + 5.0)                                  # |res| > 2 dropped
    assert len(gp.X_train) == 1
    for _ in range(10):
        gp.add_training_data(s, u, nxt)
    assert len(gp.X_train) == 5                                            # deque(maxlen)
    m, v = gp.predict_residual(s, u)
    assert not m.any() and (v == 1).all()                                  # untrained contract
    p = tmp_path / "d.csv"
    gp.save_dataset(p)
    X, Y = callers.load_csv_dataset(p)
    assert X.shape == (5, 10) and Y.shape == (5, 6)
    assert gp.get_stats()["data_points"] == 5 and not gp.get_stats()["is_trained"]
    D = callers.build_gp_residuals(gp, np.zeros((6, 26)), np.zeros((4, 25)), 0.02, 0.1)
    assert D.shape == (6, 25) and not D.any()


@pytest.mark.gpu
def test_wrapper_trains_like_reference_and_batches(golden_dir, tmp_path):
    z = np.load(os.path.join(golden_dir, "simple_gp_csv.npz"))
    gp = callers.SimpleQuadrotorGP(max_data_points=10000, random_state=0)
    gp.extend(z["X"], z["Y"])
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        gp.train_gp()
    assert gp.is_trained
    np.testing.assert_allclose(gp.gp_model.kernel_.theta, z["theta"], atol=1e-4)
    m, v = gp.predict_residual(z["state"], z["control"])
    np.testing.assert_allclose(m, z["pr_mean"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(v, z["pr_var"], rtol=1e-3, atol=1e-7)
    # batched horizon == 25 sequential one-row calls (mpc.py:1490-1509)
    rng = np.random.default_rng(0)
    Xg = z["X"][:26, :6].T + 0.01 * rng.standard_normal((6, 26)); Ug = z["X"][:25, 6:].T
    D = callers.build_gp_residuals(gp, Xg, Ug, 0.02, 0.1)
    Dloop = np.zeros((6, 25))
    for k in range(25):
        mk, _ = gp.predict_residual(Xg[:, k], Ug[:, k])
        Dloop[3:6, k] = 0.1 * (mk / 0.02)[3:6]
    np.testing.assert_allclose(D, Dloop, rtol=1e-9, atol=1e-12)
    Ds = callers.build_gp_residuals_scenarios(gp, np.stack([Xg, Xg + 0.1]), np.stack([Ug, Ug]), 0.02, 0.1)
    np.testing.assert_allclose(Ds[0], D, rtol=1e-9, atol=1e-12)
    # pickle schema A round trip through load_model (train_gp_offline.py:188-197 / simple_gp.py:50-73)
    path = tmp_path / "gp_model.pkl"
    with open(path, "wb") as f:
        pickle.dump({"gp_model": gp.gp_model, "training_count": 1, "data_points_used": 1000, "is_trained": True}, f)
    gp2 = callers.SimpleQuadrotorGP()
    assert gp2.load_model(str(path))
    m2, v2 = gp2.predict_residual(z["state"], z["control"])
    np.testing.assert_allclose(m2, m, rtol=1e-12); np.testing.assert_allclose(v2, v, rtol=0, atol=1e-12)
    # evaluate_gp_offline.py:198: whole-CSV mean-only predict + metrics
    rows = callers.evaluate_residual_metrics(z["Y"], gp.gp_model.predict(z["X"]))
    assert len(rows) == 6 and all(np.isfinite(r["mse_gp"]) for r in rows)


@pytest.mark.gpu
def test_per_output_schema_b(golden_dir):
    """gp_trainer.py:152-176 + pretrained_gp.py:52-98 with sklearn-fitted members adopted onto the GPU."""
    from sklearn.gaussian_process import GaussianProcessRegressor as SkGPR
    from sklearn.gaussian_process.kernels import RBF, ConstantKernel as CK, WhiteKernel
    from sklearn.preprocessing import StandardScaler
    z = np.load(os.path.join(golden_dir, "simple_gp_csv.npz"))
    X, Y = z["X"][:300], z["Y"][:300]
    names = ["x_residual", "y_residual", "z_residual", "vx_residual", "vy_residual", "vz_residual"]
    models, sX, sY = {}, {}, {}
    for i, n in enumerate(names[3:], start=3):
        sX[n] = StandardScaler().fit(X); sY[n] = StandardScaler().fit(Y[:, i:i + 1])
        k = CK(1.0, constant_value_bounds="fixed") * RBF([1.0] * 10, (0.1, 10)) + WhiteKernel(0.01, (1e-5, 10))
        models[n] = SkGPR(kernel=k, alpha=1e-6, optimizer=None).fit(sX[n].transform(X), sY[n].transform(Y[:, i:i + 1]).ravel())
    blob = pickle.dumps({"gp_models": models, "scalers_X": sX, "scalers_y": sY, "training_stats": {}})
    pg = callers.adopt_pickle(pickle.loads(blob))
    pg.names = names
    Xq = X[:40] + 0.01
    mean, std = pg.predict_batch(Xq)
    for i, n in enumerate(names):
        if n not in models:
            assert not mean[:, i].any() and (std[:, i] == 1e6).all(); continue
        ms, ss = models[n].predict(sX[n].transform(Xq), return_std=True)
        np.testing.assert_allclose(mean[:, i], sY[n].inverse_transform(ms.reshape(-1, 1)).ravel(), rtol=1e-6, atol=1e-9)
        np.testing.assert_allclose(std[:, i], np.abs(ss * sY[n].scale_[0]), rtol=0, atol=1e-8)
    m1, s1 = pg.predict_residual(Xq[0, :6], Xq[0, 6:])
    np.testing.assert_allclose(m1, mean[0], rtol=1e-9, atol=1e-12)
