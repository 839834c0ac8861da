"""Offline tools either side of the hot path (unmanned_aerial_vehicles_b200/offline.py) on CPU: data collector npz
schema, CSV training flow, evaluation metrics.  Where the reference tree is present (build container only) its own
modules are imported unmodified and run on the same inputs; the committed goldens cover the rest.  The device engine
is the oracle-backed double of test_estimator_host_logic.py (tests may use the oracle; the product never does)."""
import os
import pickle
import sys
import warnings

import numpy as np
import pytest

from test_estimator_host_logic import OracleEngine
from unmanned_aerial_vehicles_b200 import callers, gpr as gpr_mod, offline

REF_PX4 = "/root/reference/src/px4"
HAVE_REF = os.path.exists(os.path.join(REF_PX4, "evaluate_gp_offline.py"))


@pytest.fixture(autouse=True)
def _oracle_engine(monkeypatch):
    monkeypatch.setattr(gpr_mod, "DeviceGP", OracleEngine)


def _ref(module):
    sys.path.insert(0, REF_PX4)
    try:
        return __import__(module)
    finally:
        sys.path.remove(REF_PX4)


def _flight(n, seed):
    rng = np.random.default_rng(seed)
    prev = rng.standard_normal((n, 8))                      # longer than 6: the collector keeps [:6]
    ctrl = rng.standard_normal((n, 5))
    nxt = prev + 0.02 * rng.standard_normal((n, 8))
    return prev, ctrl, nxt, 0.02 + 0.001 * rng.random(n)


def test_data_collector_writes_the_schema_the_trainer_ingests(tmp_path):
    prev, ctrl, nxt, dt = _flight(12, 0)
    col = offline.GPDataCollector(max_samples=10, data_dir=str(tmp_path / "mine"))
    taken = [col.add_sample(prev[i], ctrl[i], nxt[i], dt[i]) for i in range(12)]
    assert taken == [True] * 10 + [False] * 2 and col.get_stats()["fill_percentage"] == 100.0
    path = col.save_data()
    z = np.load(path)
    assert os.path.basename(path) == f"flight_data_{col.flight_session}.npz"
    assert set(z.files) == {"states_prev", "controls", "states_next", "timestamps", "dt_values", "flight_session",
                            "collection_time"}
    np.testing.assert_array_equal(z["states_prev"], prev[:10, :6])
    np.testing.assert_array_equal(z["controls"], ctrl[:10, :4])
    np.testing.assert_array_equal(z["dt_values"], dt[:10])
    col2 = offline.GPDataCollector(max_samples=10, data_dir=str(tmp_path / "batched"))
    assert col2.add_samples(prev, ctrl, nxt, dt) == 10 and col2.add_samples(prev, ctrl, nxt, dt) == 0
    np.testing.assert_array_equal(np.array(col2.states_next), nxt[:10, :6])
    col2.stop_collection(); col2.clear_data()
    assert col2.get_stats()["num_samples"] == 0 and not col2.add_sample(prev[0], ctrl[0], nxt[0], dt[0])
    X, y = callers.GPTrainer(data_dir=str(tmp_path / "mine"), model_dir=str(tmp_path / "m")).load_training_data()
    assert X.shape == (10, 10) and y.shape == (10, 6)
    if HAVE_REF:                                             # the reference's collector + trainer on the same samples
        rc = _ref("gp_data_collector").GPDataCollector(max_samples=10, data_dir=str(tmp_path / "ref"))
        for i in range(12):
            rc.add_sample(prev[i], ctrl[i], nxt[i], dt[i])
        zr = np.load(rc.save_data())
        assert set(zr.files) == set(z.files)
        for k in ("states_prev", "controls", "states_next", "dt_values"):
            np.testing.assert_array_equal(zr[k], z[k])
        Xr, yr = _ref("gp_trainer").GPTrainer(data_dir=str(tmp_path / "ref"), model_dir=str(tmp_path / "m2")).load_training_data()
        np.testing.assert_array_equal(X, Xr)
        np.testing.assert_allclose(y, yr, rtol=0, atol=1e-15)


def test_evaluate_gp_reproduces_the_reference_artifact_and_function(golden_dir, tmp_path):
    """`*_nom` columns of the metrics CSV the reference committed (gp_mpc_data_20251124_225535_metrics.csv) from the
    residuals of its dataset; with the reference present, the whole return structure and the written CSV against
    evaluate_gp_offline.evaluate_gp itself for the same predictor."""
    z = np.load(os.path.join(golden_dir, "metrics_nominal.npz"))
    R = z["R_true"]
    rng = np.random.default_rng(1)
    X_feat = rng.standard_normal((len(R), 10))

    class Pred:                                               # any object with .predict (evaluate_gp_offline.py:91)
        def predict(self, X):
            return 0.3 * R + 0.01 * np.sin(X[:, :6])

    out = offline.evaluate_gp(Pred(), X_feat, R, X_feat[:, :6], X_feat[:, 6:], save_prefix=tmp_path / "run")
    df = out["per_component"]
    assert list(df.columns) == ["component", "mse_nom", "mse_gp", "rmse_nom", "rmse_gp", "improvement_%", "r2_nom",
                                "r2_gp", "frac_better"]
    np.testing.assert_allclose(df["mse_nom"], z["mse_nom"], rtol=1e-10)   # (xdot_nom + R) - xdot_nom rounds with the states
    np.testing.assert_allclose(df["rmse_nom"], z["rmse_nom"], rtol=1e-10)
    np.testing.assert_allclose(df["r2_nom"], z["r2_nom"], rtol=1e-8)
    assert os.path.exists(tmp_path / "run_metrics.csv")
    assert abs(sum(out["fractions"].values()) - 1.0) < 1e-12 and out["global"]["mse_gp"] < out["global"]["mse_nom"]
    if HAVE_REF:
        import pandas as pd
        ref = _ref("evaluate_gp_offline")
        want = ref.evaluate_gp(Pred(), X_feat, R, X_feat[:, :6], X_feat[:, 6:], save_prefix=tmp_path / "ref")
        for k in ("global", "acc_only", "fractions"):
            assert out[k].keys() == want[k].keys()
            for kk in want[k]:
                np.testing.assert_allclose(out[k][kk], want[k][kk], rtol=1e-12, atol=1e-15)
        pd.testing.assert_frame_equal(df, want["per_component"], rtol=1e-12)
        pd.testing.assert_frame_equal(pd.read_csv(tmp_path / "run_metrics.csv"), pd.read_csv(tmp_path / "ref_metrics.csv"),
                                      rtol=1e-12)
        x, u = X_feat[3, :6], X_feat[3, 6:]
        np.testing.assert_array_equal(offline.f_nominal(x, u), ref.f_nominal(x, u))
        assert offline.r2_score(R[:, 0], 0.5 * R[:, 0]) == pytest.approx(ref.r2_score(R[:, 0], 0.5 * R[:, 0]), rel=1e-14)
    assert np.isnan(offline.r2_score(np.ones(5), np.zeros(5)))
    # output-width handling (evaluate_gp_offline.py:200-222)
    class Narrow:
        def predict(self, X): return np.zeros((len(X), 3))
    assert offline.evaluate_gp(Narrow(), X_feat, R, X_feat[:, :6], X_feat[:, 6:])["per_component"].shape == (6, 9)
    class Bad:
        def predict(self, X): return np.zeros(5)
    with pytest.raises(RuntimeError):
        offline.evaluate_gp(Bad(), X_feat, R, X_feat[:, :6], X_feat[:, 6:])


def test_offline_training_flow_and_evaluation_round_trip(golden_dir, tmp_path):
    """train_gp_offline.py main flow as a function: CSVs -> SimpleQuadrotorGP.train_gp -> schema-A pickle + latest
    link -> evaluate_gp_offline's loaders -> metrics.  The wrapper reaches the reference trainer's optimum on the
    reference's CSV rows (golden simple_gp_csv.npz)."""
    z = np.load(os.path.join(golden_dir, "simple_gp_csv.npz"))
    X, Y = z["X"][:150], z["Y"][:150]
    gp = callers.SimpleQuadrotorGP(max_data_points=10000)
    gp.extend(X, Y)
    (tmp_path / "data").mkdir()
    gp.save_dataset(str(tmp_path / "data" / "gp_mpc_data_a.csv"))
    open(tmp_path / "data" / "broken.csv", "w").write("x,y\n1,2\n")           # missing columns: reported and skipped
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        path = offline.train_offline(str(tmp_path / "data"), str(tmp_path / "models"), model_name="unit", random_state=0)
    assert path and os.path.islink(tmp_path / "models" / "gp_model_latest.pkl")
    blob = pickle.load(open(tmp_path / "models" / "gp_model_latest.pkl", "rb"))
    assert set(blob) == {"gp_model", "training_count", "data_points_used", "timestamp", "is_trained"}
    assert blob["data_points_used"] == 150 and blob["is_trained"]
    model = offline.load_gp_model(path)
    df, X_feat, R_true, X_state, U_ctrl = offline.load_dataset(tmp_path / "data" / "gp_mpc_data_a.csv")
    np.testing.assert_allclose(X_feat, X, rtol=1e-11); np.testing.assert_allclose(R_true, Y, rtol=1e-11)   # read_csv is not round-trip exact
    out = offline.evaluate_gp(model, X_feat, R_true, X_state, U_ctrl)
    assert out["global"]["mse_gp"] < out["global"]["mse_nom"]
    if HAVE_REF:                                             # the reference's own wrapper, seed injected, same rows
        from sklearn.gaussian_process import GaussianProcessRegressor as SkGPR
        sg = _ref("simple_gp")
        sg.GaussianProcessRegressor = lambda **kw: SkGPR(random_state=0, **kw)
        rg = sg.SimpleQuadrotorGP(max_data_points=10000)
        for xi, yi in zip(X, Y):
            rg.X_train.append(xi); rg.Y_train.append(yi)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            rg.train_gp()
        np.testing.assert_allclose(model.kernel_.theta, rg.gp_model.kernel_.theta, rtol=0, atol=1e-4)
        lml = rg.gp_model.log_marginal_likelihood_value_
        assert abs(model.log_marginal_likelihood_value_ - lml) <= 1e-6 * abs(lml)
        np.testing.assert_allclose(model.predict(X_feat[:20]), rg.gp_model.predict(X_feat[:20]), rtol=1e-5, atol=1e-8)
        # uncertainty-gated wrapper (simple_gp.py:225-260): same decisions and the same corrected step
        mine_gp = callers.SimpleQuadrotorGP(); mine_gp.gp_model, mine_gp.is_trained = model, True
        for thr in (1e-9, 10.0):
            a, b = offline.SimpleGPEnhancedMPC(mine_gp, thr), sg.SimpleGPEnhancedMPC(rg, thr)
            for i in (0, 5):
                sa = a.enhanced_dynamics_function(X[i, :6], X[i, 6:], 0.02)
                sb = b.enhanced_dynamics_function(X[i, :6], X[i, 6:], 0.02)
                np.testing.assert_allclose(sa, sb, rtol=1e-5, atol=1e-8)
            assert a.get_usage_ratio() == b.get_usage_ratio()
    assert offline.train_offline(str(tmp_path / "empty"), str(tmp_path / "models")) is None
    with pytest.raises(RuntimeError):
        p = tmp_path / "bad.pkl"; pickle.dump({"gp_model": 3}, open(p, "wb")); offline.load_gp_model(p)
    with pytest.raises(ValueError):
        offline.load_dataset(tmp_path / "data" / "broken.csv")


def test_dataset_csv_is_written_like_the_reference_writes_it(tmp_path):
    """SimpleQuadrotorGP.save_dataset (simple_gp.py:75-115): header, '%.18e' rows, append mode; byte-identical to the
    reference's own method where the reference tree is present; the row filters of add_training_data (:118-141) agree."""
    rng = np.random.default_rng(8)
    mine = callers.SimpleQuadrotorGP(max_data_points=50)
    ref = _ref("simple_gp").SimpleQuadrotorGP(max_data_points=50) if HAVE_REF else None
    for i in range(40):
        s = rng.standard_normal(6) * np.array([1, 1, 1, 2.5, 2.5, 2.5])
        u = rng.standard_normal(4) * 1.5
        nxt = s + 0.02 * np.concatenate([s[3:], u[:3]]) + rng.standard_normal(6) * 0.7
        mine.add_training_data(s, u, nxt, 0.02)
        if ref is not None:
            ref.add_training_data(s, u, nxt, 0.02)
    assert 0 < len(mine.X_train) < 40                                     # some rows fail each filter
    a = tmp_path / "sub" / "mine.csv"
    mine.save_dataset(str(a))
    mine.save_dataset(str(a), overwrite=False)                             # appended without a second header
    lines = open(a).read().splitlines()
    assert lines[0] == "x,y,z,vx,vy,vz,ax,ay,az,yaw_rate,res_dx,res_dy,res_dz,res_dvx,res_dvy,res_dvz"
    assert len(lines) == 1 + 2 * len(mine.X_train)
    X, Y = callers.load_csv_dataset(a)
    np.testing.assert_allclose(X[: len(mine.X_train)], np.array(mine.X_train), rtol=1e-14)   # pandas' fast float parser
    mine.save_dataset(str(tmp_path / "nohdr.csv"), include_header=False)
    assert len(open(tmp_path / "nohdr.csv").read().splitlines()) == len(mine.X_train)
    callers.SimpleQuadrotorGP().save_dataset(str(tmp_path / "empty.csv"))
    assert not os.path.exists(tmp_path / "empty.csv")
    if ref is not None:
        assert len(ref.X_train) == len(mine.X_train)
        np.testing.assert_array_equal(np.array(ref.Y_train), np.array(mine.Y_train))
        b = tmp_path / "ref.csv"
        ref.save_dataset(str(b)); ref.save_dataset(str(b), overwrite=False)
        assert open(a, "rb").read() == open(b, "rb").read()


def test_wrapper_error_behaviour_follows_the_reference(tmp_path):
    """simple_gp.py:50-73 / :187-201: load_model answers False for a missing file, a pickle without 'gp_model' and a
    corrupt file; predict_residual answers (zeros, ones) when the estimator raises -- but never hides a missing
    CUDA library."""
    from unmanned_aerial_vehicles_b200 import GpkUnavailable
    gp = callers.SimpleQuadrotorGP()
    assert gp.load_model(str(tmp_path / "none.pkl")) is False
    pickle.dump({"model": 1}, open(tmp_path / "nokey.pkl", "wb"))
    assert gp.load_model(str(tmp_path / "nokey.pkl")) is False
    open(tmp_path / "corrupt.pkl", "wb").write(b"not a pickle")
    assert gp.load_model(str(tmp_path / "corrupt.pkl")) is False and not gp.is_trained

    class Raises:
        def __init__(self, exc): self.exc = exc
        def predict(self, X, return_std=False): raise self.exc

    pickle.dump({"gp_model": 5, "training_count": 3}, open(tmp_path / "odd.pkl", "wb"))
    assert gp.load_model(str(tmp_path / "odd.pkl")) is True and gp.training_count == 3   # the reference does not validate either
    gp.gp_model = Raises(ValueError("X has 9 features"))
    m, v = gp.predict_residual(np.zeros(6), np.zeros(4))
    assert not m.any() and (v == 1).all() and gp.prediction_count == 0
    gp.gp_model = Raises(GpkUnavailable("libgpk.so not found"))
    with pytest.raises(GpkUnavailable):
        gp.predict_residual(np.zeros(6), np.zeros(4))
    if HAVE_REF:
        rg = _ref("simple_gp").SimpleQuadrotorGP()
        assert rg.load_model(str(tmp_path / "nokey.pkl")) is False and rg.load_model(str(tmp_path / "corrupt.pkl")) is False
        assert rg.load_model(str(tmp_path / "odd.pkl")) is True and rg.training_count == 3
        rg.gp_model = Raises(ValueError("x"))
        rm, rv = rg.predict_residual(np.zeros(6), np.zeros(4))
        np.testing.assert_array_equal(rm, m); np.testing.assert_array_equal(rv, v)


def _ref_gp_evaluation():
    """the reference's gp_evaluation.py imports matplotlib/seaborn at module level (absent here); only its model-facing
    methods are exercised, so inert stand-ins for those two packages are enough to import it unmodified"""
    from unittest import mock
    stubs = {m: mock.MagicMock() for m in ("matplotlib", "matplotlib.pyplot", "seaborn")}
    with mock.patch.dict(sys.modules, stubs):
        return _ref("gp_evaluation")


def test_model_evaluator_formats_generators_and_predictions(tmp_path):
    """offline.GPModelEvaluator (gp_evaluation.py:53-330): both pickle formats, identity scalers when a multi-output
    pickle has none, the seeded generators (array-equal to the reference's), prediction bands."""
    from test_estimator_host_logic import _schema_pickles
    X, pB, pC, models, sX, sY, namesB = _schema_pickles(tmp_path)
    ev = offline.GPModelEvaluator(pC)
    assert ev.mode == "multi" and list(ev.gp_models) == ["pos_x", "pos_y", "pos_z", "vel_vx", "vel_vy", "vel_vz"]
    data = ev.generate_physical_test_data(50)
    assert all(len(v) == 350 for v in data.values()) and list(data) == callers.FEATURE_COLS
    pred = ev.predict_on_test_data(data)
    Xt = np.column_stack([data[k] for k in callers.FEATURE_COLS])
    for c_name, b_name in zip(ev.gp_models, namesB):
        ms, ss = models[b_name].predict(sX[b_name].transform(Xt), return_std=True)
        np.testing.assert_allclose(pred[c_name]["mean"], sY[b_name].inverse_transform(ms.reshape(-1, 1)).ravel(), rtol=1e-6, atol=1e-9)
        np.testing.assert_allclose(pred[c_name]["std"], np.abs(ss * sY[b_name].scale_[0]), rtol=0, atol=1e-8)
        np.testing.assert_array_equal(pred[c_name]["upper"], pred[c_name]["mean"] + 2 * pred[c_name]["std"])
    # multi-output pickle without scalers -> identity scalers (gp_evaluation.py:92-102)
    bare = tmp_path / "bare.pkl"
    pickle.dump({"models": {"pos_x": models["x_residual"]}}, open(bare, "wb"))
    evb = offline.GPModelEvaluator(str(bare))
    ms, ss = models["x_residual"].predict(Xt, return_std=True)
    np.testing.assert_allclose(evb.predict_on_test_data(data)["pos_x"]["mean"], ms, rtol=1e-6, atol=1e-9)
    # single container (schema A), generic features when the model is not 10-dimensional
    rng = np.random.default_rng(2)
    Xs = rng.standard_normal((60, 4)); ys = np.sin(Xs[:, 0]) + 0.05 * rng.standard_normal(60)
    single = SkGPR(kernel=RBF(1.0) + WhiteKernel(0.05), alpha=1e-6, optimizer=None).fit(Xs, ys)
    pA = tmp_path / "a.pkl"
    pickle.dump({"gp_model": single, "training_count": 2, "is_trained": True}, open(pA, "wb"))
    eva = offline.GPModelEvaluator(str(pA))
    assert eva.mode == "single" and eva.n_features == 4 and eva.training_stats["training_count"] == 2
    gen = eva.generate_generic_test_data(30)
    out = eva.predict_on_test_data(gen)["output"]
    m1, s1 = single.predict(np.column_stack([gen[f"feature_{i}"] for i in range(4)]), return_std=True)
    np.testing.assert_allclose(out["mean"], m1, rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(out["std"], s1, rtol=0, atol=1e-8)
    for bad, exc in (({"other": 1}, KeyError), ({"gp_model": 3}, KeyError)):
        pickle.dump(bad, open(tmp_path / "bad.pkl", "wb"))
        with pytest.raises(exc):
            offline.GPModelEvaluator(str(tmp_path / "bad.pkl"))
    if HAVE_REF:
        ref = _ref_gp_evaluation()
        rev = ref.GPModelEvaluator(pC)
        rdata = rev.generate_physical_test_data(50)
        for k in data:
            np.testing.assert_array_equal(data[k], rdata[k])
        rpred = rev.predict_on_test_data(rdata)
        for k in pred:
            np.testing.assert_allclose(pred[k]["mean"], rpred[k]["mean"], rtol=1e-6, atol=1e-9)
            np.testing.assert_allclose(pred[k]["std"], rpred[k]["std"], rtol=0, atol=1e-8)
        ra = ref.GPModelEvaluator(str(pA))
        assert (ra.mode, ra.n_features, ra.training_stats) == (eva.mode, eva.n_features, eva.training_stats)
        rgen = ra.generate_generic_test_data(30)
        for k in gen:
            np.testing.assert_array_equal(gen[k], rgen[k])


from sklearn.gaussian_process import GaussianProcessRegressor as SkGPR  # noqa: E402
from sklearn.gaussian_process.kernels import RBF, WhiteKernel  # noqa: E402
