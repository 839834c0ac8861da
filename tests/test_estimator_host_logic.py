"""Host logic of the sklearn-surface estimator (gpr.py) on CPU: the device engine is replaced by a test double that
answers from the CPU oracle, so the restart draw, theta ordering, normalisation, output shapes, error behaviour and
pickling can be compared with stock scikit-learn without a GPU.  The double lives here (tests may use the oracle;
the product never does) -- the GPU tests in test_gpu_parity.py run the same comparisons through libgpk.so."""
import pickle
import warnings

import numpy as np
import pytest
from sklearn.gaussian_process import GaussianProcessRegressor as SkGPR
from sklearn.gaussian_process.kernels import RBF, ConstantKernel as CK, WhiteKernel

from oracle import gp_oracle as orc
from unmanned_aerial_vehicles_b200 import gpr as gpr_mod


class OracleEngine:
    """same interface as engine.DeviceGP (lml / factor / predict / alpha_host / L_host), NumPy arithmetic"""

    def __init__(self, X, y2d, jitter, device=None):
        self.X, self.y, self.jitter = np.array(X), np.array(y2d), float(jitter)
        self.N, self.D = self.X.shape
        self.T = self.y.shape[1]
        self.factored, self.a_holds_L = None, False

    def lml(self, inv_ls, c, noise, with_grad):
        self.factored, self.a_holds_L = None, False
        lml, g = orc.lml_and_grad(self.X, self.y, 1.0 / np.asarray(inv_ls), c, noise, self.jitter, eval_gradient=with_grad)
        if not np.isfinite(lml):
            return -np.inf, 0.0, np.zeros(self.D), 0.0
        if not with_grad:
            return lml, 0.0, np.zeros(self.D), 0.0
        return lml, g["c"], g["ls"], g["noise"]

    def factor(self, inv_ls, c, noise):
        try:
            self.L, self.alpha = orc.factor(self.X, self.y, 1.0 / np.asarray(inv_ls), c, noise, self.jitter)
        except np.linalg.LinAlgError:
            self.factored = None
            return 1
        self.factored, self.a_holds_L = (tuple(float(v) for v in inv_ls), float(c), float(noise)), True
        return 0

    def alpha_host(self): return self.alpha.copy()
    def L_host(self): return self.L.copy()

    def predict(self, Xq, y_mean, y_std, return_std, return_cov=False, chunk_rows=None):
        inv_ls, c, noise = self.factored
        ls = 1.0 / np.asarray(inv_ls)
        if return_cov:
            from scipy.linalg import solve_triangular
            Ks = c * orc.rbf_cross(Xq, self.X, ls)
            V = solve_triangular(self.L, Ks.T, lower=True)
            cov = c * orc.rbf_cross(Xq, Xq, ls) + noise * np.eye(len(Xq)) - V.T @ V
            return (Ks @ self.alpha) * y_std + y_mean, None, cov, 0
        m, s = orc.predict(Xq, self.X, self.L, self.alpha, ls, c, noise, y_mean, y_std, return_std)
        return m, s, None, 0


    def predict_with_metrics(self, Xq, y_mean, y_std, R_true, chunk_rows=None):
        """NumPy statement of gpk_residual_metrics' 46 sums (include/gpk.h) behind the mean-only predict"""
        mean = self.predict(Xq, y_mean, y_std, False)[0]
        t, e = R_true, R_true - mean
        raw = np.zeros(46)
        for j in range(t.shape[1]):
            raw[5 * j: 5 * j + 5] = [t[:, j].sum(), (t[:, j] ** 2).sum(), (e[:, j] ** 2).sum(),
                                     (t[:, j] ** 2 > e[:, j] ** 2).sum(), ((t[:, j] - t[:, j].mean()) ** 2).sum()]
        sn, sg = (t ** 2).sum(1), (e ** 2).sum(1)
        raw[40:46] = [sn.sum(), sg.sum(), (sn > sg).sum(), (sn < sg).sum(), (t[:, 3:6] ** 2).sum(), (e[:, 3:6] ** 2).sum()]
        return mean, raw


@pytest.fixture(autouse=True)
def _oracle_engine(monkeypatch):
    monkeypatch.setattr(gpr_mod, "DeviceGP", OracleEngine)


def _data(n, d, t, seed=0):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, d))
    Y = np.sin(X @ rng.standard_normal((d, t))) + 0.1 * rng.standard_normal((n, t))
    return X, (Y if t > 1 else Y[:, 0])


def _pair(kernel, X, Y, **kw):
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        s = SkGPR(kernel=kernel, **kw).fit(X, Y)
        g = gpr_mod.GaussianProcessRegressor(kernel=kernel, **kw).fit(X, Y)
    return s, g


@pytest.mark.parametrize("kernel,t,normalize", [
    (RBF(0.5) + WhiteKernel(0.1), 6, True),                                                       # simple_gp.py:167
    (CK(1.0, constant_value_bounds="fixed") * RBF([1.0] * 4, (0.1, 10.0)) + WhiteKernel(0.01, (1e-5, 10.0)), 1, False),  # gp_trainer.py:163
    (WhiteKernel(0.05) + CK(2.0) * RBF(1.5), 2, True),                                            # free constant, operands swapped
])
def test_fit_with_restarts_matches_sklearn(kernel, t, normalize):
    X, Y = _data(70, 4, t, seed=3)
    s, g = _pair(kernel, X, Y, alpha=1e-6, n_restarts_optimizer=2, normalize_y=normalize, random_state=0)
    np.testing.assert_allclose(g.kernel_.theta, s.kernel_.theta, rtol=0, atol=2e-4)   # same starts, same optimiser
    assert abs(g.log_marginal_likelihood_value_ - s.log_marginal_likelihood_value_) <= 1e-6 * abs(s.log_marginal_likelihood_value_)
    assert str(g.kernel_)[:12] == str(s.kernel_)[:12] and g.n_features_in_ == 4
    Xq = X[:9] + 0.1
    ms, ss = s.predict(Xq, return_std=True)
    mg, sg = g.predict(Xq, return_std=True)
    assert mg.shape == ms.shape and sg.shape == ss.shape                               # squeeze rules (_gpr.py:453,497)
    np.testing.assert_allclose(mg, ms, rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(sg, ss, rtol=1e-4, atol=1e-6)
    assert g.predict(Xq).shape == s.predict(Xq).shape


def test_fixed_theta_matches_sklearn_tightly():
    X, Y = _data(90, 5, 3, seed=1)
    k = CK(1.3) * RBF(np.linspace(0.7, 1.6, 5)) + WhiteKernel(0.03)
    s, g = _pair(k, X, Y, alpha=1e-8, optimizer=None, normalize_y=True)
    np.testing.assert_allclose(g.alpha_, s.alpha_, rtol=1e-9, atol=1e-10)
    np.testing.assert_allclose(g.L_, s.L_, rtol=0, atol=1e-12)
    th = s.kernel_.theta + 0.1
    ls_, gs_ = s.log_marginal_likelihood(th, eval_gradient=True)
    lg_, gg_ = g.log_marginal_likelihood(th, eval_gradient=True)
    assert abs(lg_ - ls_) <= 1e-10 * abs(ls_)
    np.testing.assert_allclose(gg_, gs_, rtol=1e-8, atol=1e-9)                          # theta order: c, ls[0..5), noise
    assert g.log_marginal_likelihood() == g.log_marginal_likelihood_value_
    with pytest.raises(ValueError):
        g.log_marginal_likelihood(None, eval_gradient=True)
    mc, cc = g.predict(X[:5], return_cov=True)
    ms, cs = s.predict(X[:5], return_cov=True)
    np.testing.assert_allclose(cc, cs, rtol=0, atol=1e-9)
    ys = g.sample_y(X[:4], n_samples=3, random_state=1)
    assert ys.shape == s.sample_y(X[:4], n_samples=3, random_state=1).shape


def test_error_behaviour_is_sklearns():
    X, Y = _data(40, 2, 1, seed=2)
    Xd = np.vstack([X, X[:5]])                       # duplicated rows, no noise, no jitter -> singular Gram
    Yd = np.concatenate([Y, Y[:5]])
    g = gpr_mod.GaussianProcessRegressor(kernel=RBF(1.0), alpha=0.0, optimizer=None)
    with pytest.raises(np.linalg.LinAlgError, match="not returning a positive definite matrix"):
        g.fit(Xd, Yd)
    g2 = gpr_mod.GaussianProcessRegressor(kernel=RBF(1.0, (0.1, 10)), alpha=1e-6, optimizer=None).fit(X, Y)
    g2._engine.jitter = 0.0
    g2._engine.X = np.vstack([X[:1]] * len(X))       # force a non-PD evaluation
    lml, grad = g2.log_marginal_likelihood(np.array([0.0]), eval_gradient=True)
    assert lml == -np.inf and np.all(grad == 0)      # _gpr.py:592-593: no exception
    with pytest.raises(ValueError, match="n_targets"):
        gpr_mod.GaussianProcessRegressor(kernel=RBF(1.0), n_targets=2).fit(X, Y)
    with pytest.raises(ValueError, match="features"):
        gpr_mod.GaussianProcessRegressor(kernel=RBF(1.0), alpha=1e-6, optimizer=None).fit(X, Y).predict(np.zeros((3, 5)))
    with pytest.raises(RuntimeError):
        g2.predict(X[:2], return_std=True, return_cov=True)
    with pytest.raises(NotImplementedError):
        gpr_mod.GaussianProcessRegressor(kernel=RBF(1.0), alpha=np.full(len(X), 1e-3)).fit(X, Y)
    with pytest.raises(ValueError, match="finite"):
        gpr_mod.GaussianProcessRegressor(kernel=RBF(1.0, (1e-2, np.inf)), n_restarts_optimizer=1, alpha=1e-6).fit(X, Y)


def test_pickle_holds_host_arrays_only_and_round_trips():
    X, Y = _data(50, 3, 2, seed=4)
    g = gpr_mod.GaussianProcessRegressor(kernel=RBF(0.8) + WhiteKernel(0.05), alpha=1e-6, optimizer=None,
                                         normalize_y=True).fit(X, Y)
    m0, s0 = g.predict(X[:7] + 0.2, return_std=True)
    blob = pickle.dumps({"gp_model": g, "training_count": 1})           # schema A (train_gp_offline.py:188-197)
    assert b"OracleEngine" not in blob and b"DeviceGP" not in blob
    g2 = pickle.loads(blob)["gp_model"]
    assert g2._engine is None and isinstance(g2.L_, np.ndarray) and g2.L_.shape == (50, 50)
    m1, s1 = g2.predict(X[:7] + 0.2, return_std=True)                    # device state rebuilt lazily
    np.testing.assert_array_equal(m1, m0)
    np.testing.assert_array_equal(s1, s0)
    sk = SkGPR(kernel=RBF(0.8) + WhiteKernel(0.05), alpha=1e-6, optimizer=None, normalize_y=True).fit(X, Y)
    g3 = gpr_mod.GaussianProcessRegressor.from_sklearn(pickle.loads(pickle.dumps(sk)))
    np.testing.assert_allclose(g3.predict(X[:7] + 0.2), sk.predict(X[:7] + 0.2), rtol=1e-9, atol=1e-12)


def test_reference_callers_on_the_engine_double(golden_dir, tmp_path):
    """The caller mirrors (SimpleQuadrotorGP wrapper trained on the reference CSV rows to the reference trainer's theta*,
    batched horizon == 25 one-row calls, pickle schema A, per-output schema B) with the engine double: the same bodies
    the GPU suite runs through libgpk.so."""
    import test_callers
    test_callers.test_wrapper_trains_like_reference_and_batches(golden_dir, tmp_path)
    test_callers.test_per_output_schema_b(golden_dir)


def test_sklearn_estimator_checks_pass():
    """scikit-learn's own estimator conformance suite (clone/get_params/set_params, input validation, dtype and
    shape handling, pickling, fit idempotence, n_features_in_, multi-output ...) run on this estimator: what a caller
    that swaps the import relies on without knowing it."""
    from sklearn.utils.estimator_checks import check_estimator
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        results = check_estimator(gpr_mod.GaussianProcessRegressor(), on_fail=None)
    failed = [(r["check_name"], str(r["exception"])[:200]) for r in results if r["status"] == "failed"]
    assert not failed, failed
    assert sum(r["status"] == "passed" for r in results) >= 50


@pytest.mark.parametrize("return_std", [False, True])
def test_predict_stacked_equals_the_per_model_loop(monkeypatch, return_std):
    """gpr.predict_stacked (one device call for the per-output members of schema B / C pickles): same values and
    shapes as calling predict on each member, through the stacked engine hook when the engines offer one and through
    the loop when they do not or when a member does not qualify."""
    X, Y = _data(120, 4, 3, seed=3)
    k = CK(1.0, constant_value_bounds="fixed") * RBF([1.0] * 4, (0.1, 10)) + WhiteKernel(0.05, (1e-5, 10))
    models = []
    for t in range(3):
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            models.append(gpr_mod.GaussianProcessRegressor(kernel=k, alpha=1e-6, optimizer=None,
                                                           normalize_y=bool(t % 2)).fit(X * (1 + t), Y[:, t]))
    Xs = [np.ascontiguousarray(X[:17] * (1 + t) + 0.01) for t in range(3)]
    want = [m.predict(x, return_std=return_std) for m, x in zip(models, Xs)]

    def same(got):
        for g, w in zip(got, want):
            if return_std:
                assert g[0].shape == w[0].shape and g[1].shape == w[1].shape
                np.testing.assert_array_equal(g[0], w[0]); np.testing.assert_array_equal(g[1], w[1])
            else:
                assert g.shape == w.shape
                np.testing.assert_array_equal(g, w)

    same(gpr_mod.predict_stacked(models, Xs, return_std=return_std))          # OracleEngine has no hook: the loop
    calls = []

    def hook(engines, Xqs, scales, rs):
        calls.append(len(engines))
        return [e.predict(x, s[0], s[1], rs)[:2] + (0,) for e, x, s in zip(engines, Xqs, scales)]

    monkeypatch.setattr(OracleEngine, "predict_stacked", staticmethod(hook), raising=False)
    same(gpr_mod.predict_stacked(models, Xs, return_std=return_std))
    assert calls == [3]
    same(gpr_mod.predict_stacked(models, [x.tolist() for x in Xs], return_std=return_std))   # lists: validated loop
    assert calls == [3]
    monkeypatch.setattr(OracleEngine, "predict_stacked", staticmethod(lambda *a: None), raising=False)
    same(gpr_mod.predict_stacked(models, Xs, return_std=return_std))          # hook declines: the loop


def test_per_output_members_with_scalers_match_the_reference_loop():
    """PerOutputGP.predict_batch (pretrained_gp.py:52-98 / gp_enhanced_mpc.py:150-183): sklearn-fitted members and
    StandardScalers from a schema-B pickle; the scaler shortcut must be bit-identical to scaler.transform /
    inverse_transform, the values equal to the reference's per-output loop run with stock sklearn."""
    from sklearn.preprocessing import MinMaxScaler, StandardScaler

    from unmanned_aerial_vehicles_b200 import callers

    X, Y = _data(150, 5, 3, seed=11)
    X = X * np.array([1.0, 5.0, 0.2, 3.0, 1.0]) + np.array([0.0, 2.0, -1.0, 0.5, 10.0])
    names = ["vx_residual", "vy_residual", "vz_residual"]
    models, sX, sY = {}, {}, {}
    for i, n in enumerate(names):
        sX[n] = (StandardScaler() if i < 2 else MinMaxScaler()).fit(X)       # third member: not the shortcut
        sY[n] = StandardScaler().fit(Y[:, i:i + 1])
        k = CK(1.0, constant_value_bounds="fixed") * RBF([1.0] * 5, (0.1, 10)) + WhiteKernel(0.01, (1e-5, 10))
        models[n] = SkGPR(kernel=k, alpha=1e-6, optimizer=None).fit(sX[n].transform(X), sY[n].transform(Y[:, i:i + 1]).ravel())
    pg = callers.adopt_pickle(pickle.loads(pickle.dumps({"gp_models": models, "scalers_X": sX, "scalers_y": sY})))
    pg.names = ["x_residual"] + names
    Xq = X[:25] + 0.01
    for n in names[:2]:
        np.testing.assert_array_equal(pg._scale_in(sX[n], Xq), sX[n].transform(Xq))
        m = np.linspace(-1, 1, 25)
        np.testing.assert_array_equal(pg._scale_out(sY[n], m), sY[n].inverse_transform(m.reshape(-1, 1)).ravel())
    mean, std = pg.predict_batch(Xq)
    assert not mean[:, 0].any() and (std[:, 0] == 1e6).all()
    for i, n in enumerate(names, start=1):
        ms, ss = models[n].predict(sX[n].transform(Xq), return_std=True)
        np.testing.assert_allclose(mean[:, i], sY[n].inverse_transform(ms.reshape(-1, 1)).ravel(), rtol=1e-6, atol=1e-9)
        np.testing.assert_allclose(std[:, i], np.abs(ss * sY[n].scale_[0]), rtol=0, atol=1e-8)
    mean2, none = pg.predict_batch(Xq, return_std=False)
    assert none is None
    np.testing.assert_allclose(mean2, mean, rtol=1e-12, atol=1e-14)
    with pytest.raises(ValueError):
        pg.predict_batch(np.full((2, 5), np.nan))


from _trainer_check import check_trainer_against_reference as _check_trainer_against_reference  # noqa: E402


def test_trainer_mirror_matches_the_reference_trainer(golden_dir, tmp_path):
    z, tr = _check_trainer_against_reference(golden_dir, tmp_path)
    # vectorised npz ingest == the reference's per-sample loop (gp_trainer.py:48-103) on the same flight files
    os_ = __import__("os")
    os_.makedirs(tr.data_dir, exist_ok=True)
    for j in (0, 1):
        np.savez(os_.path.join(tr.data_dir, f"flight_data_{j}.npz"), states_prev=z[f"flight{j}_prev"],
                 controls=z[f"flight{j}_ctrl"], states_next=z[f"flight{j}_next"], dt_values=z[f"flight{j}_dt"])
    X, y = tr.load_training_data()
    order = np.lexsort(X.T[::-1])
    np.testing.assert_array_equal(X[order], z["ingest_X"])
    np.testing.assert_allclose(y[order], z["ingest_y"], rtol=0, atol=1e-15)
    Xs, ys = tr.load_training_data(max_samples=10)
    assert Xs.shape == (10, 10) and ys.shape == (10, 6)
    with pytest.raises(FileNotFoundError):
        type(tr)(data_dir=str(tmp_path / "none"), model_dir=str(tmp_path / "models")).load_training_data()


def _schema_pickles(tmp_path, drop=None):
    """schema-B and schema-C pickles with stock-sklearn members (fixed theta), as gp_trainer.py:214-221 writes them"""
    from sklearn.preprocessing import StandardScaler

    rng = np.random.default_rng(21)
    X = rng.standard_normal((90, 10)) * np.linspace(0.5, 3.0, 10) + np.linspace(-1, 1, 10)
    Y = np.stack([np.sin(X[:, i]) + 0.3 * X[:, (i + 3) % 10] for i in range(6)], 1) + 0.05 * rng.standard_normal((90, 6))
    namesB = ["x_residual", "y_residual", "z_residual", "vx_residual", "vy_residual", "vz_residual"]
    namesC = ["pos_x", "pos_y", "pos_z", "vel_vx", "vel_vy", "vel_vz"]
    models, sX, sY = {}, {}, {}
    for i, n in enumerate(namesB):
        if drop and n in drop:
            continue
        sX[n], sY[n] = StandardScaler().fit(X), StandardScaler().fit(Y[:, i:i + 1])
        k = CK(1.0, constant_value_bounds="fixed") * RBF(np.linspace(0.8, 2.0, 10) + 0.1 * i, (0.1, 10)) + WhiteKernel(0.02 + 0.01 * i, (1e-5, 10))
        models[n] = SkGPR(kernel=k, alpha=1e-6, optimizer=None).fit(sX[n].transform(X), sY[n].transform(Y[:, i:i + 1]).ravel())
    pB = tmp_path / "b.pkl"
    pickle.dump({"gp_models": models, "scalers_X": sX, "scalers_y": sY, "training_stats": {"note": 1}}, open(pB, "wb"))
    pC = tmp_path / "c.pkl"
    b2c = dict(zip(namesB, namesC))
    pickle.dump({"models": {b2c[n]: m for n, m in models.items()}, "scalers_input": {b2c[n]: s for n, s in sX.items()},
                 "scalers_output": {b2c[n]: s for n, s in sY.items()}}, open(pC, "wb"))
    return X, str(pB), str(pC), models, sX, sY, namesB


def test_pretrained_loader_equals_the_reference_class(tmp_path):
    """callers.PreTrainedGP against the reference's own PreTrainedGP (src/px4/pretrained_gp.py, imported unmodified
    when the reference tree is present -- build container only) and against its loop restated with stock sklearn."""
    import os
    import sys

    from unmanned_aerial_vehicles_b200 import callers

    X, pB, _, models, sX, sY, names = _schema_pickles(tmp_path, drop={"z_residual"})
    mine = callers.PreTrainedGP(pB)
    assert mine.is_loaded and mine.get_stats()["available_models"] == [n for n in names if n != "z_residual"]
    rows = [(X[i, :6] + 0.01, X[i, 6:] - 0.01) for i in (0, 7, 33)]
    ref = None
    if os.path.exists("/root/reference/src/px4/pretrained_gp.py"):
        sys.path.insert(0, "/root/reference/src/px4")
        try:
            import pretrained_gp
            ref = pretrained_gp.PreTrainedGP(pB)
        finally:
            sys.path.remove("/root/reference/src/px4")
    for st, ct in rows:
        m, s = mine.predict_residual(st, ct)
        x = np.concatenate([st, ct]).reshape(1, -1)
        for i, n in enumerate(names):
            if n not in models:
                assert m[i] == 0.0 and s[i] == 1e6
                continue
            ms, ss = models[n].predict(sX[n].transform(x), return_std=True)
            np.testing.assert_allclose(m[i], sY[n].inverse_transform(ms.reshape(-1, 1)).ravel()[0], rtol=1e-6, atol=1e-9)
            np.testing.assert_allclose(s[i], abs(ss[0] * sY[n].scale_[0]), rtol=0, atol=1e-8)
        if ref is not None:
            rm, rs = ref.predict_residual(st, ct)
            np.testing.assert_allclose(m, rm, rtol=1e-6, atol=1e-9)
            np.testing.assert_allclose(s, rs, rtol=0, atol=1e-8)
            assert abs(mine.get_uncertainty(st, ct) - ref.get_uncertainty(st, ct)) <= 1e-8 * 1e6
    mb, sb = mine.predict_residual_batch(np.array([r[0] for r in rows]), np.array([r[1] for r in rows]))
    np.testing.assert_allclose(mb[1], mine.predict_residual(*rows[1])[0], rtol=1e-12, atol=1e-14)
    missing = callers.PreTrainedGP(str(tmp_path / "nope.pkl"))
    m, s = missing.predict_residual(*rows[0])
    assert not missing.is_loaded and not m.any() and (s == 1e6).all()


def test_enhanced_mpc_residuals_follow_the_reference_loop(tmp_path):
    """callers.EnhancedMPCResiduals = gp_enhanced_mpc.py:142-183 (rclpy node, not importable here): 12-d state, input
    through pos_x's scaler for every member, mean only, zeros when no complete model set is loaded."""
    from unmanned_aerial_vehicles_b200 import callers

    X, _, pC, models, sX, sY, namesB = _schema_pickles(tmp_path)
    g = callers.EnhancedMPCResiduals(pC)
    assert g.gp_available
    state = np.concatenate([X[3, :3], [9.0, 9.0, 9.0], X[3, 3:6], [7.0, 7.0, 7.0]])   # attitude slots are ignored
    control = X[3, 6:]
    pos, vel = g.predict_gp_residuals(state, control)
    xin = sX["x_residual"].transform(np.concatenate([state[:3], state[6:9], control]).reshape(1, -1))
    want = [sY[n].inverse_transform(models[n].predict(xin).reshape(-1, 1)).ravel()[0] for n in namesB]
    np.testing.assert_allclose(np.concatenate([pos, vel]), want, rtol=1e-6, atol=1e-9)
    P, V = g.predict_gp_residuals_batch(np.tile(state, (4, 1)), np.tile(control, (4, 1)))
    assert P.shape == (4, 3) and V.shape == (4, 3)
    np.testing.assert_allclose(P[2], pos, rtol=1e-12, atol=1e-14)
    _, _, pC2, *_ = _schema_pickles(tmp_path, drop={"vy_residual"})
    g2 = callers.EnhancedMPCResiduals(pC2)
    assert not g2.gp_available and not np.any(g2.predict_gp_residuals(state, control))


def test_fit_from_theta_rebuilds_the_fitted_estimator_without_the_optimiser():
    """parallel.py ships only (theta*, lml*) between ranks: the rebuilt estimator must equal the fitted one, with the
    factorisation deferred until alpha_ / L_ / predict is touched (no N x N matrix crosses ranks)."""
    X, Y = _data(60, 3, 2, seed=5)
    k = CK(1.0, constant_value_bounds="fixed") * RBF([1.0] * 3, (0.1, 10.0)) + WhiteKernel(0.01, (1e-5, 10.0))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        g = gpr_mod.GaussianProcessRegressor(kernel=k, alpha=1e-6, n_restarts_optimizer=1, normalize_y=True,
                                             random_state=0).fit(X, Y)
    assert g.n_lml_evals_ > 2                                  # objective evaluations of the optimiser runs are counted
    r = gpr_mod.GaussianProcessRegressor(kernel=k, alpha=1e-6, n_restarts_optimizer=1, normalize_y=True,
                                         random_state=0).fit_from_theta(X, Y, g.kernel_.theta,
                                                                        g.log_marginal_likelihood_value_)
    assert r._engine is None and "_alpha_cache" not in r.__dict__ and r.n_lml_evals_ == 0     # nothing computed yet
    assert str(r.kernel_) == str(g.kernel_) and r.log_marginal_likelihood() == g.log_marginal_likelihood()
    np.testing.assert_array_equal(r.y_train_, g.y_train_)
    np.testing.assert_array_equal(r.alpha_, g.alpha_)         # lazy: first access factors
    np.testing.assert_array_equal(r.L_, g.L_)
    for a, b in zip(r.predict(X[:5] + 0.1, return_std=True), g.predict(X[:5] + 0.1, return_std=True)):
        np.testing.assert_array_equal(a, b)
    r2 = gpr_mod.GaussianProcessRegressor(kernel=k, alpha=1e-6, normalize_y=True).fit_from_theta(X, Y, g.kernel_.theta)
    assert r2.log_marginal_likelihood() == pytest.approx(g.log_marginal_likelihood(), rel=1e-9)   # lml=None evaluates it
    back = pickle.loads(pickle.dumps(r))
    np.testing.assert_array_equal(back.alpha_, g.alpha_)


@pytest.mark.parametrize("n_targets", [None, 3])
def test_prior_covariance_matches_sklearn(n_targets):
    """_gpr.py:417-443: predict before fit answers from the prior, including return_cov"""
    k = CK(2.0) * RBF([0.5, 1.5]) + WhiteKernel(0.1)
    Xq = np.random.default_rng(2).standard_normal((7, 2))
    ms, cs = SkGPR(kernel=k, n_targets=n_targets).predict(Xq, return_cov=True)
    mg, cg = gpr_mod.GaussianProcessRegressor(kernel=k, n_targets=n_targets).predict(Xq, return_cov=True)
    assert mg.shape == ms.shape and cg.shape == cs.shape
    np.testing.assert_allclose(cg, cs, rtol=0, atol=1e-15)


def test_evaluate_gp_from_device_side_sums_equals_the_host_route(tmp_path):
    """offline.evaluate_gp takes the estimator's predict_with_residual_metrics (sums reduced behind the batched
    predict) when the model has the reference's six components; the dict and the metrics table must equal what the
    host route (the reference's own arithmetic, evaluate_gp_offline.py:225-312) computes from the same predictions."""
    from unmanned_aerial_vehicles_b200 import offline
    X, Y = _data(150, 10, 6, seed=9)
    Y = 0.05 * Y
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        g = gpr_mod.GaussianProcessRegressor(kernel=RBF(0.5) + WhiteKernel(0.1), alpha=1e-4, normalize_y=True,
                                             optimizer=None).fit(X, Y)
    Xe = X + 0.05
    Re = Y + 0.01 * np.random.default_rng(1).standard_normal(Y.shape)
    fused = offline.evaluate_gp(g, Xe, Re, Xe[:, :6], Xe[:, 6:], save_prefix=tmp_path / "dev")

    class HostOnly:                                   # same predictions, no fused hook -> the host route
        def predict(self, Xq): return g.predict(Xq)
    host = offline.evaluate_gp(HostOnly(), Xe, Re, Xe[:, :6], Xe[:, 6:], save_prefix=tmp_path / "host")
    for blk in ("global", "acc_only", "fractions"):
        for k, v in host[blk].items():
            assert fused[blk][k] == pytest.approx(v, rel=1e-9, abs=1e-12), (blk, k)
    assert list(fused["per_component"].columns) == list(host["per_component"].columns)
    for col in host["per_component"].columns[1:]:
        np.testing.assert_allclose(fused["per_component"][col], host["per_component"][col], rtol=1e-9, atol=1e-12)
    assert (tmp_path / "dev_metrics.csv").read_text().splitlines()[0] == (tmp_path / "host_metrics.csv").read_text().splitlines()[0]
    mean, sums = g.predict_with_residual_metrics(Xe, Re)
    np.testing.assert_array_equal(mean, g.predict(Xe))
    assert sums["n"] == 150 and len(sums["component"]) == 6 and len(sums["rows"]) == 6
