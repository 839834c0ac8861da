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
    got_m = gpr.predict_stacked(models, Xs, return_std=False)              # mean only: the fused kernel, in both
    for gm, m, x in zip(got_m, models, Xs):
        np.testing.assert_array_equal(gm, m.predict(x, return_std=False))


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


@pytest.mark.gpu
@pytest.mark.parametrize("n,NB", [(1000, 256), (3000, 1024)])
def test_distributed_factor_predicts_like_the_single_gpu_engine(n, NB):
    """DistGP.predict (queries as one more row tile of the block-cyclic factor; every product a DMMA GEMM) on one
    rank with the real CUDA ops against the single-GPU engine and the oracle: mean 1e-6 rel, std 1e-8 abs."""
    from oracle import gp_oracle as orc
    from unmanned_aerial_vehicles_b200.dist_chol import DistGP
    from unmanned_aerial_vehicles_b200.engine import DeviceGP
    rng = np.random.default_rng(1)
    X = rng.standard_normal((n, 9)); Y = np.sin(X @ rng.standard_normal((9, 3)) / 3) + 0.1 * rng.standard_normal((n, 3))
    ls = np.linspace(0.8, 2.0, 9)
    Xq = np.vstack([X[::29], X[5::31] + 1e-3 * rng.standard_normal(X[5::31].shape), 4.0 * rng.standard_normal((7, 9))])
    g = DistGP(X, Y, 1e-6, NB=NB)
    assert g.factor(1.0 / ls, 1.0, 0.01) == 0
    g.solve()
    mean, std = g.predict(Xq, return_std=True, chunk_rows=128)
    eng = DeviceGP(X, Y, 1e-6)
    assert eng.factor(1.0 / ls, 1.0, 0.01) == 0
    m_e, s_e, _, _ = eng.predict(Xq, np.zeros(3), np.ones(3), True)
    np.testing.assert_allclose(mean, m_e, rtol=1e-6, atol=1e-9)
    assert np.max(np.abs(std - s_e)) <= 1e-8
    L, a = orc.factor(X, Y, ls, 1.0, 0.01, 1e-6)
    m_o, s_o = orc.predict(Xq, X, L, a, ls, 1.0, 0.01)
    np.testing.assert_allclose(mean, m_o, rtol=1e-6, atol=1e-9)
    assert np.max(np.abs(std - s_o)) <= 1e-8
    np.testing.assert_allclose(g.predict(Xq[:5]), m_o[:5], rtol=1e-6, atol=1e-9)


@pytest.mark.gpu
@pytest.mark.parametrize("m", [1, 700, 5000])
def test_residual_metrics_reduced_on_device_match_numpy(m):
    """gpk_residual_metrics behind the fused mean-only predict (evaluate_gp_offline.py:277-312):
    the 46 sums against NumPy on the same predictions, and the evaluate_gp dict built from them against the host
    route."""
    import warnings
    from sklearn.gaussian_process.kernels import RBF, WhiteKernel
    from unmanned_aerial_vehicles_b200 import GaussianProcessRegressor, offline
    rng = np.random.default_rng(8)
    X = rng.standard_normal((900, 10))
    Y = 0.05 * np.sin(X @ rng.standard_normal((10, 6)) / 3) + 0.005 * rng.standard_normal((900, 6))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        g = GaussianProcessRegressor(kernel=RBF(0.8) + WhiteKernel(0.05), alpha=1e-4, normalize_y=True,
                                     optimizer=None).fit(X, Y)
    Xe = X[rng.integers(0, 900, m)] + 0.05 * rng.standard_normal((m, 10))
    Re = 0.05 * np.sin(Xe @ rng.standard_normal((10, 6)) / 3)
    mean, sums = g.predict_with_residual_metrics(Xe, Re)
    want_mean = g.predict(Xe)
    np.testing.assert_allclose(mean, want_mean, rtol=0, atol=1e-13)
    e = Re - mean
    for j in range(6):
        t = Re[:, j]
        want = [t.sum(), (t ** 2).sum(), (e[:, j] ** 2).sum(), float((t ** 2 > e[:, j] ** 2).sum()), ((t - t.mean()) ** 2).sum()]
        np.testing.assert_allclose(sums["component"][j], want, rtol=1e-11, atol=1e-15)
    sn, sg = (Re ** 2).sum(1), (e ** 2).sum(1)
    want = [sn.sum(), sg.sum(), float((sn > sg).sum()), float((sn < sg).sum()), (Re[:, 3:] ** 2).sum(), (e[:, 3:] ** 2).sum()]
    np.testing.assert_allclose(sums["rows"], want, rtol=1e-11, atol=1e-15)
    if m > 1:
        class HostOnly:
            def predict(self, Xq): return g.predict(Xq)
        dev = offline.evaluate_gp(g, Xe, Re, Xe[:, :6], Xe[:, 6:])
        host = offline.evaluate_gp(HostOnly(), Xe, Re, Xe[:, :6], Xe[:, 6:])
        for col in host["per_component"].columns[1:]:
            np.testing.assert_allclose(dev["per_component"][col], host["per_component"][col], rtol=1e-8, atol=1e-12)
        assert dev["global"]["mse_gp"] == pytest.approx(host["global"]["mse_gp"], rel=1e-9)
        assert dev["fractions"] == pytest.approx(host["fractions"], abs=1e-12)


@pytest.mark.gpu
def test_copy_kernels_take_more_than_65535_rows():
    """the row index of the copy kernels lives in grid.x (grid.y stops at 65535): a 70 000-row pad-copy must launch and
    be exact -- the same indexing serves gpk_extract_lower (the L_ download) and the L21 copy of the recursion"""
    import ctypes as C
    import torch
    from unmanned_aerial_vehicles_b200 import _lib
    lib = _lib.load()
    R, Cc, RP, CP = 70000, 5, 70016, 128
    src = torch.arange(R * Cc, dtype=torch.float64, device="cuda").reshape(R, Cc)
    dst = torch.full((RP, CP), -1.0, dtype=torch.float64, device="cuda")
    rc = lib.gpk_pad_copy(C.c_void_p(src.data_ptr()), R, Cc, C.c_void_p(dst.data_ptr()), RP, CP,
                          C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.equal(dst[:R, :Cc], src) and float(dst[R:].abs().sum()) == 0.0 and float(dst[:, Cc:].abs().sum()) == 0.0


@pytest.mark.gpu
def test_distributed_fit_single_rank_on_device_equals_the_single_gpu_fit():
    """dist_fit.DistributedGPR through libgpk.so (one rank: the block-cyclic layout with a 1x1 grid) against the
    single-GPU estimator with the same optimiser: same theta*, LML and predictions"""
    import warnings
    from sklearn.gaussian_process.kernels import RBF, ConstantKernel as CK, WhiteKernel
    from unmanned_aerial_vehicles_b200 import GaussianProcessRegressor
    from unmanned_aerial_vehicles_b200.dist_fit import DistributedGPR
    rng = np.random.default_rng(12)
    X = rng.standard_normal((1500, 5))
    y = np.sin(X @ rng.standard_normal(5) / 2) + 0.1 * rng.standard_normal(1500) + 3.0
    k = CK(1.0, constant_value_bounds="fixed") * RBF([1.0] * 5, (0.1, 10.0)) + WhiteKernel(0.05, (1e-5, 10.0))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        e = GaussianProcessRegressor(kernel=k, alpha=1e-6, normalize_y=True).fit(X, y)
        g = DistributedGPR(k, alpha=1e-6, normalize_y=True, NB=512).fit(X, y)
    np.testing.assert_allclose(g.kernel_.theta, e.kernel_.theta, rtol=0, atol=1e-5)
    assert abs(g.log_marginal_likelihood_value_ - e.log_marginal_likelihood_value_) <= 1e-8 * abs(e.log_marginal_likelihood_value_)
    me, se = e.predict(X[:100] + 0.03, return_std=True)
    mg, sg = g.predict(X[:100] + 0.03, return_std=True)
    assert mg.shape == me.shape and sg.shape == se.shape
    np.testing.assert_allclose(mg, me, rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(sg, se, rtol=0, atol=1e-6)
