"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle, stock sklearn on the same
arrays, and the golden vectors made by the reference's own trainer.  Tolerances are BASELINE.json's:
mean 1e-6 relative, std 1e-8 absolute, LML 1e-6 relative (tests assert tighter where the math allows)."""
import ctypes as C
import os
import pickle
import warnings

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from sklearn.gaussian_process import GaussianProcessRegressor as SkGPR  # noqa: E402
from sklearn.gaussian_process.kernels import RBF, ConstantKernel as CK, WhiteKernel  # noqa: E402

from oracle import gp_oracle as orc  # noqa: E402
from unmanned_aerial_vehicles_b200 import GaussianProcessRegressor, _lib  # noqa: E402
from unmanned_aerial_vehicles_b200.engine import DeviceGP  # noqa: E402


def _p(t):
    return C.c_void_p(t.data_ptr())


def _st():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _data(n, d, t, seed=0):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, d))
    Y = np.sin(X @ (rng.standard_normal((d, t)) / np.sqrt(d))) + 0.1 * rng.standard_normal((n, t))
    return X, Y


# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("a_tri,b_tri,lower_only,acc", [(0, 0, 0, 0), (0, 1, 0, 0), (2, 0, 0, 0), (1, 0, 0, 0),
                                                        (2, 2, 1, 0), (0, 0, 1, 1)])
@pytest.mark.parametrize("size", [384, 1024, 2048])
def test_gemm_nt(a_tri, b_tri, lower_only, acc, size):
    lib = _lib.load()
    rng = np.random.default_rng(1)
    M = N = K = size
    A = rng.standard_normal((M, K)); B = rng.standard_normal((N, K)); C0 = rng.standard_normal((M, N))
    if a_tri == 1: A = np.tril(A)
    if a_tri == 2: A = np.triu(A)
    if b_tri == 1: B = np.tril(B)
    if b_tri == 2: B = np.triu(B)
    Ad, Bd, Cd = (torch.from_numpy(x).cuda() for x in (A, B, C0.copy()))
    rc = lib.gpk_gemm_nt(_p(Ad), K, _p(Bd), K, _p(Cd), N, M, N, K, -0.5, acc, a_tri, b_tri, lower_only, _st())
    assert rc == 0
    got = Cd.cpu().numpy()
    ref = -0.5 * (A @ B.T) + (C0 if acc else 0.0)
    if lower_only:  # contract: every element with col <= row is produced; 128-tiles strictly above are untouched
        np.testing.assert_allclose(np.tril(got), np.tril(ref), rtol=0, atol=1e-11)
        for i in range(0, M, 128):
            for j in range(i + 128, N, 128):
                blk = (slice(i, i + 128), slice(j, j + 128))
                np.testing.assert_array_equal(got[blk], C0[blk])
    else:
        np.testing.assert_allclose(got, ref, rtol=0, atol=1e-11)


@pytest.mark.parametrize("n,d,iso", [(64, 3, True), (300, 10, False), (257, 9, False), (130, 17, False)])
def test_gram_matches_oracle(n, d, iso):
    X, Y = _data(n, d, 1)
    ls = np.full(d, 0.7) if iso else np.linspace(0.5, 2.0, d)
    eng = DeviceGP(X, Y, 1e-6)
    K = eng.gram_host(1.0 / ls, 1.3, 0.2, jitter=1e-6)
    Ko = orc.gram(X, ls, 1.3, 0.2)
    Ko[np.diag_indices_from(Ko)] += 1e-6
    np.testing.assert_allclose(K, Ko, rtol=0, atol=1e-13)


@pytest.mark.parametrize("n", [100, 128, 700, 1500])
def test_cholesky_and_inverse(n):
    X, Y = _data(n, 5, 2, seed=n)
    ls = np.full(5, 1.2)
    eng = DeviceGP(X, Y, 1e-8)
    assert eng.factor(1.0 / ls, 1.0, 0.01) == 0
    L = eng.L_host()
    K = orc.gram(X, ls, 1.0, 0.01); K[np.diag_indices_from(K)] += 1e-8
    assert np.linalg.norm(L @ L.T - K) / np.linalg.norm(K) < 1e-14
    assert not np.triu(L, 1).any()
    Lo, ao = orc.factor(X, Y, ls, 1.0, 0.01, 1e-8)
    np.testing.assert_allclose(L, Lo, rtol=0, atol=1e-10)
    np.testing.assert_allclose(eng.alpha_host(), ao, rtol=1e-8, atol=1e-8 * np.abs(ao).max())
    Ml = eng.Mlow_d[:n, :n].cpu().numpy(); Mu = eng.Mup_d[:n, :n].cpu().numpy()
    np.testing.assert_allclose(np.tril(Ml) @ L, np.eye(n), rtol=0, atol=1e-9)
    np.testing.assert_allclose(np.triu(Mu), np.tril(Ml).T, rtol=0, atol=0)


def _golden_lml_case(z, key=""):
    return z["theta_grid" + key], z["lml_grid" + key], z["grad_grid" + key]


def test_lml_grad_reference_csv_golden(golden_dir):
    """sklearn LML+grad on the reference's CSV/trainer (iso RBF+White, D=10, T=6, normalize_y)."""
    z = np.load(os.path.join(golden_dir, "simple_gp_csv.npz"))
    g = GaussianProcessRegressor(kernel=RBF(0.5) + WhiteKernel(0.1), alpha=1e-4, normalize_y=True, optimizer=None)
    g.fit(z["X"], z["Y"])
    for th, lml_ref, grad_ref in zip(*_golden_lml_case(z)):
        lml, grad = g.log_marginal_likelihood(th, eval_gradient=True)
        assert abs(lml - lml_ref) <= 1e-8 * abs(lml_ref), (th, lml, lml_ref)
        np.testing.assert_allclose(grad, grad_ref, rtol=1e-6, atol=1e-6 * np.abs(grad_ref).max())
        assert abs(g.log_marginal_likelihood(th) - lml_ref) <= 1e-8 * abs(lml_ref)


@pytest.mark.parametrize("t", [0, 1, 2])
def test_lml_grad_ard_trainer_golden(golden_dir, t):
    z = np.load(os.path.join(golden_dir, "ard_trainer.npz"))
    k = CK(1.0, constant_value_bounds="fixed") * RBF([1.0] * 9, (0.1, 10.0)) + WhiteKernel(0.01, (1e-5, 10.0))
    g = GaussianProcessRegressor(kernel=k, alpha=1e-6, optimizer=None).fit(z["Xs"], z[f"ys{t}"])
    for th, lml_ref, grad_ref in zip(*_golden_lml_case(z, str(t))):
        lml, grad = g.log_marginal_likelihood(th, eval_gradient=True)
        assert abs(lml - lml_ref) <= 1e-8 * abs(lml_ref)
        np.testing.assert_allclose(grad, grad_ref, rtol=1e-6, atol=1e-6 * np.abs(grad_ref).max())


def test_lml_grad_free_constant_multi_output_golden(golden_dir):
    z = np.load(os.path.join(golden_dir, "lml_grid_const_ard.npz"))
    k = CK(2.0, (1e-2, 1e2)) * RBF([1.0, 2.0, 0.5, 1.5, 1.0], (1e-2, 1e2)) + WhiteKernel(0.05, (1e-5, 1e1))
    g = GaussianProcessRegressor(kernel=k, alpha=1e-8, optimizer=None, normalize_y=True).fit(z["X"], z["Y"])
    np.testing.assert_allclose(g.alpha_, z["alpha_"], rtol=1e-7, atol=1e-9)
    for th, lml_ref, grad_ref in zip(*_golden_lml_case(z)):
        lml, grad = g.log_marginal_likelihood(th, eval_gradient=True)
        assert abs(lml - lml_ref) <= 1e-8 * abs(lml_ref)
        np.testing.assert_allclose(grad, grad_ref, rtol=1e-6, atol=1e-6 * np.abs(grad_ref).max())
    m, s = g.predict(z["Xq"], return_std=True)
    np.testing.assert_allclose(m, z["mean"], rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(s, z["std"], rtol=0, atol=1e-8)


def test_predict_reference_csv_golden(golden_dir):
    """theta fixed at the reference trainer's optimum: alpha_, L_, mean, std and predict_residual contract."""
    z = np.load(os.path.join(golden_dir, "simple_gp_csv.npz"))
    th = z["theta"]
    k = RBF(float(np.exp(th[0]))) + WhiteKernel(float(np.exp(th[1])))
    g = GaussianProcessRegressor(kernel=k, alpha=1e-4, normalize_y=True, optimizer=None).fit(z["X"], z["Y"])
    assert abs(g.log_marginal_likelihood_value_ - z["lml"]) <= 1e-8 * abs(z["lml"])
    np.testing.assert_allclose(g.alpha_, z["alpha_"], rtol=1e-7, atol=1e-9)
    np.testing.assert_allclose(np.diag(g.L_), z["L_diag"], rtol=1e-11)
    np.testing.assert_allclose(g.L_.sum(1), z["L_rowsum"], rtol=1e-9, atol=1e-11)
    m, s = g.predict(z["Xq"], return_std=True)
    assert m.shape == z["mean"].shape and s.shape == z["std"].shape
    np.testing.assert_allclose(m, z["mean"], rtol=1e-6, atol=1e-10)
    np.testing.assert_allclose(s, z["std"], rtol=0, atol=1e-8)
    # SimpleQuadrotorGP.predict_residual (simple_gp.py:187-201): one row, (mean.flatten(), std.flatten()**2)
    x = np.concatenate([z["state"], z["control"]]).reshape(1, -1)
    mean, std = g.predict(x, return_std=True)
    np.testing.assert_allclose(mean.flatten(), z["pr_mean"], rtol=1e-6, atol=1e-10)
    np.testing.assert_allclose(std.flatten() ** 2, z["pr_var"], rtol=0, atol=1e-8)
    np.testing.assert_allclose(g.predict(z["Xq"]), z["mean"], rtol=1e-6, atol=1e-10)


def test_fit_with_optimizer_matches_reference_trainer(golden_dir):
    """SimpleQuadrotorGP.train_gp configuration end to end, same seed as the golden run (G5)."""
    z = np.load(os.path.join(golden_dir, "simple_gp_csv.npz"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        g = GaussianProcessRegressor(kernel=RBF(length_scale=0.5) + WhiteKernel(noise_level=0.1), alpha=1e-4,
                                     normalize_y=True, n_restarts_optimizer=1, random_state=0).fit(z["X"], z["Y"])
    np.testing.assert_allclose(g.kernel_.theta, z["theta"], rtol=0, atol=1e-4)
    assert abs(g.log_marginal_likelihood_value_ - z["lml"]) <= 1e-6 * abs(z["lml"])
    assert abs(g.log_marginal_likelihood() - z["lml"]) <= 1e-6 * abs(z["lml"])
    assert g.n_features_in_ == 10 and g.X_train_.shape == (1000, 10)
    m, s = g.predict(z["Xq"], return_std=True)
    np.testing.assert_allclose(m, z["mean"], rtol=1e-4, atol=1e-6)   # theta agrees to 1e-4 only
    blob = pickle.dumps({"gp_model": g, "is_trained": True})          # train_gp_offline.py:188-197 schema
    g2 = pickle.loads(blob)["gp_model"]
    m2, s2 = g2.predict(z["Xq"], return_std=True)
    np.testing.assert_array_equal(m2, m)
    np.testing.assert_allclose(s2, s, rtol=0, atol=1e-12)


@pytest.mark.parametrize("n,d,t,iso,normalize", [(333, 9, 3, False, False), (1000, 9, 1, False, True),
                                                 (2100, 10, 6, True, True)])
def test_against_live_sklearn(n, d, t, iso, normalize):
    X, Y = _data(n, d, t, seed=7)
    if t == 1:
        Y = Y[:, 0]
    ls = 1.3 if iso else list(np.linspace(0.8, 2.5, d))
    mk = lambda: CK(1.0, constant_value_bounds="fixed") * RBF(ls, (0.1, 10)) + WhiteKernel(0.01, (1e-5, 10))
    s = SkGPR(kernel=mk(), alpha=1e-6, optimizer=None, normalize_y=normalize).fit(X, Y)
    g = GaussianProcessRegressor(kernel=mk(), alpha=1e-6, optimizer=None, normalize_y=normalize).fit(X, Y)
    assert abs(g.log_marginal_likelihood_value_ - s.log_marginal_likelihood_value_) <= \
        1e-9 * abs(s.log_marginal_likelihood_value_)
    np.testing.assert_allclose(g.alpha_, s.alpha_, rtol=1e-6, atol=1e-7 * np.abs(s.alpha_).max())
    np.testing.assert_allclose(g.L_, s.L_, rtol=0, atol=1e-9)
    rng = np.random.default_rng(3)
    Xq = np.vstack([X[:50], X[:50] + 1e-4, rng.standard_normal((200, d)) * 2])
    ms, ss = s.predict(Xq, return_std=True)
    mg, sg = g.predict(Xq, return_std=True)
    assert mg.shape == ms.shape and sg.shape == ss.shape
    np.testing.assert_allclose(mg, ms, rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(sg, ss, rtol=0, atol=1e-8)
    lml_s, gr_s = s.log_marginal_likelihood(s.kernel_.theta + 0.3, eval_gradient=True)
    lml_g, gr_g = g.log_marginal_likelihood(g.kernel_.theta + 0.3, eval_gradient=True)
    assert abs(lml_g - lml_s) <= 1e-9 * abs(lml_s)
    np.testing.assert_allclose(gr_g, gr_s, rtol=1e-6, atol=1e-7 * np.abs(gr_s).max())
    _, cs = s.predict(Xq[:40], return_cov=True)
    _, cg = g.predict(Xq[:40], return_cov=True)
    assert cg.shape == cs.shape
    np.testing.assert_allclose(cg, cs, rtol=0, atol=1e-8)


def test_non_positive_definite_behaviour():
    """_gpr.py:592-593: (-inf, zeros) during optimisation; :353-361 LinAlgError from fit."""
    X = np.zeros((40, 3)); y = np.ones(40)
    g = GaussianProcessRegressor(kernel=RBF(1.0, (1e-2, 1e2)), alpha=0.0, optimizer=None)
    with pytest.raises(np.linalg.LinAlgError, match="not returning a positive definite matrix"):
        g.fit(X, y)
    g2 = GaussianProcessRegressor(kernel=RBF(1.0, (1e-2, 1e2)), alpha=1e-3, optimizer=None).fit(X + np.arange(40)[:, None], y)
    g2.alpha = 0.0
    g2._engine = None
    g2.X_train_ = X
    lml, grad = g2.log_marginal_likelihood(np.array([0.0]), eval_gradient=True)
    assert lml == -np.inf and not grad.any()


def test_negative_variance_clamped_with_warning():
    X, Y = _data(200, 2, 1, seed=5)
    g = GaussianProcessRegressor(kernel=RBF(5.0), alpha=1e-12, optimizer=None).fit(X, Y[:, 0])
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        _, s = g.predict(X, return_std=True)
    assert (s >= 0).all() and np.isfinite(s).all()
    if (s == 0).any():
        assert any("Predicted variances smaller than 0" in str(x.message) for x in w)


@pytest.mark.parametrize("n", [6000, 20000])
def test_large_n_properties(n):
    """Size-independent checks up to BASELINE's full N=20000: L L^T = K on sampled rows, K alpha = y,
    predictive mean at training rows = y - (noise + alpha) alpha_, finite-difference gradient."""
    d, t = 9, 3
    X, Y = _data(n, d, t, seed=11)
    ls = np.full(d, 1.5)
    eng = DeviceGP(X, Y, 1e-6)
    assert eng.factor(1.0 / ls, 1.0, 0.01) == 0
    a = eng.alpha_host()
    rows = np.arange(0, n, 97)
    Krows = orc.rbf_cross(X[rows], X, ls); Krows[np.arange(len(rows)), rows] += 0.01 + 1e-6
    np.testing.assert_allclose(Krows @ a, Y[rows], rtol=0, atol=1e-8)
    L = eng.L_host()
    np.testing.assert_allclose(L[rows] @ L.T, Krows, rtol=0, atol=1e-12)
    del L
    mean, std, _, _ = eng.predict(X[rows], 0.0, 1.0, True)
    np.testing.assert_allclose(mean, Y[rows] - (0.01 + 1e-6) * a[rows], rtol=0, atol=1e-8)
    assert np.all(std > 0) and np.all(std < np.sqrt(1.01))
    lml, g_c, g_ls, g_noise = eng.lml(1.0 / ls, 1.0, 0.01, True)
    eps = 1e-5
    lp, *_ = eng.lml(1.0 / ls, 1.0, 0.01 * np.exp(eps), False)
    lm, *_ = eng.lml(1.0 / ls, 1.0, 0.01 * np.exp(-eps), False)
    assert abs((lp - lm) / (2 * eps) - g_noise) <= 1e-4 * abs(g_noise)
    lp, *_ = eng.lml(1.0 / (ls * np.exp(eps)), 1.0, 0.01, False)
    lm, *_ = eng.lml(1.0 / (ls * np.exp(-eps)), 1.0, 0.01, False)
    assert abs((lp - lm) / (2 * eps) - g_ls.sum()) <= 1e-4 * abs(g_ls.sum())


def test_from_sklearn_pickle_adoption():
    X, Y = _data(400, 10, 6, seed=2)
    s = SkGPR(kernel=RBF(0.5) + WhiteKernel(0.1), alpha=1e-4, normalize_y=True, optimizer=None).fit(X, Y)
    g = GaussianProcessRegressor.from_sklearn(pickle.loads(pickle.dumps(s)))
    ms, ss = s.predict(X[:64] + 0.01, return_std=True)
    mg, sg = g.predict(X[:64] + 0.01, return_std=True)
    np.testing.assert_allclose(mg, ms, rtol=1e-6, atol=1e-10)
    np.testing.assert_allclose(sg, ss, rtol=0, atol=1e-8)


@pytest.mark.parametrize("n", [700, 3000])
def test_small_batch_variance_path_matches_gemm_path(n):
    """Query batches of <= 32 rows take the triangular mat-vec variance kernel, larger ones the DMMA contraction
    with the row-sum-of-squares epilogue: same numbers, and both equal the oracle (std 1e-8 abs)."""
    X, Y = _data(n, 9, 3, seed=5)
    ls = np.linspace(0.7, 1.9, 9)
    eng = DeviceGP(X, Y, 1e-6)
    assert eng.factor(1.0 / ls, 1.0, 0.02) == 0
    Xq = X[:100] + 0.03
    m_big, s_big, _, _ = eng.predict(Xq, 0.0, 1.0, True)
    for lo, hi in [(0, 1), (1, 3), (3, 11), (11, 43 - 11), (32, 64), (64, 96), (96, 100)]:
        m, s, _, _ = eng.predict(Xq[lo:hi], 0.0, 1.0, True)
        np.testing.assert_allclose(m, m_big[lo:hi], rtol=0, atol=1e-12)
        np.testing.assert_allclose(s, s_big[lo:hi], rtol=0, atol=1e-11)
    L, a = orc.factor(X, Y, ls, 1.0, 0.02, 1e-6)
    m_o, s_o = orc.predict(Xq, X, L, a, ls, 1.0, 0.02)
    np.testing.assert_allclose(m_big, m_o, rtol=1e-6, atol=1e-9)
    assert np.max(np.abs(s_big - s_o)) <= 1e-8


@pytest.mark.parametrize("n,d,t,m", [(333, 3, 1, 1), (1000, 10, 6, 25), (700, 9, 3, 300), (2100, 9, 3, 5000),
                                     (500, 14, 10, 130), (900, 20, 2, 77), (400, 8, 4, 64)])
def test_mean_only_predict_fused_kernel(n, d, t, m):
    """return_std=False (evaluate_gp_offline.py:198, gp_enhanced_mpc.py:165,175) takes the fused kernel: no cross-Gram
    in HBM.  Against the oracle's K* alpha (_gpr.py:446-454) at 1e-6 relative (north star) -- in fact ~1e-13 -- and
    against the two-pass route (cross-Gram, then K* alpha) the return_std=True call takes, over every feature padding
    (4/8/10/12/16/32), both target widths (4, 8, two passes beyond 8), one query tile or many, splits of the
    training rows or one."""
    X, Y = _data(n, d, t, seed=11)
    if t == 1:
        Y = Y[:, 0]
    ls = np.linspace(0.7, 2.2, d)
    k = CK(1.7) * RBF(ls, (0.1, 10)) + WhiteKernel(0.02, (1e-5, 10))
    g = GaussianProcessRegressor(kernel=k, alpha=1e-6, optimizer=None, normalize_y=True).fit(X, Y)
    rng = np.random.default_rng(5)
    Xq = np.vstack([X[: m // 2] + 1e-3, rng.standard_normal((m - m // 2, d)) * 1.5])
    got = g.predict(Xq)
    two_pass, _ = g.predict(Xq, return_std=True)
    assert got.shape == two_pass.shape
    L, a = orc.factor(X, g.y_train_, ls, 1.7, 0.02, 1e-6)
    want, _ = orc.predict(Xq, X, L, a, ls, 1.7, 0.02, g._y_train_mean, g._y_train_std, return_std=False)
    want = want[:, 0] if got.ndim == 1 else want
    scale = np.abs(want).max()
    np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-9 * scale)
    np.testing.assert_allclose(got, two_pass, rtol=0, atol=1e-11 * scale)
    # many chunks == one, up to the summation order (the number of training-row splits follows the chunk height)
    chunked = g._engine.predict(Xq, g._yscale_b[0], g._yscale_b[1], False, chunk_rows=128)[0]
    np.testing.assert_allclose(chunked if got.ndim == 2 else chunked[:, 0], got, rtol=0, atol=1e-12 * scale)
    again = g.predict(Xq)
    np.testing.assert_array_equal(again, got)                          # fixed reduction order: run-to-run identical
