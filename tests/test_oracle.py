"""Pin the CPU oracle: against stock sklearn 1.9.0 executed in-process, and against the golden vectors the
reference's own trainer produced (tests/golden, made by oracle/make_golden.py)."""
import os

import numpy as np
import pytest
from sklearn.gaussian_process import GaussianProcessRegressor as SkGPR
from sklearn.gaussian_process.kernels import RBF, ConstantKernel as C, WhiteKernel

from oracle import gp_oracle as orc


def _data(n=120, d=4, t=2, seed=0):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, d))
    Y = np.sin(X @ rng.standard_normal((d, t))) + 0.1 * rng.standard_normal((n, t))
    return X, Y


def test_gram_matches_sklearn():
    X, _ = _data()
    k = C(1.7) * RBF([0.5, 1.0, 2.0, 0.8]) + WhiteKernel(0.2)
    np.testing.assert_allclose(orc.gram(X, np.array([0.5, 1.0, 2.0, 0.8]), 1.7, 0.2), k(X), rtol=0, atol=1e-14)


@pytest.mark.parametrize("iso", [True, False])
def test_lml_grad_matches_sklearn(iso):
    X, Y = _data()
    ls = 0.7 if iso else np.array([0.5, 1.0, 2.0, 0.8])
    k = C(1.7) * RBF(ls) + WhiteKernel(0.2)
    g = SkGPR(kernel=k, alpha=1e-6, optimizer=None).fit(X, Y)
    lml_s, grad_s = g.log_marginal_likelihood(g.kernel_.theta, eval_gradient=True)
    lml_o, go = orc.lml_and_grad(X, Y, ls, 1.7, 0.2, 1e-6)
    grad_o = np.concatenate([[go["c"]], [go["ls"].sum()] if iso else go["ls"], [go["noise"]]])
    assert abs(lml_o - lml_s) <= 1e-10 * abs(lml_s)
    np.testing.assert_allclose(grad_o, grad_s, rtol=1e-9, atol=1e-9)


def test_non_pd_returns_minus_inf():
    X = np.zeros((5, 2))
    lml, g = orc.lml_and_grad(X, np.ones(5), 1.0, 1.0, 0.0, 0.0)
    assert lml == -np.inf and not g["ls"].any()


def test_factor_predict_matches_sklearn():
    X, Y = _data()
    k = RBF(0.9) + WhiteKernel(0.05)
    g = SkGPR(kernel=k, alpha=1e-4, optimizer=None, normalize_y=True).fit(X, Y)
    L, a = orc.factor(X, g.y_train_, 0.9, 1.0, 0.05, 1e-4)
    np.testing.assert_allclose(L, g.L_, atol=1e-12)
    np.testing.assert_allclose(a, g.alpha_, rtol=1e-9, atol=1e-12)
    Xq = X[:20] * 1.05
    m_s, s_s = g.predict(Xq, return_std=True)
    m_o, s_o = orc.predict(Xq, X, L, a, 0.9, 1.0, 0.05, g._y_train_mean, g._y_train_std)
    np.testing.assert_allclose(m_o, m_s, rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(s_o, s_s, rtol=0, atol=1e-10)


def test_oracle_against_reference_golden(golden_dir):
    """Reference trainer output (SimpleQuadrotorGP.train_gp on its CSV): LML grid + predictions."""
    z = np.load(os.path.join(golden_dir, "simple_gp_csv.npz"))
    X, Y = z["X"], z["Y"]
    yn = (Y - z["y_mean"]) / z["y_std"]
    for th, lml_ref, grad_ref in list(zip(z["theta_grid"], z["lml_grid"], z["grad_grid"]))[:3]:
        lml, g = orc.lml_and_grad(X, yn, np.exp(th[0]), 1.0, np.exp(th[1]), 1e-4)
        assert abs(lml - lml_ref) <= 1e-9 * abs(lml_ref)
        np.testing.assert_allclose([g["ls"].sum(), g["noise"]], grad_ref, rtol=1e-7, atol=1e-7)
    th = z["theta"]
    L, a = orc.factor(X, yn, np.exp(th[0]), 1.0, np.exp(th[1]), 1e-4)
    np.testing.assert_allclose(a, z["alpha_"], rtol=1e-8, atol=1e-10)
    np.testing.assert_allclose(np.diag(L), z["L_diag"], rtol=1e-12)
    m, s = orc.predict(z["Xq"], X, L, a, np.exp(th[0]), 1.0, np.exp(th[1]), z["y_mean"], z["y_std"])
    np.testing.assert_allclose(m, z["mean"], rtol=1e-8, atol=1e-11)
    np.testing.assert_allclose(s, z["std"], rtol=0, atol=1e-9)
