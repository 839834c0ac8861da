"""Host-side logic: kernel lowering, theta ordering, gradient assembly, estimator surface."""
import numpy as np
import pytest
from sklearn.base import clone
from sklearn.gaussian_process.kernels import RBF, ConstantKernel as C, Matern, WhiteKernel

from unmanned_aerial_vehicles_b200 import GaussianProcessRegressor
from unmanned_aerial_vehicles_b200.kernel_lowering import assemble_gradient, lower


def test_lower_reference_production_kernel():
    low = lower(RBF(0.5) + WhiteKernel(0.1), 10)   # simple_gp.py:167
    assert low.iso and low.c == 1.0 and low.noise == 0.1
    assert np.all(low.length_scale == 0.5) and low.length_scale.shape == (10,)
    assert low.theta_map == [("ls_iso", 0), ("noise", 0)]


def test_lower_reference_ard_kernel():
    k = C(1.0, constant_value_bounds="fixed") * RBF([1.0] * 9, (0.1, 10)) + WhiteKernel(0.01, (1e-5, 10))
    low = lower(k, 9)                               # gp_trainer.py:163-166
    assert not low.iso and len(low.theta_map) == k.n_dims == 10
    assert low.theta_map[-1] == ("noise", 0) and low.theta_map[0] == ("ls", 0)


@pytest.mark.parametrize("k", [
    WhiteKernel(0.3) + RBF(2.0), RBF(2.0) * C(3.0), C(3.0) * RBF([1, 2]) + WhiteKernel(0.1), RBF([1.0, 2.0])])
def test_lower_orderings(k):
    low = lower(k, 2)
    assert len(low.theta_map) == k.n_dims
    g = assemble_gradient(low, 10.0, np.array([1.0, 2.0]), 20.0)
    names = [hp.name for hp in k.hyperparameters if not hp.fixed]
    pos = 0
    for n in names:
        if n.endswith("constant_value"):
            assert g[pos] == 10.0; pos += 1
        elif n.endswith("noise_level"):
            assert g[pos] == 20.0; pos += 1
        else:
            if low.iso:
                assert g[pos] == 3.0; pos += 1
            else:
                assert list(g[pos:pos + 2]) == [1.0, 2.0]; pos += 2


def test_unsupported_kernel_raises():
    with pytest.raises(NotImplementedError):
        lower(Matern(1.0), 3)
    with pytest.raises(NotImplementedError):
        lower(RBF(1.0) + RBF(2.0), 3)


def test_estimator_params_and_clone():
    g = GaussianProcessRegressor(kernel=RBF(0.5) + WhiteKernel(0.1), alpha=1e-4, normalize_y=True,
                                 n_restarts_optimizer=1)
    p = g.get_params()
    assert p["alpha"] == 1e-4 and p["normalize_y"] and p["n_restarts_optimizer"] == 1
    g2 = clone(g)
    assert g2.get_params()["kernel"].get_params() == p["kernel"].get_params()
    assert hasattr(g, "predict") and hasattr(g, "fit") and hasattr(g, "log_marginal_likelihood")


def test_prior_predict_needs_no_device():
    g = GaussianProcessRegressor(kernel=C(2.0) * RBF(1.0) + WhiteKernel(0.5))
    m, s = g.predict(np.zeros((4, 3)), return_std=True)
    assert m.shape == (4,) and np.allclose(s, np.sqrt(2.5))


@pytest.mark.parametrize("k,d", [
    (RBF(0.5) + WhiteKernel(0.1), 10),                                                   # simple_gp.py:167
    (C(1.0, constant_value_bounds="fixed") * RBF([1.0] * 9, (0.1, 10.0)) + WhiteKernel(0.01, (1e-5, 10.0)), 9),
    (WhiteKernel(0.05) + C(2.0) * RBF([1.5, 0.7, 3.0]), 3),
    (RBF([1.0, 2.0], length_scale_bounds="fixed") + WhiteKernel(0.3), 2),               # fixed length-scales
    (C(2.0) * RBF(1.5), 4),
])
def test_with_theta_equals_lowering_the_cloned_kernel(k, d):
    """the optimiser's shortcut (theta -> flat parameters) gives bit for bit what sklearn's theta setter gives"""
    rng = np.random.default_rng(0)
    template = lower(k, d)
    for _ in range(20):
        theta = k.theta + rng.uniform(-3, 3, size=k.theta.shape)
        fast, slow = template.with_theta(theta), lower(k.clone_with_theta(theta), d)
        assert fast.c == slow.c and fast.noise == slow.noise and fast.iso == slow.iso
        assert np.array_equal(fast.length_scale, slow.length_scale) and fast.theta_map == slow.theta_map
    assert np.array_equal(template.length_scale, lower(k, d).length_scale)               # template untouched
    with pytest.raises(ValueError):
        template.with_theta(np.zeros(len(k.theta) + 1))


@pytest.mark.parametrize("G", [1, 3, 8, 16, 24])
@pytest.mark.parametrize("ntm,ntn", [(1, 1), (7, 5), (8, 8), (19, 11), (128, 128)])
def test_gemm_raster_map_is_a_bijection(G, ntm, ntn):
    """Model of the tile raster of gpk_gemm_nt_kernel (csrc/gpk_gemm.cu: groups of G tile-rows, column-major inside a
    group; G = 8 unless GPK_GEMM_RASTER is set): every (tm, tn) is visited exactly once for any group height,
    including a ragged last group."""
    seen = set()
    for ident in range(ntm * ntn):
        grp, within = divmod(ident, G * ntn)
        gsz = min(G, ntm - grp * G)
        tm, tn = grp * G + within % gsz, within // gsz
        assert 0 <= tm < ntm and 0 <= tn < ntn
        seen.add((tm, tn))
    assert len(seen) == ntm * ntn
