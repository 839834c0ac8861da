"""DistributedGPR (dist_fit.py): hyper-parameter fit on the distributed factor -- every rank runs the same L-BFGS-B on
objective values that are identical after the all-reduces.  CPU: NumPy/gloo ops double against stock scikit-learn's
fit on the same data (same optimiser, same starts)."""
import os
import socket
import sys
import warnings

import numpy as np
import pytest
from sklearn.gaussian_process import GaussianProcessRegressor as SkGPR
from sklearn.gaussian_process.kernels import RBF, ConstantKernel as CK, WhiteKernel

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))

from test_dist_grad import NumpyGradOps  # noqa: E402


def _case(n=260, d=3, t=2, seed=4):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, d))
    Y = np.sin(X @ rng.standard_normal((d, t))) + 0.1 * rng.standard_normal((n, t)) + np.array([2.0, -1.0])[:t]
    k = CK(1.0, constant_value_bounds="fixed") * RBF([1.0] * d, (0.1, 10.0)) + WhiteKernel(0.05, (1e-5, 10.0))
    return X, Y, k


def _compare(g, s, X):
    np.testing.assert_allclose(g.kernel_.theta, s.kernel_.theta, rtol=0, atol=2e-4)
    assert abs(g.log_marginal_likelihood_value_ - s.log_marginal_likelihood_value_) <= 1e-6 * abs(s.log_marginal_likelihood_value_)
    Xq = X[:11] + 0.05
    ms, ss = s.predict(Xq, return_std=True)
    mg, sg = g.predict(Xq, return_std=True)
    assert mg.shape == ms.shape and sg.shape == ss.shape
    np.testing.assert_allclose(mg, ms, rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(sg, ss, rtol=1e-4, atol=1e-6)


def test_distributed_fit_single_rank_matches_sklearn():
    from unmanned_aerial_vehicles_b200.dist_fit import DistributedGPR
    X, Y, k = _case()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        s = SkGPR(kernel=k, alpha=1e-6, n_restarts_optimizer=1, normalize_y=True, random_state=0).fit(X, Y)
    g = DistributedGPR(k, alpha=1e-6, n_restarts_optimizer=1, normalize_y=True, random_state=0, NB=128,
                       ops=NumpyGradOps()).fit(X, Y)
    _compare(g, s, X)
    assert g.n_lml_evals_ > 5 and len(g.evaluation_seconds_) == g.n_lml_evals_
    th = s.kernel_.theta + 0.2
    lml_s, gr_s = s.log_marginal_likelihood(th, eval_gradient=True)
    lml_g, gr_g = g.log_marginal_likelihood(th, eval_gradient=True)
    assert abs(lml_g - lml_s) <= 1e-9 * abs(lml_s)
    np.testing.assert_allclose(gr_g, gr_s, rtol=1e-6, atol=1e-8)
    assert g.log_marginal_likelihood(th) == pytest.approx(lml_s, rel=1e-9)
    np.testing.assert_allclose(g.predict(X[:3]), s.predict(X[:3]), rtol=1e-4, atol=1e-6)   # fitted factor restored
    # pickle: host NumPy only, no factor inside; the loaded model re-factors on first use and answers the same
    import pickle
    blob = pickle.dumps({"gp_model": g, "is_trained": True})             # train_gp_offline.py:188-197 schema
    assert len(blob) < 40 * X.nbytes                                     # X, y, kernel -- not an N x N matrix
    back = pickle.loads(blob)["gp_model"]
    assert back._gp is None and back.ops is None
    back.ops = NumpyGradOps()
    for a_, b_ in zip(back.predict(X[:9] + 0.05, return_std=True), g.predict(X[:9] + 0.05, return_std=True)):
        np.testing.assert_allclose(a_, b_, rtol=1e-12, atol=1e-14)
    assert back.log_marginal_likelihood() == g.log_marginal_likelihood() and str(back.kernel_) == str(g.kernel_)
    # truncated run and 1-D targets (single-output models of gp_trainer.py:169-176)
    g1 = DistributedGPR(k, alpha=1e-6, maxiter=2, NB=128, ops=NumpyGradOps()).fit(X, Y[:, 0])
    assert g1.predict(X[:4]).shape == (4,) and g1.predict(X[:4], return_std=True)[1].shape == (4,)
    assert g1.n_lml_evals_ <= 8
    # maxfun is a hard cap (SciPy's own is only checked between iterations): exactly 2 evaluations, best one kept
    seen = []
    g2 = DistributedGPR(k, alpha=1e-6, maxfun=2, on_eval=lambda kk, lml, sec: seen.append((kk, lml)), NB=128,
                        ops=NumpyGradOps()).fit(X, Y[:, 0])
    assert g2.n_lml_evals_ == 2 and [kk for kk, _ in seen] == [1, 2]
    assert g2.log_marginal_likelihood_value_ == max(l for _, l in seen)
    assert g2.predict(X[:4]).shape == (4,)                                        # the final factor is in place
    fixed = DistributedGPR(k, alpha=1e-6, optimizer=None, NB=128, ops=NumpyGradOps()).fit(X, Y)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        s0 = SkGPR(kernel=k, alpha=1e-6, optimizer=None).fit(X, Y)
    assert fixed.log_marginal_likelihood() == pytest.approx(s0.log_marginal_likelihood_value_, rel=1e-10)
    with pytest.raises(np.linalg.LinAlgError, match="not returning a positive definite matrix"):
        DistributedGPR(RBF(1.0, (1e-2, 1e2)), alpha=0.0, optimizer=None, NB=128, ops=NumpyGradOps()).fit(np.zeros((40, 3)), np.ones(40))


def _worker(rank, world, port, q, grid):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from test_dist_fit import _case
    from test_dist_grad import NumpyGradOps
    from unmanned_aerial_vehicles_b200.dist_fit import DistributedGPR
    X, Y, k = _case()
    g = DistributedGPR(k, alpha=1e-6, normalize_y=True, NB=128, grid=grid, ops=NumpyGradOps(dist)).fit(X, Y)
    m, sd = g.predict(X[:6] + 0.05, return_std=True)
    q.put((rank, g.kernel_.theta.tolist(), float(g.log_marginal_likelihood_value_), g.n_lml_evals_, m.tolist(), sd.tolist()))
    dist.barrier(); dist.destroy_process_group()


@pytest.mark.parametrize("world,grid", [(2, (1, 2)), (4, (2, 2))])
def test_distributed_fit_on_gloo_grids_matches_sklearn(world, grid):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    s_ = socket.socket(); s_.bind(("127.0.0.1", 0)); port = s_.getsockname()[1]; s_.close()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q, grid)) for r in range(world)]
    [p.start() for p in procs]
    outs = sorted(q.get(timeout=600) for _ in range(world))
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert all(o[1:] == outs[0][1:] for o in outs)            # identical iterates and results on every rank
    X, Y, k = _case()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        s = SkGPR(kernel=k, alpha=1e-6, normalize_y=True).fit(X, Y)
    np.testing.assert_allclose(outs[0][1], s.kernel_.theta, rtol=0, atol=2e-4)
    assert abs(outs[0][2] - s.log_marginal_likelihood_value_) <= 1e-6 * abs(s.log_marginal_likelihood_value_)
    ms, ss = s.predict(X[:6] + 0.05, return_std=True)
    np.testing.assert_allclose(outs[0][4], ms, rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(outs[0][5], ss, rtol=1e-4, atol=1e-6)
