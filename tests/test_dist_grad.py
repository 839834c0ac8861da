"""Gradient of the log marginal likelihood from the DISTRIBUTED factor (dist_grad.py: block-cyclic TRTRI + LAUUM +
tile-wise gradient contraction): host logic on CPU with the NumPy/gloo ops double on grids 1x1 ... 2x4 against the
oracle (sklearn _gpr.py:629-651), and the CUDA ops on one GPU against the single-GPU engine."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))

from test_dist_chol import NumpyOps, _problem  # noqa: E402


class NumpyGradOps(NumpyOps):
    """adds the two gradient entry points of dist_chol.CudaOps (gpk_grad_block / gpk_grad_block_sum) in NumPy"""

    def grad_ws(self, NB, D):
        return np.zeros(D + 2)

    def grad_block(self, Xr, nr, gi0, Xc, nc, gj0, D, inv_ls, c, ar, ac, T, Kinv, rows, cols, ws, accum):
        if not accum:
            ws[...] = 0.0
        inv_ls = np.asarray(inv_ls)
        A, B = Xr[:nr] * inv_ls, Xc[:nc] * inv_ls
        gi, gj = gi0 + np.arange(nr)[:, None], gj0 + np.arange(nc)[None, :]
        W = ar[:nr] @ ac[:nc].T - T * Kinv[:nr, :nc]
        d2 = ((A[:, None, :] - B[None, :, :]) ** 2)
        Kc = c * np.exp(-0.5 * d2.sum(-1))
        off = np.where(gj < gi, 2.0, 0.0) * W * Kc                 # strictly lower entries count twice
        ws[:D] += np.einsum("ij,ijd->d", off, d2)
        dg = (gj == gi)
        ws[D] += off.sum() + (W * dg).sum() * c
        ws[D + 1] += (W * dg).sum()

    def grad_sum(self, ws, NB, D, empty):
        return np.zeros((D + 2, 1)) if empty else ws.reshape(-1, 1).copy()


def _check_grad(g, X, Y, ls, c, noise, jit):
    from oracle import gp_oracle as orc
    lml_o, go = orc.lml_and_grad(X, Y, ls, c, noise, jit)
    lml, g_c, g_ls, g_noise = g.lml_grad(1.0 / ls, c, noise)
    assert abs(lml - lml_o) <= 1e-10 * abs(lml_o)
    np.testing.assert_allclose(g_ls, go["ls"], rtol=1e-7, atol=1e-9)
    assert g_c == pytest.approx(go["c"], rel=1e-7, abs=1e-9) and g_noise == pytest.approx(go["noise"], rel=1e-7, abs=1e-9)
    # the factor survives the inverse: alpha and predict still answer from L
    a = orc.factor(X, Y, ls, c, noise, jit)[1]
    np.testing.assert_allclose(g.alpha_host(), a, rtol=1e-8, atol=1e-9)
    m = g.predict(X[:5] + 0.01)
    assert m.shape == (5, Y.shape[1])


def test_single_rank_numpy_ops_gradient():
    from unmanned_aerial_vehicles_b200.dist_chol import DistGP
    X, Y, ls, c, noise, jit = _problem()
    for NB, look in ((128, True), (256, False)):
        g = DistGP(X, Y, jit, NB=NB, ops=NumpyGradOps(), lookahead=look)
        _check_grad(g, X, Y, ls, c, noise, jit)
        # K^-1 tiles themselves
        K = __import__("oracle.gp_oracle", fromlist=["x"]).gram(X, ls, c, noise) + jit * np.eye(len(X))
        Kinv = np.linalg.inv(K)
        for j in g.my_cols:
            for i in g.my_rows:
                if i >= j:
                    r0 = (i // g.Pr - g.lfirst(j)) * NB
                    n_i, n_j = min(NB, len(X) - i * NB), min(NB, len(X) - j * NB)
                    np.testing.assert_allclose(g.kinv[j][r0:r0 + n_i, :n_j], Kinv[i * NB:i * NB + n_i, j * NB:j * NB + n_j],
                                               rtol=0, atol=1e-8 * np.abs(Kinv).max())
    bad = DistGP(np.zeros((200, 4)), Y[:200], 0.0, NB=128, ops=NumpyGradOps())
    lml, g_c, g_ls, g_noise = bad.lml_grad(1.0 / ls, 1.0, 0.0)
    assert lml == -np.inf and g_c == 0.0 and not g_ls.any() and g_noise == 0.0      # _gpr.py:592-593


def _worker(rank, world, port, q, grid, n):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from test_dist_chol import _problem
    from test_dist_grad import NumpyGradOps, _check_grad
    from unmanned_aerial_vehicles_b200.dist_chol import DistGP
    X, Y, ls, c, noise, jit = _problem(n=n)
    g = DistGP(X, Y, jit, NB=128, ops=NumpyGradOps(dist), grid=grid)
    _check_grad(g, X, Y, ls, c, noise, jit)
    lml, g_c, g_ls, g_noise = g.lml_grad(1.0 / (ls * 1.3), c, 2 * noise)          # a second theta on the same object
    q.put((rank, lml, g_c, g_ls.tolist(), g_noise))
    dist.barrier(); dist.destroy_process_group()


def _run(world, grid, n):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q, grid, n)) for r in range(world)]
    [p.start() for p in procs]
    outs = sorted(q.get(timeout=300) for _ in range(world))
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert all(o[1:] == outs[0][1:] for o in outs)       # every rank holds the same lml and gradient
    from oracle import gp_oracle as orc
    X, Y, ls, c, noise, jit = _problem(n=n)
    lml_o, go = orc.lml_and_grad(X, Y, ls * 1.3, c, 2 * noise, jit)
    assert abs(outs[0][1] - lml_o) <= 1e-10 * abs(lml_o)
    np.testing.assert_allclose(outs[0][3], go["ls"], rtol=1e-7, atol=1e-9)


@pytest.mark.parametrize("world,grid,n", [(2, (1, 2), 600), (2, (2, 1), 600), (4, (2, 2), 1000), (8, (2, 4), 900)])
def test_gradient_on_gloo_grids(world, grid, n):
    _run(world, grid, n)


@pytest.mark.gpu
@pytest.mark.parametrize("n,NB,t", [(1000, 256, 3), (3000, 1024, 1), (2048, 512, 2)])
def test_cuda_ops_single_rank_gradient_matches_engine(n, NB, t):
    from unmanned_aerial_vehicles_b200.dist_chol import DistGP
    from unmanned_aerial_vehicles_b200.engine import DeviceGP
    rng = np.random.default_rng(1)
    X = rng.standard_normal((n, 9)); Y = np.sin(X @ rng.standard_normal((9, t)) / 3) + 0.1 * rng.standard_normal((n, t))
    ls = np.linspace(0.8, 2.0, 9)
    g = DistGP(X, Y, 1e-6, NB=NB)
    lml, g_c, g_ls, g_noise = g.lml_grad(1.0 / ls, 1.3, 0.02)
    eng = DeviceGP(X, Y, 1e-6)
    lml_e, gc_e, gls_e, gn_e = eng.lml(1.0 / ls, 1.3, 0.02, True)
    assert abs(lml - lml_e) <= 1e-10 * abs(lml_e)
    np.testing.assert_allclose(g_ls, gls_e, rtol=1e-7, atol=1e-8)
    assert g_c == pytest.approx(gc_e, rel=1e-7, abs=1e-8) and g_noise == pytest.approx(gn_e, rel=1e-7, abs=1e-8)
    m, s = g.predict(X[:64] + 0.01, return_std=True)                          # L is intact after the inverse
    assert eng.factor(1.0 / ls, 1.3, 0.02) == 0
    me, se, _, _ = eng.predict(X[:64] + 0.01, 0.0, 1.0, True)
    np.testing.assert_allclose(m, me, rtol=1e-6, atol=1e-9)
    assert np.max(np.abs(s - se)) <= 1e-8
