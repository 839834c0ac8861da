"""Distributed (1-D block-column-cyclic) Cholesky: host logic on CPU with a NumPy/gloo ops object (world sizes 1
and 2), and the real CUDA ops on one GPU against the single-GPU engine."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class NumpyOps:
    """test double for dist_chol.CudaOps: same interface, NumPy arithmetic, torch.distributed(gloo) collectives"""

    def __init__(self, dist=None):
        self.dist = dist

    def rank_world(self):
        return (self.dist.get_rank(), self.dist.get_world_size()) if self.dist else (0, 1)

    def zeros(self, r, c): return np.zeros((r, c))
    def empty(self, r, c): return np.full((r, c), np.nan)
    def from_host(self, a): return np.array(a, dtype=np.float64)
    def to_host(self, a): return np.array(a)
    def zero(self, a): a[...] = 0.0
    def copy(self, src, dst): dst[...] = src
    def sync(self): pass

    def gram_block(self, Xq, mq, Xc, nc, D, inv_ls, c, out, MP, NB):
        from oracle import gp_oracle as orc
        out[...] = 0.0
        if mq and nc:
            out[:mq, :nc] = c * orc.rbf_cross(Xq[:mq], Xc[:nc], 1.0 / np.asarray(inv_ls))

    def fix_diag(self, A, n, n_valid, add, pad):
        for i in range(n):
            A[i, i] = A[i, i] + add if i < n_valid else pad

    def potrf_inv(self, A, Ml, Mu, NB):
        try:
            L = np.linalg.cholesky(np.tril(A) + np.tril(A, -1).T)
        except np.linalg.LinAlgError:
            return 1
        A[...] = L
        Ml[...] = np.linalg.inv(L)
        Mu[...] = Ml.T
        return 0

    def gemm_nt(self, A, B, Cm, M, N, K, alpha, acc, b_tri=0):
        Bm = np.tril(B[:N, :K]) if b_tri == 1 else B[:N, :K]
        r = alpha * (A[:M, :K] @ Bm.T)
        Cm[:M, :N] = Cm[:M, :N] + r if acc else r

    def gemv(self, A, v, out, alpha, beta, transpose):
        if transpose:
            out += alpha * (A.T @ v)
        else:
            out[...] = (beta * out if beta else 0.0) + alpha * (A @ v)

    def trimv(self, Mtri, v, out, upper):
        out[...] = (np.triu(Mtri) if upper else np.tril(Mtri)) @ v

    def logdiag(self, A, n, acc):
        acc[0, 0] += np.log(np.diag(A)[:n]).sum()

    def _t(self, a):
        import torch
        return torch.from_numpy(a)

    def bcast(self, buf, src):
        if self.dist:
            t = self._t(np.ascontiguousarray(buf)); self.dist.broadcast(t, src=src); buf[...] = t.numpy()

    def allreduce_sum(self, buf):
        if self.dist:
            t = self._t(np.ascontiguousarray(buf)); self.dist.all_reduce(t); buf[...] = t.numpy()

    def allreduce_max_int(self, v):
        if self.dist:
            import torch
            t = torch.tensor([v]); self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX); return int(t[0])
        return v


def _problem(n=300, d=4, t=2):
    rng = np.random.default_rng(0)
    X = rng.standard_normal((n, d)); Y = np.sin(X @ rng.standard_normal((d, t))) + 0.1 * rng.standard_normal((n, t))
    return X, Y, np.array([0.8, 1.2, 1.0, 1.5]), 1.3, 0.05, 1e-6


def _check(g, X, Y, ls, c, noise, jit):
    from oracle import gp_oracle as orc
    L, a = orc.factor(X, Y, ls, c, noise, jit)
    lml, _ = orc.lml_and_grad(X, Y, ls, c, noise, jit, eval_gradient=False)
    np.testing.assert_allclose(g.alpha_host(), a, rtol=1e-8, atol=1e-9)
    assert abs(g.lml() - lml) <= 1e-10 * abs(lml)


def test_single_rank_numpy_ops():
    from unmanned_aerial_vehicles_b200.dist_chol import DistGP
    X, Y, ls, c, noise, jit = _problem()
    g = DistGP(X, Y, jit, NB=128, ops=NumpyOps())
    assert g.nb == 3 and g.NP == 384 and g.owned == [0, 1, 2]
    assert g.factor(1.0 / ls, c, noise) == 0
    g.solve()
    _check(g, X, Y, ls, c, noise, jit)
    bad = DistGP(np.zeros((200, 4)), Y[:200], 0.0, NB=128, ops=NumpyOps())
    assert bad.factor(1.0 / ls, 1.0, 0.0) != 0


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from test_dist_chol import NumpyOps, _check, _problem
    from unmanned_aerial_vehicles_b200.dist_chol import DistGP
    X, Y, ls, c, noise, jit = _problem(n=600)
    g = DistGP(X, Y, jit, NB=128, ops=NumpyOps(dist))
    info = g.factor(1.0 / ls, c, noise)
    g.solve()
    _check(g, X, Y, ls, c, noise, jit)
    q.put((rank, info, g.owned, g.lml()))
    dist.barrier(); dist.destroy_process_group()


def test_two_ranks_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    outs = sorted(q.get(timeout=300) for _ in range(2))
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert outs[0][1] == 0 and outs[0][2] == [0, 2, 4] and outs[1][2] == [1, 3]
    assert outs[0][3] == outs[1][3]


@pytest.mark.gpu
@pytest.mark.parametrize("n,NB", [(1000, 256), (3000, 1024), (2048, 512)])
def test_cuda_ops_single_rank_matches_engine(n, NB):
    from unmanned_aerial_vehicles_b200.dist_chol import DistGP
    from unmanned_aerial_vehicles_b200.engine import DeviceGP
    rng = np.random.default_rng(1)
    X = rng.standard_normal((n, 9)); Y = np.sin(X @ rng.standard_normal((9, 3)) / 3) + 0.1 * rng.standard_normal((n, 3))
    ls = np.linspace(0.8, 2.0, 9)
    g = DistGP(X, Y, 1e-6, NB=NB)
    assert g.factor(1.0 / ls, 1.0, 0.01) == 0
    g.solve()
    eng = DeviceGP(X, Y, 1e-6)
    assert eng.factor(1.0 / ls, 1.0, 0.01) == 0
    np.testing.assert_allclose(g.alpha_host(), eng.alpha_host(), rtol=1e-7, atol=1e-8)
    lml_ref = eng.lml(1.0 / ls, 1.0, 0.01, False)[0]
    assert abs(g.lml() - lml_ref) <= 1e-10 * abs(lml_ref)
    _check(g, X, Y, ls, 1.0, 0.01, 1e-6)
