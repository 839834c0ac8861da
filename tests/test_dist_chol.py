"""Distributed 2-D block-cyclic Cholesky with look-ahead: host logic on CPU with a NumPy/gloo ops object (world
sizes 1, 2 and 4; grids 1x1, 1x2, 2x1, 2x2, 1x4), and the real CUDA ops on one GPU against the single-GPU engine."""
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
        self.groups = {"world": None, "row": None, "col": None}
        self.log = []

    def rank_world(self):
        return (self.dist.get_rank(), self.dist.get_world_size()) if self.dist else (0, 1)

    def make_grid(self, Pr, Pc):
        rank, P = self.rank_world()
        assert Pr * Pc == P
        pr, pc = rank // Pc, rank % Pc
        if P > 1:
            for r in range(Pr):
                g = self.dist.new_group([r * Pc + c for c in range(Pc)]) if Pc > 1 else None
                if r == pr: self.groups["row"] = g
            for c in range(Pc):
                g = self.dist.new_group([r * Pc + c for r in range(Pr)]) if Pr > 1 else None
                if c == pc: self.groups["col"] = g
        return pr, pc

    def info_slots(self, n): self.infos = np.zeros(max(n, 1), dtype=np.int32)
    def read_infos(self): return self.infos

    import contextlib as _cl

    @_cl.contextmanager
    def side(self):
        self.log.append("side+")
        yield
        self.log.append("side-")

    def join_side(self): self.log.append("join")

    def zeros(self, r, c): return np.zeros((r, c))
    def empty(self, r, c): return np.full((r, c), np.nan)
    def from_host(self, a): return np.array(a, dtype=np.float64)
    def to_host(self, a): return np.array(a)
    def zero(self, a): a[...] = 0.0
    def copy(self, src, dst): dst[...] = src
    def copy_triu(self, src, dst): dst[...] = np.triu(src)
    def sync(self): pass

    def gram_block(self, Xq, mq, Xc, nc, D, inv_ls, c, out, MP, NB):
        from oracle import gp_oracle as orc
        out[...] = 0.0
        if mq and nc:
            out[:mq, :nc] = c * orc.rbf_cross(Xq[:mq], Xc[:nc], 1.0 / np.asarray(inv_ls))

    def fix_diag(self, A, n, n_valid, add, pad):
        for i in range(n):
            A[i, i] = A[i, i] + add if i < n_valid else pad

    def potrf_inv(self, A, Ml, Mu, NB, slot):
        try:
            L = np.linalg.cholesky(np.tril(A) + np.tril(A, -1).T)
        except np.linalg.LinAlgError:
            self.infos[slot] = 1
            A[...] = np.nan; Ml[...] = np.nan; Mu[...] = np.nan
            return
        A[...] = L
        Ml[...] = np.linalg.inv(L)
        Mu[...] = Ml.T
        # like the device kernel, leave garbage (not zeros) in the 128-blocks the recursion never writes:
        # strictly above Mlow's block diagonal, strictly below Mup's
        nb = NB // 128
        for bi in range(nb):
            for bj in range(nb):
                blk = (slice(128 * bi, 128 * bi + 128), slice(128 * bj, 128 * bj + 128))
                if bj > bi: Ml[blk] = np.nan
                if bj < bi: Mu[blk] = np.nan

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

    def bcast(self, buf, src, kind="world"):
        if not self.dist or (kind != "world" and self.groups[kind] is None):
            return
        t = self._t(np.ascontiguousarray(buf)); self.dist.broadcast(t, src=src, group=self.groups[kind])
        buf[...] = t.numpy()

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


def _check_predict(g, X, Y, ls, c, noise, jit):
    """DistGP.predict against the oracle (sklearn _gpr.py:446-500): mean 1e-6 rel, std 1e-8 abs (north star), on
    training rows (variance ~ noise: worst cancellation), nearby rows and far-field rows; ragged chunking"""
    from oracle import gp_oracle as orc
    rng = np.random.default_rng(9)
    Xq = np.vstack([X[::37], X[3::41] + 1e-3 * rng.standard_normal(X[3::41].shape), 4.0 * rng.standard_normal((5, X.shape[1]))])
    L, a = orc.factor(X, Y, ls, c, noise, jit)
    m_o, s_o = orc.predict(Xq, X, L, a, ls, c, noise)
    mean, std = g.predict(Xq, return_std=True, chunk_rows=128 if len(Xq) > 140 else 256)
    assert mean.shape == (len(Xq), Y.shape[1]) and std.shape == mean.shape
    np.testing.assert_allclose(mean, m_o, rtol=1e-6, atol=1e-9)
    assert np.max(np.abs(std - s_o.reshape(len(Xq), -1))) <= 1e-8
    np.testing.assert_allclose(g.predict(Xq[:7]), m_o[:7], rtol=1e-6, atol=1e-9)      # mean only: no exchange of W
    with pytest.raises(ValueError):
        g.predict(Xq[:, :-1])


def test_single_rank_numpy_ops():
    from unmanned_aerial_vehicles_b200.dist_chol import DistGP, default_grid
    assert [default_grid(p) for p in (1, 2, 4, 6, 8)] == [(1, 1), (1, 2), (2, 2), (2, 3), (2, 4)]
    X, Y, ls, c, noise, jit = _problem()
    g = DistGP(X, Y, jit, NB=128, ops=NumpyOps())
    assert g.nb == 3 and g.NP == 384 and g.owned == [0, 1, 2] and g.diag == [0, 1, 2]
    assert g.factor(1.0 / ls, c, noise) == 0
    # look-ahead: the broadcasts of panel k+1 are issued on the side stream, joined before panel k+1 is used
    assert g.ops.log[:4] == ["side+", "side-", "join", "side+"] and g.ops.log[-1] == "join"
    g.solve()
    _check(g, X, Y, ls, c, noise, jit)
    _check_predict(g, X, Y, ls, c, noise, jit)
    g2 = DistGP(X, Y, jit, NB=128, ops=NumpyOps(), lookahead=False)
    assert g2.factor(1.0 / ls, c, noise) == 0 and "side+" not in g2.ops.log
    g2.solve()
    _check(g2, X, Y, ls, c, noise, jit)
    bad = DistGP(np.zeros((200, 4)), Y[:200], 0.0, NB=128, ops=NumpyOps())
    assert bad.factor(1.0 / ls, 1.0, 0.0) == 1          # info = k*NB + block info (the NumPy double reports 1)
    with pytest.raises(ValueError):
        DistGP(X, Y, jit, NB=128, ops=NumpyOps(), grid=(2, 1))


def test_tile_ownership_maps():
    """every lower tile has exactly one owner and the local stacking matches (i % Pr, j % Pc)"""
    from unmanned_aerial_vehicles_b200.dist_chol import DistGP

    class FakeOps(NumpyOps):
        def __init__(self, rank, P): super().__init__(); self.r, self.P = rank, P
        def rank_world(self): return self.r, self.P
        def make_grid(self, Pr, Pc): return self.r // Pc, self.r % Pc

    X, Y, *_ = _problem(n=1100)
    for Pr, Pc in [(2, 4), (4, 2), (1, 8), (2, 2), (3, 2)]:
        seen = {}
        for r in range(Pr * Pc):
            g = DistGP(X, Y, 1e-6, NB=128, ops=FakeOps(r, Pr * Pc), grid=(Pr, Pc))
            assert sorted(g.colpos.values()) == list(range(g.nlc))
            for j in g.my_cols:
                rows = [i for i in g.my_rows if i >= j]
                assert g.cols[j].shape == (len(rows) * 128, 128)
                assert g.lfirst(j) == len([i for i in g.my_rows if i < j])
                for i in rows:
                    assert (i, j) not in seen and g.owner(i, j) == r
                    seen[(i, j)] = r
        assert len(seen) == g.nb * (g.nb + 1) // 2


def _worker(rank, world, port, q, grid, n):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from test_dist_chol import NumpyOps, _check, _check_predict, _problem
    from unmanned_aerial_vehicles_b200.dist_chol import DistGP
    X, Y, ls, c, noise, jit = _problem(n=n)
    g = DistGP(X, Y, jit, NB=128, ops=NumpyOps(dist), grid=grid)
    info = g.factor(1.0 / ls, c, noise)
    g.solve()
    _check(g, X, Y, ls, c, noise, jit)
    _check_predict(g, X, Y, ls, c, noise, jit)
    bad = DistGP(np.zeros((300, 4)), Y[:300], 0.0, NB=128, ops=NumpyOps(dist), grid=grid)
    q.put((rank, info, g.owned, g.lml(), bad.factor(1.0 / ls, 1.0, 0.0)))
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
    assert all(o[1] == 0 for o in outs) and all(o[3] == outs[0][3] for o in outs)
    assert all(o[4] == 1 for o in outs)                 # non-PD: same info on every rank
    return outs


def test_two_ranks_gloo_1x2():
    outs = _run(2, (1, 2), 600)
    assert outs[0][2] == [0, 2, 4] and outs[1][2] == [1, 3]


def test_two_ranks_gloo_2x1():
    outs = _run(2, (2, 1), 600)
    assert outs[0][2] == [0, 1, 2, 3, 4] == outs[1][2]


@pytest.mark.parametrize("grid,n", [((2, 2), 1000), ((1, 4), 700), ((4, 1), 700)])
def test_four_ranks_gloo(grid, n):
    _run(4, grid, n)


def test_eight_ranks_gloo_2x4():
    """the grid the 8-GPU runs use by default (lcm(Pr, Pc) = 4: half of the transposed-exchange pieces are empty)"""
    outs = _run(8, (2, 4), 1300)
    assert [o[2] for o in outs[:4]] == [[0, 4, 8], [1, 5, 9], [2, 6, 10], [3, 7]]


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
