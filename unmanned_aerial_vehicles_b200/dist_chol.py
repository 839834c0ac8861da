"""Exact GP factorisation for training sets too large for one GPU (BASELINE config 5: N=200k, 320 GB Gram).

1-D block-column-cyclic right-looking Cholesky over the ranks of a torch.distributed group (NCCL over
NVLink/NVSwitch on GPUs).  Block column j (width NB, rows j*NB..NP) lives on rank j % P; only the lower
trapezoid is stored (N^2/2/P doubles per rank: 20 GB at N=200k, P=8).  K is generated in place from the
replicated X (14 MB) so the Gram matrix is never communicated.  Per panel k:

    owner:  L_kk, L_kk^-1 = potrf_inv(A_kk)  (the single-GPU recursive kernel);  L_{>k,k} = A_{>k,k} L_kk^-T
    all  :  ncclBroadcast of the factored panel ((NP - k NB) x NB doubles)
    all  :  A_{>=j, j} -= L_{>=j,k} L_{j,k}^T   for every owned block column j > k   (DMMA NT GEMMs, K = NB)

Total broadcast volume is N^2/2 doubles = 160 GB at N=200k, i.e. ~0.25 s at the measured 700 GB/s against
~11 s of FP64 tensor work per rank, so a 2-D grid (smaller volume, more complex) is not needed on NVSwitch.
alpha = K^-1 y by distributed forward/back substitution (one tiny all-reduce / broadcast per panel), then
lml = -1/2 y.alpha - T sum log L_ii - T N/2 log 2 pi  (sklearn _gpr.py:601-617).

The numerical work goes through an `ops` object: `CudaOps` (libgpk.so + NCCL) is the product; the tests inject
a NumPy/gloo implementation to exercise the distributed logic on CPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib


class CudaOps:
    """libgpk.so kernels on torch-allocated HBM; collectives through torch.distributed (NCCL)."""

    def __init__(self, group=None):
        import torch
        import torch.distributed as dist

        if not torch.cuda.is_available():
            raise _lib.GpkUnavailable("no CUDA device: the distributed GP path has no CPU fallback")
        self.t, self.dist, self.group = torch, dist, group
        self.lib = _lib.load()
        self.dev = torch.device("cuda", torch.cuda.current_device())
        self.info = torch.zeros(1, dtype=torch.int32, device=self.dev)

    def _st(self):
        return C.c_void_p(self.t.cuda.current_stream().cuda_stream)

    @staticmethod
    def _p(x):
        return C.c_void_p(x.data_ptr())

    def rank_world(self):
        d = self.dist
        return (d.get_rank(self.group), d.get_world_size(self.group)) if d.is_initialized() else (0, 1)

    def zeros(self, rows, cols):
        return self.t.zeros((rows, cols), dtype=self.t.float64, device=self.dev)

    def empty(self, rows, cols):
        return self.t.empty((rows, cols), dtype=self.t.float64, device=self.dev)

    def from_host(self, a):
        return self.t.as_tensor(np.ascontiguousarray(a, dtype=np.float64), device=self.dev)

    def to_host(self, a):
        return a.cpu().numpy()

    def gram_block(self, Xq, mq, Xc, nc, D, inv_ls, c, out, MP, NB):
        """out (MP x NB) = c*rbf(Xq[:mq], Xc[:nc]) (zero padded)"""
        _lib.check(self.lib.gpk_cross_gram(self._p(Xq), mq, self._p(Xc), nc, D, _lib.darr(inv_ls), c, self._p(out),
                                           MP, NB, self._st()), "gpk_cross_gram")

    def fix_diag(self, A, n, n_valid, add, pad):
        _lib.check(self.lib.gpk_fix_diag(self._p(A), A.stride(0), n, n_valid, add, pad, self._st()), "gpk_fix_diag")

    def potrf_inv(self, A, Ml, Mu, NB):
        _lib.check(self.lib.gpk_potrf_inv(self._p(A), self._p(Ml), self._p(Mu), NB, self._p(self.info), self._st()),
                   "gpk_potrf_inv")
        return int(self.info.cpu()[0])

    def gemm_nt(self, A, B, Cm, M, N, K, alpha, acc, b_tri=0):
        _lib.check(self.lib.gpk_gemm_nt(self._p(A), A.stride(0), self._p(B), B.stride(0), self._p(Cm), Cm.stride(0),
                                        M, N, K, alpha, acc, 0, b_tri, 0, self._st()), "gpk_gemm_nt")

    def copy(self, src, dst):
        dst.copy_(src)

    def zero(self, a):
        a.zero_()

    def gemv(self, A, v, out, alpha, beta, transpose):
        M, K = A.shape
        _lib.check(self.lib.gpk_gemv(self._p(A), A.stride(0), M, K, self._p(v), v.shape[1], self._p(out), alpha, beta,
                                     int(transpose), self._st()), "gpk_gemv")

    def trimv(self, Mtri, v, out, upper):
        _lib.check(self.lib.gpk_trimv(self._p(Mtri), Mtri.shape[0], self._p(v), self._p(out), v.shape[1], int(upper),
                                      self._st()), "gpk_trimv")

    def logdiag(self, A, n, acc):
        _lib.check(self.lib.gpk_logdiag(self._p(A), A.stride(0), n, self._p(acc), self._st()), "gpk_logdiag")

    def bcast(self, buf, src):
        if self.dist.is_initialized() and self.rank_world()[1] > 1:
            self.dist.broadcast(buf, src=src, group=self.group)

    def allreduce_sum(self, buf):
        if self.dist.is_initialized() and self.rank_world()[1] > 1:
            self.dist.all_reduce(buf, group=self.group)

    def allreduce_max_int(self, v):
        if self.dist.is_initialized() and self.rank_world()[1] > 1:
            tt = self.t.tensor([v], dtype=self.t.int64, device=self.dev)
            self.dist.all_reduce(tt, op=self.dist.ReduceOp.MAX, group=self.group)
            return int(tt[0])
        return v

    def sync(self):
        self.t.cuda.synchronize()


class DistGP:
    """Distributed exact GP: factor K(theta) + alpha*I, solve for alpha_, evaluate the LML."""

    def __init__(self, X, y2d, jitter, NB=2048, ops=None):
        self.ops = ops if ops is not None else CudaOps()
        self.rank, self.P = self.ops.rank_world()
        X = np.ascontiguousarray(X, dtype=np.float64)
        y2d = np.ascontiguousarray(y2d, dtype=np.float64)
        self.N, self.D = X.shape
        self.T = y2d.shape[1]
        assert NB % 128 == 0
        self.NB = NB
        self.nb = (self.N + NB - 1) // NB
        self.NP = self.nb * NB
        self.jitter = float(jitter)
        o = self.ops
        self.X_d = o.from_host(X)
        ypad = np.zeros((self.NP, self.T))
        ypad[: self.N] = y2d
        self.y_d = o.from_host(ypad)
        self.owned = [j for j in range(self.nb) if j % self.P == self.rank]
        self.cols = {j: o.empty(self.NP - j * NB, NB) for j in self.owned}       # lower trapezoid only
        self.Ml = {j: o.empty(NB, NB) for j in self.owned}                        # L_jj^-1
        self.Mu = {j: o.empty(NB, NB) for j in self.owned}                        # L_jj^-T
        self.buf = o.empty(self.NP, NB)                                           # panel broadcast buffer
        self.gemm_flops = 0.0
        self.factored = False

    def owner(self, j):
        return j % self.P

    def local_bytes(self):
        return 8 * (sum((self.NP - j * self.NB) * self.NB for j in self.owned) + 2 * len(self.owned) * self.NB ** 2
                    + self.NP * self.NB)

    # -- K generated in place ---------------------------------------------------------------------
    def assemble(self, inv_ls, c, noise):
        o, NB, N = self.ops, self.NB, self.N
        for j in self.owned:
            r0 = j * NB
            rows = self.NP - r0
            mq = max(0, N - r0)
            nc = max(0, min(NB, N - r0))
            o.gram_block(self.X_d[r0:], mq, self.X_d[r0:], nc, self.D, inv_ls, c, self.cols[j], rows, NB)
            # diagonal: (c + noise) + jitter on real rows (gram gives c there), identity in the padding
            o.fix_diag(self.cols[j], NB, nc, noise, 1.0)          # (c + noise), padding -> 1
            o.fix_diag(self.cols[j], nc, nc, self.jitter, 1.0)    # ... + alpha  (_gpr.py:589)

    # -- right-looking factorisation ----------------------------------------------------------------
    def factor(self, inv_ls, c, noise):
        o, NB, nb = self.ops, self.NB, self.nb
        self.assemble(inv_ls, c, noise)
        info = 0
        for k in range(nb):
            rows = self.NP - k * NB
            buf = self.buf[:rows]
            if self.owner(k) == self.rank:
                Ck = self.cols[k]
                bi = o.potrf_inv(Ck[:NB], self.Ml[k], self.Mu[k], NB)
                if bi and not info:
                    info = k * NB + bi
                o.copy(Ck[:NB], buf[:NB])
                if rows > NB:
                    o.gemm_nt(Ck[NB:], self.Ml[k], buf[NB:], rows - NB, NB, NB, 1.0, 0, b_tri=1)
                    o.copy(buf[NB:], Ck[NB:])
                    self.gemm_flops += (rows - NB) * float(NB) * NB
            o.bcast(buf, self.owner(k))
            for j in self.owned:
                if j <= k:
                    continue
                off = (j - k) * NB
                # A[j.., j] -= L[j.., k] L[j, k]^T
                o.gemm_nt(buf[off:], buf[off:off + NB], self.cols[j], rows - off, NB, NB, -1.0, 1)
                self.gemm_flops += 2.0 * (rows - off) * NB * NB
        info = o.allreduce_max_int(info)
        self.factored = info == 0
        return info

    # -- alpha = K^-1 y, log-det, lml -----------------------------------------------------------------
    def solve(self):
        assert self.factored
        o, NB, nb, T = self.ops, self.NB, self.nb, self.T
        z = o.zeros(self.NP, T)            # this rank's partial sums  sum_k L[:,k] u_k  over owned k
        u = o.zeros(self.NP, T)
        t = o.empty(NB, T)
        for k in range(nb):                # forward: L u = y
            blk = slice(k * NB, (k + 1) * NB)
            o.copy(z[blk], t)
            o.allreduce_sum(t)
            if self.owner(k) == self.rank:
                rhs = self.y_d[blk] - t
                o.trimv(self.Ml[k], rhs, u[blk], upper=False)
                if k + 1 < nb:
                    o.gemv(self.cols[k][NB:], u[blk], z[(k + 1) * NB:], 1.0, 1.0, transpose=False)
            o.bcast(u[blk], self.owner(k))
        alpha = o.zeros(self.NP, T)
        s = o.zeros(NB, T)
        for k in range(nb - 1, -1, -1):    # backward: L^T alpha = u
            blk = slice(k * NB, (k + 1) * NB)
            if self.owner(k) == self.rank:
                if k + 1 < nb:
                    o.zero(s)
                    o.gemv(self.cols[k][NB:], alpha[(k + 1) * NB:], s, 1.0, 0.0, transpose=True)
                    rhs = u[blk] - s
                else:
                    rhs = u[blk] + 0.0
                o.trimv(self.Mu[k], rhs, alpha[blk], upper=True)
            o.bcast(alpha[blk], self.owner(k))
        self.alpha_d, self.u_d = alpha, u
        return alpha

    def lml(self):
        o = self.ops
        acc = o.zeros(1, 1)
        for j in self.owned:
            o.logdiag(self.cols[j], self.NB, acc)
        o.allreduce_sum(acc)
        logdet = float(o.to_host(acc)[0, 0])
        y = o.to_host(self.y_d)[: self.N]
        a = o.to_host(self.alpha_d)[: self.N]
        lml = -0.5 * np.einsum("ik,ik->k", y, a) - logdet - self.N / 2 * np.log(2 * np.pi)
        return float(lml.sum())

    def alpha_host(self):
        return self.ops.to_host(self.alpha_d)[: self.N]
