"""Exact GP factorisation for training sets too large for one GPU (BASELINE config 5: N=200k, 320 GB Gram).

2-D block-cyclic right-looking Cholesky with one-panel look-ahead over a Pr x Pc process grid of a
torch.distributed group (NCCL over NVLink/NVSwitch on GPUs).  Tile (i, j) (NB x NB, i >= j) lives on grid
position (i % Pr, j % Pc); rank r sits at (r // Pc, r % Pc).  Only the lower trapezoid is stored, one tensor per
owned block column holding that column's owned row tiles stacked (N^2/2/P doubles per rank: 20 GB at N=200k,
P=8).  K is generated in place from the replicated X (14 MB) so the Gram matrix is never communicated.

Per panel k (process column pck = k % Pc):

    diag owner        L_kk, L_kk^-1 = potrf_inv(A_kk)          (the single-GPU recursive kernel)
    process column    ncclBroadcast of L_kk^-1 down the column group;  L_ik = A_ik L_kk^-T for the owned rows i > k
    process rows      ncclBroadcast of the owned L_ik along each row group (root: the rank in column pck)
    process columns   transposed exchange: rank (j % Pr, pc) broadcasts L_jk (j % Pc == pc) down column group pc
    all               A_ij -= L_ik L_jk^T for every owned tile i >= j > k   (DMMA NT GEMMs, K = NB)

Look-ahead: block column k+1 is updated first, panel k+1 is factored, and its two broadcasts are issued on a side
stream into the second panel buffer while the main stream finishes the trailing update of panel k.
Received volume per rank is 8 (N^2/2)(1/Pr + 1/Pc) bytes (120 GB on 2x4 at N=200k, 160 GB on 1x8).  A 1 x P grid
degenerates to the 1-D block-column-cyclic layout (no column exchange).
alpha = K^-1 y by distributed forward/back substitution (one small all-reduce + broadcast per panel), then
lml = -1/2 y.alpha - T sum log L_ii - T N/2 log 2 pi  (sklearn _gpr.py:601-617).

The numerical work goes through an `ops` object: `CudaOps` (libgpk.so + NCCL) is the product; the tests inject
a NumPy/gloo implementation to exercise the distributed logic on CPU.
"""
from __future__ import annotations

import contextlib
import ctypes as C

import numpy as np

from . import _lib
from .dist_grad import DistGradMixin


def default_grid(P):
    """most square Pr x Pc with Pr <= Pc (8 -> 2x4, 4 -> 2x2, 2 -> 1x2)"""
    pr = int(np.floor(np.sqrt(P)))
    while P % pr:
        pr -= 1
    return pr, P // pr


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
        self.groups = {"world": group, "row": None, "col": None}
        self.side_stream = torch.cuda.Stream(device=self.dev)
        self.bcast_bytes = 0          # bytes this rank sent or received through broadcasts (NCCL traffic of the path)
        self.bcast_calls = 0
        self.trace = None             # list of (phase, panel, start event, end event) while tracing

    def _st(self):
        return C.c_void_p(self.t.cuda.current_stream().cuda_stream)

    @staticmethod
    def _p(x):
        return C.c_void_p(x.data_ptr())

    def rank_world(self):
        d = self.dist
        return (d.get_rank(self.group), d.get_world_size(self.group)) if d.is_initialized() else (0, 1)

    def make_grid(self, Pr, Pc):
        """row / column sub-communicators of the Pr x Pc grid (every rank creates every group, same order)"""
        rank, P = self.rank_world()
        assert Pr * Pc == P
        pr, pc = rank // Pc, rank % Pc
        if P > 1:
            for r in range(Pr):
                g = self.dist.new_group([r * Pc + c for c in range(Pc)]) if Pc > 1 else None
                if r == pr:
                    self.groups["row"] = g
            for c in range(Pc):
                g = self.dist.new_group([r * Pc + c for r in range(Pr)]) if Pr > 1 else None
                if c == pc:
                    self.groups["col"] = g
            # NCCL builds a communicator at its first collective: do that here, not inside the first timed panel
            one = self.t.zeros(1, dtype=self.t.float64, device=self.dev)
            for kind in ("row", "col"):
                if self.groups[kind] is not None:
                    self.dist.all_reduce(one, group=self.groups[kind])
            self.t.cuda.synchronize()
        self.grid = (Pr, Pc)
        return pr, pc

    def zeros(self, rows, cols):
        return self.t.zeros((rows, cols), dtype=self.t.float64, device=self.dev)

    def empty(self, rows, cols):
        return self.t.empty((rows, cols), dtype=self.t.float64, device=self.dev)

    def from_host(self, a):
        return self.t.as_tensor(np.ascontiguousarray(a, dtype=np.float64), device=self.dev)

    def to_host(self, a):
        return a.cpu().numpy()

    def info_slots(self, n):
        self.infos = self.t.zeros(max(n, 1), dtype=self.t.int32, device=self.dev)

    def gram_block(self, Xq, mq, Xc, nc, D, inv_ls, c, out, MP, NB):
        """out (MP x NB) = c*rbf(Xq[:mq], Xc[:nc]) (zero padded)"""
        _lib.check(self.lib.gpk_cross_gram(self._p(Xq), mq, self._p(Xc), nc, D, _lib.darr(inv_ls), c, self._p(out),
                                           MP, NB, self._st()), "gpk_cross_gram")

    def fix_diag(self, A, n, n_valid, add, pad):
        _lib.check(self.lib.gpk_fix_diag(self._p(A), A.stride(0), n, n_valid, add, pad, self._st()), "gpk_fix_diag")

    def potrf_inv(self, A, Ml, Mu, NB, slot):
        """info (LAPACK convention, relative to this block) is left in device slot `slot`: no host sync here"""
        _lib.check(self.lib.gpk_potrf_inv(self._p(A), self._p(Ml), self._p(Mu), NB, self._p(self.infos[slot:]),
                                          self._st()), "gpk_potrf_inv")

    def read_infos(self):
        return self.infos.cpu().numpy()

    def gemm_nt(self, A, B, Cm, M, N, K, alpha, acc, b_tri=0):
        _lib.check(self.lib.gpk_gemm_nt(self._p(A), A.stride(0), self._p(B), B.stride(0), self._p(Cm), Cm.stride(0),
                                        M, N, K, alpha, acc, 0, b_tri, 0, self._st()), "gpk_gemm_nt")

    def copy(self, src, dst):
        dst.copy_(src)

    def zero(self, a):
        a.zero_()

    def copy_triu(self, src, dst):
        """dst = upper triangle of src, zeros below the diagonal"""
        dst.copy_(self.t.triu(src))

    def gemv(self, A, v, out, alpha, beta, transpose):
        M, K = A.shape
        _lib.check(self.lib.gpk_gemv(self._p(A), A.stride(0), M, K, self._p(v), v.shape[1], self._p(out), alpha, beta,
                                     int(transpose), self._st()), "gpk_gemv")

    def trimv(self, Mtri, v, out, upper):
        _lib.check(self.lib.gpk_trimv(self._p(Mtri), Mtri.shape[0], self._p(v), self._p(out), v.shape[1], int(upper),
                                      self._st()), "gpk_trimv")

    def logdiag(self, A, n, acc):
        _lib.check(self.lib.gpk_logdiag(self._p(A), A.stride(0), n, self._p(acc), self._st()), "gpk_logdiag")

    # -- gradient from the distributed inverse (dist_grad.py) ------------------------------------------
    def grad_ws(self, NB, D):
        n = int(self.lib.gpk_grad_block_workspace_doubles(NB, NB, D))
        return self.t.zeros(n, dtype=self.t.float64, device=self.dev)

    def grad_block(self, Xr, nr, gi0, Xc, nc, gj0, D, inv_ls, c, ar, ac, T, Kinv, rows, cols, ws, accum):
        _lib.check(self.lib.gpk_grad_block(self._p(Xr), nr, gi0, self._p(Xc), nc, gj0, D, _lib.darr(inv_ls), c,
                                           self._p(ar), self._p(ac), T, self._p(Kinv), Kinv.stride(0), rows, cols,
                                           self._p(ws), int(accum), self._st()), "gpk_grad_block")

    def grad_sum(self, ws, NB, D, empty):
        """(D + 2, 1) device tensor: per-feature sums, sum W K, tr W (the padded feature slots dropped)"""
        DP = int(self.lib.gpk_grad_padded_dim(D))
        out = self.t.zeros(DP + 2, dtype=self.t.float64, device=self.dev)
        if not empty:
            _lib.check(self.lib.gpk_grad_block_sum(self._p(ws), NB, NB, D, self._p(out), self._st()), "gpk_grad_block_sum")
        return self.t.cat([out[:D], out[DP:]]).reshape(-1, 1).contiguous()

    def _multi(self):
        return self.dist.is_initialized() and self.rank_world()[1] > 1

    def bcast(self, buf, src, kind="world"):
        """src is the rank within the world group; kind selects world / grid-row / grid-column communicator"""
        if not self._multi():
            return
        g = self.groups[kind]
        if kind != "world" and g is None:
            return                      # group of one
        self.bcast_bytes += buf.numel() * 8
        self.bcast_calls += 1
        self.dist.broadcast(buf, src=src, group=g)

    @contextlib.contextmanager
    def span(self, phase, k):
        """CUDA-event pair around the enclosed launches on the current stream (only while self.trace is a list)"""
        if self.trace is None:
            yield
            return
        st = self.t.cuda.current_stream()
        e0, e1 = self.t.cuda.Event(enable_timing=True), self.t.cuda.Event(enable_timing=True)
        e0.record(st)
        try:
            yield
        finally:
            e1.record(st)
            self.trace.append((phase, k, e0, e1, st.cuda_stream == self.side_stream.cuda_stream))

    def trace_summary(self):
        """seconds per phase (summed over panels) from the recorded spans, and -- for the spans of the main stream, in
        issue order -- the idle time in FRONT of each phase ("idle_before_s": the stream had nothing to run, i.e. the
        host had not issued it yet or an earlier wait had not been satisfied); synchronises"""
        self.t.cuda.synchronize()
        out = {}
        prev_end = None
        for phase, k, e0, e1, on_side in self.trace or []:
            d = out.setdefault(phase, {"seconds": 0.0, "spans": 0, "max_ms": 0.0, "idle_before_s": 0.0})
            ms = e0.elapsed_time(e1)
            d["seconds"] += ms * 1e-3; d["spans"] += 1; d["max_ms"] = max(d["max_ms"], ms)
            if not on_side:
                if prev_end is not None:
                    d["idle_before_s"] += max(0.0, prev_end.elapsed_time(e0)) * 1e-3
                prev_end = e1
        return out

    def allreduce_sum(self, buf):
        if self._multi():
            self.dist.all_reduce(buf, group=self.group)

    @contextlib.contextmanager
    def side(self):
        """enqueue on the side stream, ordered after everything issued so far on the main stream"""
        main = self.t.cuda.current_stream()
        self.side_stream.wait_stream(main)
        with self.t.cuda.stream(self.side_stream):
            yield

    def join_side(self):
        self.t.cuda.current_stream().wait_stream(self.side_stream)

    def sync(self):
        self.t.cuda.synchronize()

    def allreduce_max_int(self, v):
        if self._multi():
            tt = self.t.tensor([v], dtype=self.t.int64, device=self.dev)
            self.dist.all_reduce(tt, op=self.dist.ReduceOp.MAX, group=self.group)
            return int(tt[0])
        return v


class DistGP(DistGradMixin):
    """Distributed exact GP: factor K(theta) + alpha*I, solve for alpha_, evaluate the LML; its gradient through the
    distributed inverse (dist_grad.py: inverse(), gradient(), lml_grad())."""

    def __init__(self, X, y2d, jitter, NB=2048, ops=None, grid=None, lookahead=True):
        self.ops = ops if ops is not None else CudaOps()
        self.rank, self.P = self.ops.rank_world()
        self.Pr, self.Pc = grid if grid is not None else default_grid(self.P)
        if self.Pr * self.Pc != self.P:
            raise ValueError(f"grid {self.Pr}x{self.Pc} does not match world size {self.P}")
        self.pr, self.pc = self.ops.make_grid(self.Pr, self.Pc)
        X = np.ascontiguousarray(X, dtype=np.float64)
        y2d = np.ascontiguousarray(y2d, dtype=np.float64)
        self.N, self.D = X.shape
        self.T = y2d.shape[1]
        assert NB % 128 == 0
        self.NB = NB
        self.nb = nb = (self.N + NB - 1) // NB
        self.NP = nb * NB
        self.jitter = float(jitter)
        self.lookahead = bool(lookahead)
        o, Pr, Pc, pr, pc = self.ops, self.Pr, self.Pc, self.pr, self.pc
        self.my_rows = [i for i in range(nb) if i % Pr == pr]
        self.my_cols = [j for j in range(nb) if j % Pc == pc]
        self.owned = self.my_cols                      # block columns this rank holds tiles of
        self.nlr, self.nlc = len(self.my_rows), len(self.my_cols)
        # training inputs: replicated (column operand) and this rank's row tiles stacked (row operand)
        Xpad = np.zeros((self.NP, self.D))
        Xpad[: self.N] = X
        self.X_d = o.from_host(X)
        self.Xrow_d = o.from_host(Xpad.reshape(nb, NB, self.D)[self.my_rows].reshape(-1, self.D))
        self.nvalid_loc = sum(max(0, min(NB, self.N - i * NB)) for i in self.my_rows)
        ypad = np.zeros((self.NP, self.T))
        ypad[: self.N] = y2d
        self.y_d = o.from_host(ypad)
        # lower trapezoid: per owned block column, the owned row tiles i >= j stacked
        self.cols = {j: o.empty((self.nlr - self.lfirst(j)) * NB, NB) for j in self.my_cols}
        self.diag = [k for k in self.my_cols if k % Pr == pr]
        self.Ml = {k: o.empty(NB, NB) for k in self.diag}                         # L_kk^-1
        self.Mu = {k: o.empty(NB, NB) for k in self.diag}                         # L_kk^-T
        self.mlbuf = o.empty(NB, NB)
        # panel buffers (double-buffered for the look-ahead): L_ik for my rows, L_jk for my columns.
        # Lcol is ordered by the grid row that roots each piece of the transposed exchange.
        self.Lrow = [o.empty(max(self.nlr, 1) * NB, NB) for _ in range(2)]
        self.Lcol = [o.empty(max(self.nlc, 1) * NB, NB) for _ in range(2)]
        self.seg = [[j for j in self.my_cols if j % Pr == rp] for rp in range(Pr)]
        self.colpos = {}
        for rp in range(Pr):
            for j in self.seg[rp]:
                self.colpos[j] = len(self.colpos)
        o.info_slots(nb)
        self.gemm_flops = 0.0
        self.factored = False
        self.mt = self.kinv = None        # dist_grad.py allocates them on the first inverse()

    def _span(self, phase, k=-1):
        sp = getattr(self.ops, "span", None)
        return sp(phase, k) if sp is not None else contextlib.nullcontext()

    # -- index helpers --------------------------------------------------------------------------------
    def lfirst(self, m):
        """number of this rank's row tiles with global index < m (= local index of the first owned row >= m)"""
        if m <= self.pr:
            return 0
        return min(self.nlr, (m - self.pr + self.Pr - 1) // self.Pr)

    def owner(self, i, j=None):
        """world rank of tile (i, j); owner(k) is the diagonal tile (k, k)"""
        j = i if j is None else j
        return (i % self.Pr) * self.Pc + (j % self.Pc)

    def local_bytes(self):
        NB = self.NB
        tiles = sum(self.nlr - self.lfirst(j) for j in self.my_cols) + 2 * len(self.diag) + 1
        tiles += 2 * max(self.nlr, 1) + 2 * max(self.nlc, 1)
        return 8 * tiles * NB * NB

    # -- K generated in place -------------------------------------------------------------------------
    def assemble(self, inv_ls, c, noise):
        o, NB, N = self.ops, self.NB, self.N
        for j in self.my_cols:
            l0 = self.lfirst(j)
            rows = (self.nlr - l0) * NB
            if rows == 0:
                continue
            mq = max(0, self.nvalid_loc - l0 * NB)
            nc = max(0, min(NB, N - j * NB))
            o.gram_block(self.Xrow_d[l0 * NB:], mq, self.X_d[j * NB:], nc, self.D, inv_ls, c, self.cols[j], rows, NB)
            if j % self.Pr == self.pr:
                # diagonal tile: (c + noise) + jitter on real rows (gram gives c there), identity in the padding
                o.fix_diag(self.cols[j], NB, nc, noise, 1.0)          # (c + noise), padding -> 1
                o.fix_diag(self.cols[j], nc, nc, self.jitter, 1.0)    # ... + alpha  (_gpr.py:589)

    # -- right-looking factorisation with one-panel look-ahead ----------------------------------------
    def _panel_factor(self, k, b):
        """process column k % Pc: factor the diagonal tile, form L_ik = A_ik L_kk^-T into Lrow[b] (and in place)"""
        o, NB = self.ops, self.NB
        if self.pc != k % self.Pc:
            return
        Ck = self.cols[k]
        own_diag = self.pr == k % self.Pr
        if own_diag:
            with self._span("diag_potrf_inv", k):
                o.potrf_inv(Ck[:NB], self.Ml[k], self.Mu[k], NB, k)
                o.copy(self.Ml[k], self.mlbuf)
        with self._span("bcast_diag_inverse", k):
            o.bcast(self.mlbuf, self.owner(k), "col")
        off = NB if own_diag else 0
        rows = Ck.shape[0] - off
        if rows > 0:
            l0 = self.lfirst(k + 1)
            dst = self.Lrow[b][l0 * NB: l0 * NB + rows]
            with self._span("panel_gemm", k):
                o.gemm_nt(Ck[off:], self.mlbuf, dst, rows, NB, NB, 1.0, 0, b_tri=1)
                o.copy(dst, Ck[off:])
            self.gemm_flops += rows * float(NB) * NB

    def _panel_bcast(self, k, b):
        """row broadcast of L_ik, then the transposed exchange down the process columns (fills Lrow[b], Lcol[b])"""
        o, NB, Pr, Pc = self.ops, self.NB, self.Pr, self.Pc
        l0 = self.lfirst(k + 1)
        if self.nlr - l0 > 0:
            with self._span("bcast_panel_rows", k):
                o.bcast(self.Lrow[b][l0 * NB: self.nlr * NB], self.pr * Pc + k % Pc, "row")
        row3 = self.Lrow[b].reshape(-1, NB, NB)
        col3 = self.Lcol[b].reshape(-1, NB, NB)
        for rp in range(Pr):
            js = [j for j in self.seg[rp] if j > k]
            if not js:
                continue
            c0, c1 = self.colpos[js[0]], self.colpos[js[-1]] + 1
            with self._span("bcast_panel_cols", k):
                if self.pr == rp:
                    step = (js[1] - js[0]) // Pr if len(js) > 1 else 1
                    o.copy(row3[js[0] // Pr: js[-1] // Pr + 1: step], col3[c0:c1])
                o.bcast(self.Lcol[b][c0 * NB: c1 * NB], rp * Pc + self.pc, "col")

    def _update(self, k, b, js):
        """A[i >= j, j] -= L[i >= j, k] L[j, k]^T for the owned block columns in js"""
        o, NB = self.ops, self.NB
        for j in js:
            Cj = self.cols[j]
            M = Cj.shape[0]
            if M == 0:
                continue
            l0 = self.lfirst(j)
            p = self.colpos[j]
            o.gemm_nt(self.Lrow[b][l0 * NB:], self.Lcol[b][p * NB: (p + 1) * NB], Cj, M, NB, NB, -1.0, 1)
            self.gemm_flops += 2.0 * M * NB * NB

    def factor(self, inv_ls, c, noise):
        o, NB, nb = self.ops, self.NB, self.nb
        side = o.side if self.lookahead else contextlib.nullcontext
        self.theta = (tuple(float(v) for v in inv_ls), float(c), float(noise))
        with self._span("assemble"):
            self.assemble(inv_ls, c, noise)
        self._panel_factor(0, 0)
        with side():
            self._panel_bcast(0, 0)
        for k in range(nb):
            b = k % 2
            with self._span("wait_panel", k):  # on the main stream: how long it stalls for panel k's broadcasts
                o.join_side()                  # panel k has arrived in Lrow[b] / Lcol[b]
            if k + 1 < nb:
                if self.pc == (k + 1) % self.Pc:
                    with self._span("update_next_column", k):
                        self._update(k, b, [k + 1])
                self._panel_factor(k + 1, 1 - b)
                with side():                   # overlaps the rest of the trailing update below
                    self._panel_bcast(k + 1, 1 - b)
            with self._span("update_trailing", k):
                self._update(k, b, [j for j in self.my_cols if j > k + 1])
        o.join_side()
        infos = o.read_infos()
        bad = [k * NB + int(infos[k]) for k in self.diag if infos[k] != 0]
        BIG = 1 << 40
        r = o.allreduce_max_int(BIG - min(bad) if bad else 0)
        info = BIG - r if r > 0 else 0
        self.factored = info == 0
        return info

    # -- alpha = K^-1 y, log-det, lml -----------------------------------------------------------------
    def solve(self):
        assert self.factored
        o, NB, nb, T, Pr, Pc = self.ops, self.NB, self.nb, self.T, self.Pr, self.Pc
        nl = max(self.nlr, 1) * NB
        z = o.zeros(nl, T)                 # sum_k L_ik u_k over this rank's tiles, my rows (local order)
        u = o.zeros(self.NP, T)
        t = o.empty(NB, T)
        for k in range(nb):                # forward: L u = y
            blk = slice(k * NB, (k + 1) * NB)
            if self.pr == k % Pr:
                lk = k // Pr
                o.copy(z[lk * NB: (lk + 1) * NB], t)
            else:
                o.zero(t)
            o.allreduce_sum(t)
            if self.owner(k) == self.rank:
                rhs = self.y_d[blk] - t
                o.trimv(self.Ml[k], rhs, u[blk], upper=False)
            o.bcast(u[blk], self.owner(k))
            if self.pc == k % Pc:
                off = NB if self.pr == k % Pr else 0
                if self.cols[k].shape[0] - off > 0:
                    o.gemv(self.cols[k][off:], u[blk], z[self.lfirst(k + 1) * NB: self.nlr * NB], 1.0, 1.0,
                           transpose=False)
        alpha = o.zeros(self.NP, T)
        a_loc = o.zeros(nl, T)             # alpha restricted to my rows (local order)
        s = o.zeros(NB, T)
        for k in range(nb - 1, -1, -1):    # backward: L^T alpha = u
            blk = slice(k * NB, (k + 1) * NB)
            o.zero(s)
            if self.pc == k % Pc:
                off = NB if self.pr == k % Pr else 0
                if self.cols[k].shape[0] - off > 0:
                    o.gemv(self.cols[k][off:], a_loc[self.lfirst(k + 1) * NB: self.nlr * NB], s, 1.0, 0.0,
                           transpose=True)
            o.allreduce_sum(s)
            if self.owner(k) == self.rank:
                rhs = u[blk] - s
                o.trimv(self.Mu[k], rhs, alpha[blk], upper=True)
            o.bcast(alpha[blk], self.owner(k))
            if self.pr == k % Pr:
                lk = k // Pr
                o.copy(alpha[blk], a_loc[lk * NB: (lk + 1) * NB])
        self.alpha_d, self.u_d = alpha, u
        return alpha

    def lml(self):
        o = self.ops
        acc = o.zeros(1, 1)
        for k in self.diag:
            o.logdiag(self.cols[k], self.NB, acc)
        o.allreduce_sum(acc)
        logdet = float(o.to_host(acc)[0, 0])
        y = o.to_host(self.y_d)[: self.N]
        a = o.to_host(self.alpha_d)[: self.N]
        lml = -0.5 * np.einsum("ik,ik->k", y, a) - logdet - self.N / 2 * np.log(2 * np.pi)
        return float(lml.sum())

    def alpha_host(self):
        return self.ops.to_host(self.alpha_d)[: self.N]

    # -- predict from the distributed factor ----------------------------------------------------------
    def predict(self, Xq, return_std=False, chunk_rows=256):
        """mean (M, T) [, std (M, T)] for query rows Xq (the SAME rows on every rank: the work is shared by tiles,
        not by query rows), sklearn _gpr.py:446-500 without target normalisation.

        Queries are a row tile of the factorisation: with Q = k(Xq, X) (m x N), the variance needs W = Q L^-T, block
        column k of which is  W_k = (Q_k - sum_{j<k} W_j L_kj^T) L_kk^-T  -- the panel formula of `_panel_factor`, so
        every product is one NT GEMM on the DMMA kernel.  Per block k: the owner of the diagonal tile forms Q_k in place
        from the replicated X (never communicated), adds Q_k alpha_k to the mean, forms W_k and broadcasts it; the ranks
        of process column k % Pc add W_k L_ik^T to their running sums for all owned rows i > k in ONE GEMM over the
        stacked column; one all-reduce per block brings the sums of block k together.  sum_k rowsum(W_k^2) is taken
        as the diagonal of sum_k W_k W_k^T (m x m, one more small GEMM per block).  m N^2 flops per chunk, shared by
        the P ranks."""
        assert self.factored and hasattr(self, "alpha_d"), "factor() and solve() first"
        o, NB, nb, T, Pr, Pc = self.ops, self.NB, self.nb, self.T, self.Pr, self.Pc
        inv_ls, c, noise = self.theta
        Xq = np.ascontiguousarray(Xq, dtype=np.float64)
        if Xq.ndim != 2 or Xq.shape[1] != self.D:
            raise ValueError(f"X has {Xq.shape[1] if Xq.ndim == 2 else '?'} features, the training set {self.D}")
        M = Xq.shape[0]
        chunk_rows = max(128, (int(chunk_rows) // 128) * 128)
        mean = np.empty((M, T))
        std = np.empty((M, T)) if return_std else None
        neg = 0
        for s0 in range(0, M, chunk_rows):
            m = min(chunk_rows, M - s0)
            MP = ((m + 127) // 128) * 128
            xq = o.from_host(Xq[s0:s0 + m])
            Qk, acc = o.empty(MP, NB), o.zeros(MP, T)
            if return_std:
                Z = o.zeros(MP, max(self.nlr, 1) * NB)      # running sums for this rank's row tiles (local order)
                W, t, G = o.empty(MP, NB), o.empty(MP, NB), o.zeros(MP, MP)
            for k in range(nb):
                blk = slice(k * NB, (k + 1) * NB)
                own = self.owner(k) == self.rank
                if return_std:
                    if self.pr == k % Pr:
                        lk = k // Pr
                        o.copy(Z[:, lk * NB:(lk + 1) * NB], t)
                    else:
                        o.zero(t)
                    o.allreduce_sum(t)
                if own:
                    nc = max(0, min(NB, self.N - k * NB))
                    o.gram_block(xq, m, self.X_d[k * NB:], nc, self.D, inv_ls, c, Qk, MP, NB)
                    o.gemv(Qk, self.alpha_d[blk], acc, 1.0, 1.0, transpose=False)
                    if return_std:
                        rhs = Qk - t
                        o.gemm_nt(rhs, self.Ml[k], W, MP, NB, NB, 1.0, 0, b_tri=1)
                if return_std:
                    o.bcast(W, self.owner(k))
                    o.gemm_nt(W, W, G, MP, MP, NB, 1.0, 1)
                    if self.pc == k % Pc:
                        off = NB if self.pr == k % Pr else 0
                        rows = self.cols[k].shape[0] - off
                        if rows > 0:
                            l0 = self.lfirst(k + 1)
                            o.gemm_nt(W, self.cols[k][off:], Z[:, l0 * NB:], MP, rows, NB, 1.0, 1)
            o.allreduce_sum(acc)
            mean[s0:s0 + m] = o.to_host(acc)[:m]
            if return_std:
                var = (c + noise) - np.diag(o.to_host(G))[:m]
                neg += int((var < 0).sum())
                std[s0:s0 + m] = np.sqrt(np.maximum(var, 0.0))[:, None]
        if return_std:
            if neg:
                import warnings
                warnings.warn("Predicted variances smaller than 0. Setting those variances to 0.")
            return mean, std
        return mean
