"""Hyper-parameter gradient of the log marginal likelihood from the DISTRIBUTED factor (the step
that makes an N=200k fit a fit, not one evaluation) -- sklearn _gpr.py:629-651 needs K^-1:

    d lml / d theta_k = 1/2 sum_t sum_ij (alpha_t alpha_t^T - K^-1)_ij  dK_ji / d theta_k

K^-1 = L^-T L^-1 is formed on the same Pr x Pc block-cyclic grid as the factor, tile (i, j) (i >= j) on grid position
(i % Pr, j % Pc), in two right-looking sweeps whose trailing updates are NT GEMMs on the DMMA kernel:

  TRTRI  (N^3/3)  M = L^-1 by block forward substitution, stored TRANSPOSED tile by tile (Mt(i,j) = M(i,j)^T) so that
                  every product is A B^T on row-major operands.  Row k is finished with the diagonal inverse the
                  factorisation left on the diagonal owner (Mt(k,j) = Rt(k,j) L_kk^-T), broadcast down the process
                  columns, and pushed into the rows below with the panel L(i>k, k) (broadcast along process rows):
                      Rt(i,j) -= Mt(k,j) L(i,k)^T        for every owned tile i > k >= j.
                  Column j's tiles sit side by side in one buffer (NB x R_j NB), so the update of a whole owned column
                  is ONE GEMM (M = NB, N = rows below k, K = NB).
  LAUUM  (N^3/3)  K^-1(i,j) += Mt(k,i) Mt(k,j)^T for j <= i <= k, sweep over k: row k of Mt goes down the process
                  columns (right operand) and, tile by tile, along the process rows (left operand, stacked in local row
                  order), then one GEMM per owned column into kinv[j], laid out like the factor's column buffers.

L itself is not touched (the factor stays valid for predict); M and K^-1 take two more trapezoids (at N=200k on 8
GPUs: 3 x 20 GB per rank).  The gradient contraction then runs tile by tile with K recomputed from the replicated X
(`gpk_grad_block`: the single-GPU gradient kernel with global row/column offsets), partial sums reduced on the device
in a fixed order, one all-reduce of D+2 doubles at the end.
"""
from __future__ import annotations

import contextlib

import numpy as np


class DistGradMixin:
    """inverse() and gradient() for dist_chol.DistGP (expects its layout helpers: lfirst, owner, cols, Ml, Mu, ...)"""

    # -- helpers ------------------------------------------------------------------------------------------
    def _cols_le(self, k):
        """owned block columns j <= k, ascending (position in the row buffers = index in this list)"""
        return [j for j in self.my_cols if j <= k]

    def _rows_le(self, k):
        """owned block rows i <= k, ascending; local index of row i is i // Pr"""
        return [i for i in self.my_rows if i <= k]

    def _alloc_inverse(self):
        o, NB = self.ops, self.NB
        nlc = max(self.nlc, 1)
        if getattr(self, "mt", None) is None:
            # Mt: per owned column j the tiles Mt(i, j), owned i >= j, SIDE BY SIDE: (NB, R_j * NB)
            self.mt = {j: o.empty(NB, max(self.nlr - self.lfirst(j), 1) * NB) for j in self.my_cols}
            self.kinv = {j: o.empty(max(self.nlr - self.lfirst(j), 1) * NB, NB) for j in self.my_cols}
            self.rowbuf = [o.empty(nlc * NB, NB) for _ in range(2)]            # Mt(k, j) for my columns, by position
            self.rowstack = [o.empty(max(self.nlr, 1) * NB, NB) for _ in range(2)]   # Mt(k, i) for my rows, local order
            self.mlrow = o.empty(NB, NB)
            self.tmp_tile = o.empty(NB, NB)

    def release_inverse(self):
        self.mt = self.kinv = self.rowbuf = self.rowstack = None

    # -- TRTRI ----------------------------------------------------------------------------------------------
    def _finish_row(self, k, b):
        """Mt(k, j) = Rt(k, j) L_kk^-T for the owned j <= k (process row k % Pr), into rowbuf[b] and back into mt;
        then the broadcast of the row down every process column."""
        o, NB, Pr, Pc = self.ops, self.NB, self.Pr, self.Pc
        js = self._cols_le(k)
        if self.pr == k % Pr:
            own_diag = self.pc == k % Pc
            with self._span("trtri_finish_row", k):
                if own_diag:
                    o.copy(self.Ml[k], self.mlrow)
                o.bcast(self.mlrow, self.owner(k), "row")
                lk = k // Pr
                for p, j in enumerate(js):
                    dst = self.rowbuf[b][p * NB:(p + 1) * NB]
                    slot = self.mt[j][:, (lk - self.lfirst(j)) * NB:(lk - self.lfirst(j) + 1) * NB]
                    if j == k:
                        # Mt(k,k) = (L_kk^-1)^T; potrf_inv leaves the 128-blocks below Mup's diagonal unwritten
                        # and this tile is a FULL operand of the updates below: take the triangle only
                        o.copy_triu(self.Mu[k], dst)
                    else:
                        o.gemm_nt(slot, self.mlrow, dst, NB, NB, NB, 1.0, 0, b_tri=1)
                        self.inv_flops += float(NB) ** 3
                    o.copy(dst, slot)
        if js:
            with self._span("trtri_bcast_row", k):
                o.bcast(self.rowbuf[b][: len(js) * NB], (k % Pr) * Pc + self.pc, "col")

    def _bcast_L_panel(self, k, b):
        """L(i > k, k) for my rows into Lrow[b] (local row order), broadcast along the process row"""
        o, NB, Pc = self.ops, self.NB, self.Pc
        l0 = self.lfirst(k + 1)
        if self.nlr - l0 <= 0:
            return
        with self._span("trtri_bcast_L_panel", k):
            if self.pc == k % Pc:
                off = NB if self.pr == k % self.Pr else 0
                o.copy(self.cols[k][off:], self.Lrow[b][l0 * NB: self.nlr * NB])
            o.bcast(self.Lrow[b][l0 * NB: self.nlr * NB], self.pr * Pc + k % Pc, "row")

    def _trtri(self):
        o, NB, nb = self.ops, self.NB, self.nb
        side = o.side if self.lookahead and hasattr(o, "side") else contextlib.nullcontext
        with side():
            self._bcast_L_panel(0, 0)                 # L is static: its panels are prefetched one step ahead
        for k in range(nb):
            b = k % 2
            self._finish_row(k, b)
            o.join_side()                              # panel k has arrived in Lrow[b]
            if k + 1 < nb:
                with side():
                    self._bcast_L_panel(k + 1, 1 - b)
            l0 = self.lfirst(k + 1)
            rows = (self.nlr - l0) * NB
            if rows <= 0:
                continue
            with self._span("trtri_update", k):
                for p, j in enumerate(self._cols_le(k)):
                    c0 = (l0 - self.lfirst(j)) * NB
                    o.gemm_nt(self.rowbuf[b][p * NB:(p + 1) * NB], self.Lrow[b][l0 * NB: self.nlr * NB],
                              self.mt[j][:, c0: c0 + rows], NB, rows, NB, -1.0, 0 if j == k else 1)
                    self.inv_flops += 2.0 * rows * NB * NB
        o.join_side()

    # -- LAUUM ----------------------------------------------------------------------------------------------
    def _lauum_fetch(self, k, b):
        """row k of Mt: rowbuf[b] <- Mt(k, j) for my columns j <= k (down the process columns);
        rowstack[b] <- Mt(k, i) for my rows i <= k in local order (tile by tile along the process row)"""
        o, NB, Pr, Pc = self.ops, self.NB, self.Pr, self.Pc
        js = self._cols_le(k)
        if js:
            with self._span("lauum_bcast_cols", k):
                if self.pr == k % Pr:
                    lk = k // Pr
                    for p, j in enumerate(js):
                        c0 = (lk - self.lfirst(j)) * NB
                        o.copy(self.mt[j][:, c0: c0 + NB], self.rowbuf[b][p * NB:(p + 1) * NB])
                o.bcast(self.rowbuf[b][: len(js) * NB], (k % Pr) * Pc + self.pc, "col")
        pos = {j: p for p, j in enumerate(js)}
        rows = self._rows_le(k)
        if rows:
            with self._span("lauum_bcast_rows", k):
                for i in rows:
                    li = i // Pr
                    dst = self.rowstack[b][li * NB:(li + 1) * NB]
                    if self.pc == i % Pc:
                        o.copy(self.rowbuf[b][pos[i] * NB:(pos[i] + 1) * NB], dst)
                    o.bcast(dst, self.pr * Pc + i % Pc, "row")

    def _lauum(self):
        o, NB, nb = self.ops, self.NB, self.nb
        side = o.side if self.lookahead and hasattr(o, "side") else contextlib.nullcontext
        for j in self.my_cols:
            o.zero(self.kinv[j])
        with side():
            self._lauum_fetch(0, 0)                    # Mt is static during this sweep: prefetched one step ahead
        for k in range(nb):
            b = k % 2
            o.join_side()
            if k + 1 < nb:
                with side():
                    self._lauum_fetch(k + 1, 1 - b)
            l1 = self.lfirst(k + 1)                    # my rows i <= k are the local rows [0, l1)
            with self._span("lauum_update", k):
                for p, j in enumerate(self._cols_le(k)):
                    lj = self.lfirst(j)
                    rows = (l1 - lj) * NB
                    if rows <= 0:
                        continue
                    o.gemm_nt(self.rowstack[b][lj * NB: l1 * NB], self.rowbuf[b][p * NB:(p + 1) * NB],
                              self.kinv[j][:rows], rows, NB, NB, 1.0, 1)
                    self.inv_flops += 2.0 * rows * NB * NB
            # the buffers of step k are rewritten by the prefetch of step k+2: that prefetch is issued after the next
            # join_side(), i.e. behind this step's GEMMs on the main stream (side() waits for the main stream)
        o.join_side()

    def inverse(self):
        """K^-1 tiles into self.kinv[j] (rows i >= j stacked like self.cols[j]); needs factor()"""
        assert self.factored, "factor() first"
        self._alloc_inverse()
        self.inv_flops = 0.0
        self._trtri()
        self._lauum()

    # -- gradient -------------------------------------------------------------------------------------------
    def gradient(self):
        """(g_c, g_ls[D], g_noise) = d lml / d log(c, length_scale, noise), sklearn _gpr.py:629-651 summed over the T
        targets; needs factor(), solve() and inverse()"""
        assert self.factored and hasattr(self, "alpha_d") and getattr(self, "kinv", None) is not None
        o, NB, N, D, T = self.ops, self.NB, self.N, self.D, self.T
        inv_ls, c, noise = self.theta
        ws = o.grad_ws(NB, D)
        first = True
        with self._span("gradient"):
            for j in self.my_cols:
                nc = max(0, min(NB, N - j * NB))
                for i in self.my_rows:
                    if i < j:
                        continue
                    li = i // self.Pr
                    nr = max(0, min(NB, N - i * NB))
                    if nr == 0 or nc == 0:
                        continue
                    r0 = (li - self.lfirst(j)) * NB
                    o.grad_block(self.Xrow_d[li * NB:(li + 1) * NB], nr, i * NB, self.X_d[j * NB:], nc, j * NB, D, inv_ls,
                                 c, self.alpha_d[i * NB:(i + 1) * NB], self.alpha_d[j * NB:(j + 1) * NB], T,
                                 self.kinv[j][r0: r0 + NB], NB, NB, ws, 0 if first else 1)
                    first = False
            sums = o.grad_sum(ws, NB, D, first)           # device buffer [D dims, c, trace] (zeros if nothing owned)
        o.allreduce_sum(sums)
        s = np.asarray(o.to_host(sums)).reshape(-1)
        return 0.5 * float(s[D]), 0.5 * s[:D].copy(), 0.5 * float(noise) * float(s[D + 1])

    def lml_grad(self, inv_ls, c, noise):
        """one objective evaluation at the distributed size: (lml, g_c, g_ls, g_noise); (-inf, zeros) when K is not
        positive definite (sklearn _gpr.py:592-593)"""
        info = self.factor(inv_ls, c, noise)
        if info != 0:
            return -np.inf, 0.0, np.zeros(self.D), 0.0
        self.solve()
        lml = self.lml()
        self.inverse()
        g_c, g_ls, g_noise = self.gradient()
        return lml, g_c, g_ls, g_noise
