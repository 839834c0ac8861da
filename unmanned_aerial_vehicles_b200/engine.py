"""Device-side state of one exact GP (one kernel shared by T targets) on one B200.

Host code is plumbing only: torch allocates HBM and supplies the stream; every flop runs in libgpk.so
through the C ABI (include/gpk.h).  Mirrors the arithmetic of sklearn/gaussian_process/_gpr.py
(:349-367 final factorisation, :541-656 LML + gradient, :446-500 predict).
"""
from __future__ import annotations

import ctypes as C
import threading

import numpy as np

from . import _lib


def _torch_cuda():
    import torch

    if not torch.cuda.is_available():
        raise _lib.GpkUnavailable("no CUDA device: the GP hot path has no CPU fallback")
    return torch


class DeviceGP:
    """Training set resident in HBM plus the three NP x NP factorisation buffers."""

    def __init__(self, X: np.ndarray, y2d: np.ndarray, jitter: float, device=None):
        torch = _torch_cuda()
        self.torch = torch
        self.lib = _lib.load()
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        X = np.ascontiguousarray(X, dtype=np.float64)
        y2d = np.ascontiguousarray(y2d, dtype=np.float64)
        self.N, self.D = X.shape
        self.T = y2d.shape[1]
        if self.D > 32:
            raise ValueError("libgpk supports at most 32 features")
        self.NP = int(self.lib.gpk_pad(self.N))
        self.jitter = float(jitter)
        with torch.cuda.device(self.device):
            # inputs come from pinned host memory so the copies can overlap the first launch
            self.X_d = torch.from_numpy(X).pin_memory().to(self.device, non_blocking=True)
            self.y_d = torch.zeros((self.NP, self.T), dtype=torch.float64, device=self.device)
            self.y_d[: self.N].copy_(torch.from_numpy(y2d).pin_memory(), non_blocking=True)
            self.u_d = torch.zeros((self.NP, self.T), dtype=torch.float64, device=self.device)
            self.alpha_d = torch.zeros((self.NP, self.T), dtype=torch.float64, device=self.device)
            nws = int(self.lib.gpk_grad_workspace_doubles(self.N, self.D))
            self.ws_d = torch.empty(nws, dtype=torch.float64, device=self.device)
            self.out_len = int(self.lib.gpk_lml_out_len(self.D, self.T))
            self.out_d = torch.zeros(self.out_len, dtype=torch.float64, device=self.device)
            self.out_h = torch.zeros(self.out_len + 1, dtype=torch.float64).pin_memory()
            self.info_d = torch.zeros(2, dtype=torch.int32, device=self.device)
        self.A_d = self.Mlow_d = self.Mup_d = None
        self.factored = None  # (inv_ls tuple, c, noise) of the factor currently held in Mlow/alpha
        self.a_holds_L = False
        self.launches = 0
        self._plock = threading.Lock()   # predict() shares cached staging buffers: one caller at a time

    # -- helpers ---------------------------------------------------------------------------
    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream)

    def _bufs(self):
        if self.A_d is None:
            t = self.torch
            shape = (self.NP, self.NP)
            self.A_d = t.empty(shape, dtype=t.float64, device=self.device)
            self.Mlow_d = t.empty(shape, dtype=t.float64, device=self.device)
            self.Mup_d = t.empty(shape, dtype=t.float64, device=self.device)

    def _chunk_rows_default(self) -> int:
        return int(min(16384, max(128, (2 << 30) // (self.NP * 8)))) // 128 * 128

    @staticmethod
    def _p(t):
        return C.c_void_p(t.data_ptr())

    def upload(self, X_host, y_host):
        """Replace the resident training inputs / targets from host tensors of the same shape (pinned ones copy
        asynchronously on the engine's stream).  Invalidates the factor held."""
        t = self.torch
        X_host = X_host if t.is_tensor(X_host) else t.from_numpy(np.ascontiguousarray(X_host, dtype=np.float64))
        y_host = y_host if t.is_tensor(y_host) else t.from_numpy(np.ascontiguousarray(y_host, dtype=np.float64))
        if tuple(X_host.shape) != (self.N, self.D) or tuple(y_host.reshape(self.N, -1).shape) != (self.N, self.T):
            raise ValueError("upload: shapes must match the training set the engine was built for")
        with t.cuda.device(self.device):
            self.X_d.copy_(X_host, non_blocking=True)
            self.y_d[: self.N].copy_(y_host.reshape(self.N, self.T), non_blocking=True)
        self.factored = None
        self.a_holds_L = False

    # -- log marginal likelihood (+ gradient) ------------------------------------------------
    def lml(self, inv_ls, c: float, noise: float, with_grad: bool):
        """Returns (lml, g_c, g_ls[D], g_noise); lml = -inf when K is not positive definite
        (sklearn _gpr.py:592-593).  Gradients are w.r.t. log-parameters."""
        t = self.torch
        self._bufs()
        with t.cuda.device(self.device):
            rc = self.lib.gpk_lml_eval(
                self._p(self.X_d), self.N, self.D, _lib.darr(inv_ls), c, noise, self.jitter, self._p(self.y_d), self.T,
                self._p(self.A_d), self._p(self.Mlow_d), self._p(self.Mup_d), self._p(self.u_d), self._p(self.alpha_d),
                self.NP, int(bool(with_grad)), self._p(self.ws_d), self._p(self.out_d), self._p(self.info_d),
                self._stream())
            _lib.check(rc, "gpk_lml_eval")
            # one device->host read per evaluation: [out..., info]
            self.out_h[: self.out_len].copy_(self.out_d, non_blocking=True)
            info = self.info_d[:1].to("cpu", non_blocking=False)
        self.factored = None
        self.a_holds_L = False
        out = self.out_h[: self.out_len].numpy().copy()
        T, D = self.T, self.D
        if int(info[0]) != 0 or not np.isfinite(out[0]):
            return -np.inf, 0.0, np.zeros(D), 0.0
        return float(out[0]), float(out[2 + T]), out[3 + T: 3 + T + D].copy(), float(out[3 + T + D])

    # -- final factorisation ---------------------------------------------------------------
    def factor(self, inv_ls, c: float, noise: float):
        """gram -> potrf+inverse -> potrs.  Returns LAPACK-style info (0 ok)."""
        t = self.torch
        self._bufs()
        key = (tuple(float(v) for v in inv_ls), float(c), float(noise))
        with t.cuda.device(self.device):
            st = self._stream()
            _lib.check(self.lib.gpk_gram(self._p(self.X_d), self.N, self.D, _lib.darr(inv_ls), c, noise, self.jitter,
                                         self._p(self.A_d), self.NP, 1, st), "gpk_gram")
            _lib.check(self.lib.gpk_potrf_inv(self._p(self.A_d), self._p(self.Mlow_d), self._p(self.Mup_d), self.NP,
                                              self._p(self.info_d), st), "gpk_potrf_inv")
            _lib.check(self.lib.gpk_potrs(self._p(self.Mlow_d), self._p(self.Mup_d), self.NP, self._p(self.y_d),
                                          self._p(self.u_d), self._p(self.alpha_d), self.T, st), "gpk_potrs")
            info = int(self.info_d[:1].cpu()[0])
        self.factored = key if info == 0 else None
        self.a_holds_L = info == 0
        return info

    def L_host(self) -> np.ndarray:
        assert self.a_holds_L
        t = self.torch
        with t.cuda.device(self.device):
            dst = t.empty((self.N, self.N), dtype=t.float64, device=self.device)
            _lib.check(self.lib.gpk_extract_lower(self._p(self.A_d), self.NP, self._p(dst), self.N, self._stream()),
                       "gpk_extract_lower")
            return dst.cpu().numpy()

    def alpha_host(self) -> np.ndarray:
        return self.alpha_d[: self.N].cpu().numpy()

    def gram_host(self, inv_ls, c, noise, jitter=0.0) -> np.ndarray:
        """Full K (for parity tests of the Gram kernel)."""
        t = self.torch
        with t.cuda.device(self.device):
            K = t.empty((self.NP, self.NP), dtype=t.float64, device=self.device)
            _lib.check(self.lib.gpk_gram(self._p(self.X_d), self.N, self.D, _lib.darr(inv_ls), c, noise, jitter,
                                         self._p(K), self.NP, 0, self._stream()), "gpk_gram")
            return K[: self.N, : self.N].cpu().numpy()

    # -- predict -------------------------------------------------------------------------------
    def _predict_ws(self, MPmax: int, mean_only: bool = False):
        """Device scratch for query chunks of up to MPmax rows (gpk_predict carves it).  Small ones (the 1..128-row
        control-tick calls of simple_gp.py:187-201 / mpc.py:1490-1509) are kept between calls, so a predict is
        ONE C-ABI call: H2D, the launches, D2H, one stream sync, no allocation on the way.  mean_only: the fused
        kernel's scratch (no M x N cross-Gram: MPmax*(D + T*(1 + splits)) doubles)."""
        slot = "_pws_mean" if mean_only else "_pws"
        ws = getattr(self, slot, None)
        if ws is not None and ws[0] >= MPmax:
            return ws
        t = self.torch
        if mean_only:
            n = int(self.lib.gpk_predict_mean_workspace_doubles(MPmax, self.N, self.D, self.T))
        else:
            n = int(self.lib.gpk_predict_workspace_doubles(MPmax, self.NP, self.D, self.T))
        ws = (MPmax, t.empty(n, dtype=t.float64, device=self.device), n, np.zeros(1, dtype=np.int32))
        setattr(self, slot, ws if n * 8 <= (256 << 20) else None)
        return ws

    def _target_scale(self, y_mean, y_std):
        key = (tuple(np.broadcast_to(np.asarray(y_mean, dtype=np.float64), (self.T,)).tolist()),
               tuple(np.broadcast_to(np.asarray(y_std, dtype=np.float64), (self.T,)).tolist()))
        cached = getattr(self, "_yscale", None)
        if cached is None or cached[0] != key:
            t = self.torch
            cached = (key, t.tensor(key[0], dtype=t.float64, device=self.device),
                      t.tensor(key[1], dtype=t.float64, device=self.device))
            self._yscale = cached
        return cached[1], cached[2]

    def predict(self, Xq: np.ndarray, y_mean, y_std, return_std: bool, return_cov: bool = False, chunk_rows=None):
        """mean (M,T) [, std (M,T) | cov (M,M) un-normalised] for the factor currently held."""
        assert self.factored is not None, "factor() first"
        t = self.torch
        inv_ls, c, noise = self.factored
        Xq = np.ascontiguousarray(Xq, dtype=np.float64)
        M = Xq.shape[0]
        NP, T = self.NP, self.T
        mean_only = not return_std and not return_cov
        if chunk_rows is None:   # mean only: nothing M x N is stored, so the chunk is bounded by the staging copy alone
            chunk_rows = 131072 if mean_only else self._chunk_rows_default()
        chunk_rows = max(128, (chunk_rows // 128) * 128)
        if return_cov:
            chunk_rows = ((M + 127) // 128) * 128
        mean = np.empty((M, T))
        std = np.empty((M, T)) if return_std else None
        cov = None
        neg_total = 0
        with self._plock, t.cuda.device(self.device):
            st = self._stream()
            ym, ys = self._target_scale(y_mean, y_std)
            inv_h = _lib.darr(inv_ls)
            if not return_cov:
                MPmax, ws_d, ws_n, neg_h = self._predict_ws(min(chunk_rows, ((M + 127) // 128) * 128), mean_only)
                for s in range(0, M, chunk_rows):
                    m = min(chunk_rows, M - s)
                    rc = self.lib.gpk_predict(
                        Xq[s:s + m].ctypes.data, m, self.D, self._p(self.X_d), self.N, inv_h, c, noise,
                        self._p(self.alpha_d), T, self._p(self.Mlow_d), NP, self._p(ym), self._p(ys),
                        mean[s:s + m].ctypes.data, std[s:s + m].ctypes.data if return_std else None,
                        neg_h.ctypes.data, self._p(ws_d), ws_n, st)
                    _lib.check(rc, "gpk_predict")
                    if return_std:
                        neg_total += int(neg_h[0])
                return mean, std, None, neg_total
            # covariance (sample_y): one chunk holding every query row
            m = M
            MP = ((m + 127) // 128) * 128
            Ks = t.empty((MP, NP), dtype=t.float64, device=self.device)
            mean_d = t.empty((MP, T), dtype=t.float64, device=self.device)
            xq_d = t.as_tensor(Xq, device=self.device)
            _lib.check(self.lib.gpk_cross_gram(self._p(xq_d), m, self._p(self.X_d), self.N, self.D, inv_h, c,
                                               self._p(Ks), MP, NP, st), "gpk_cross_gram")
            _lib.check(self.lib.gpk_predict_mean(self._p(Ks), m, self.N, NP, self._p(self.alpha_d), T, self._p(ym),
                                                 self._p(ys), self._p(mean_d), st), "gpk_predict_mean")
            V = t.empty((MP, NP), dtype=t.float64, device=self.device)
            _lib.check(self.lib.gpk_gemm_nt(self._p(Ks), NP, self._p(self.Mlow_d), NP, self._p(V), NP, MP, NP,
                                            NP, 1.0, 0, 0, 1, 0, st), "gpk_gemm_nt")
            Kqq = t.empty((MP, MP), dtype=t.float64, device=self.device)
            _lib.check(self.lib.gpk_cross_gram(self._p(xq_d), m, self._p(xq_d), m, self.D, inv_h, c,
                                               self._p(Kqq), MP, MP, st), "gpk_cross_gram")
            _lib.check(self.lib.gpk_gemm_nt(self._p(V), NP, self._p(V), NP, self._p(Kqq), MP, MP, MP, NP,
                                            -1.0, 1, 0, 0, 0, st), "gpk_gemm_nt")
            cov = Kqq[:m, :m].cpu().numpy()
            cov[np.diag_indices_from(cov)] += noise
            mean[...] = mean_d[:m].cpu().numpy()
        return mean, std, cov, neg_total


    def predict_with_metrics(self, Xq: np.ndarray, y_mean, y_std, R_true: np.ndarray, chunk_rows: int = 131072):
        """Mean-only predict over all of Xq with the residual-metric reductions of evaluate_gp_offline.py:277-312 run
        behind it ON THE DEVICE: the predicted residuals never leave HBM between the fused mean kernel and the
        reductions; the host receives the (M, T) mean and gpk_residual_metrics' 46 sums (see include/gpk.h)."""
        assert self.factored is not None, "factor() first"
        t = self.torch
        inv_ls, c, _ = self.factored
        Xq = np.ascontiguousarray(Xq, dtype=np.float64)
        R_true = np.ascontiguousarray(R_true, dtype=np.float64)
        M, T = Xq.shape[0], self.T
        if R_true.shape != (M, T):
            raise ValueError(f"R_true must be {(M, T)}, got {R_true.shape}")
        if T > 8:
            raise ValueError("device-side residual metrics handle at most 8 components")
        chunk_rows = max(128, (int(chunk_rows) // 128) * 128)
        with self._plock, t.cuda.device(self.device):
            st = self._stream()
            ym, ys = self._target_scale(y_mean, y_std)
            inv_h = _lib.darr(inv_ls)
            MPc = min(chunk_rows, ((M + 127) // 128) * 128)
            nws = int(self.lib.gpk_predict_mean_workspace_doubles(MPc, self.N, self.D, T))
            ws = t.empty(nws, dtype=t.float64, device=self.device)
            xq_d = t.as_tensor(Xq, device=self.device)
            rt_d = t.as_tensor(R_true, device=self.device)
            mean_d = t.empty((M + MPc, T), dtype=t.float64, device=self.device)
            for s in range(0, M, chunk_rows):
                m = min(chunk_rows, M - s)
                MP = ((m + 127) // 128) * 128
                _lib.check(self.lib.gpk_predict_mean_fused(self._p(xq_d[s:]), m, MP, self._p(self.X_d), self.N, self.D,
                                                           inv_h, c, self._p(self.alpha_d), T, self._p(ym), self._p(ys),
                                                           self._p(ws), self._p(mean_d[s:]), st), "gpk_predict_mean_fused")
            nout = int(self.lib.gpk_residual_metrics_out_len())
            mws = t.empty(int(self.lib.gpk_residual_metrics_workspace_doubles(M)), dtype=t.float64, device=self.device)
            out_d = t.zeros(nout, dtype=t.float64, device=self.device)
            _lib.check(self.lib.gpk_residual_metrics(self._p(rt_d), self._p(mean_d), M, T, self._p(mws), self._p(out_d),
                                                     st), "gpk_residual_metrics")
            return mean_d[:M].cpu().numpy(), out_d.cpu().numpy()


def predict_stacked(engines, Xqs, scales, return_std: bool):
    """Several factored engines on ONE device answer the same number of query rows each in ONE C-ABI call
    (gpk_predict_stacked: one stream sync for all of them) -- the per-output members of the reference's schema B / C
    pickles, each seeing the query through its own input scaler (pretrained_gp.py:52-98, gp_enhanced_mpc.py:150-183).

    engines: DeviceGP list; Xqs: one (M, D_k) float64 array per engine; scales: (y_mean, y_std) per engine.
    Returns [(mean (M,T_k), std (M,T_k) | None, clamped-row count)] in order, or None when the batch does not qualify
    (different devices, more rows than one chunk, more than MAX_STACKED models) -- the caller then
    loops over DeviceGP.predict, the same kernels one model at a time."""
    K = len(engines)
    if K < 1 or K > _lib.MAX_STACKED or len({id(e) for e in engines}) != K:
        return None
    dev = engines[0].device
    M = int(Xqs[0].shape[0])
    if M < 1 or any(e.device != dev or e.factored is None for e in engines):
        return None
    if any(x.shape[0] != M for x in Xqs) or any(M > e._chunk_rows_default() for e in engines):
        return None
    t = engines[0].torch
    MP = ((M + 127) // 128) * 128
    arr = (_lib.StackedModel * K)()
    keep, out = [], []
    locks = sorted(engines, key=id)          # one global order, so two stacked callers cannot deadlock
    for e in locks:
        e._plock.acquire()
    try:
        with t.cuda.device(dev):
            st = engines[0]._stream()
            for k, (e, Xq, (y_mean, y_std)) in enumerate(zip(engines, Xqs, scales)):
                Xq = np.ascontiguousarray(Xq, dtype=np.float64)
                if Xq.shape[1] != e.D:
                    raise ValueError(f"model {k}: query has {Xq.shape[1]} features, the training set {e.D}")
                inv_ls, c, noise = e.factored
                ym, ys = e._target_scale(y_mean, y_std)
                _, ws_d, ws_n, neg_h = e._predict_ws(MP)
                inv_h = _lib.darr(inv_ls)
                mean = np.empty((M, e.T))
                std = np.empty((M, e.T)) if return_std else None
                neg_h[0] = 0
                keep.append((Xq, inv_h, ym, ys, ws_d))
                out.append((mean, std, neg_h))
                m = arr[k]
                m.Xq_h, m.D, m.X_d, m.N = Xq.ctypes.data, e.D, e.X_d.data_ptr(), e.N
                m.inv_ls_h, m.c, m.noise = C.cast(inv_h, C.c_void_p), c, noise
                m.alpha_d, m.T, m.Mlow_d, m.NP = e.alpha_d.data_ptr(), e.T, e.Mlow_d.data_ptr(), e.NP
                m.y_mean_d, m.y_std_d = ym.data_ptr(), ys.data_ptr()
                m.mean_h = mean.ctypes.data
                m.std_h = std.ctypes.data if return_std else None
                m.neg_h = neg_h.ctypes.data if return_std else None
                m.ws_d, m.ws_doubles = ws_d.data_ptr(), ws_n
            _lib.check(engines[0].lib.gpk_predict_stacked(arr, K, M, st), "gpk_predict_stacked")
    finally:
        for e in locks:
            e._plock.release()
    return [(mean, std, int(neg_h[0]) if return_std else 0) for mean, std, neg_h in out]


DeviceGP.predict_stacked = staticmethod(predict_stacked)   # gpr.predict_stacked finds it on the engines' class
