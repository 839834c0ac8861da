"""``GaussianProcessRegressor.fit`` for a training set that does not fit one GPU (BASELINE config 5: "an N=200k fit
completing on 8xB200"; the reference caps its training set at 10 000 rows, train_gp_offline.py:124, because its exact
GP is O(N^3) on the CPU).

Every rank of the process group runs the SAME host optimiser (SciPy L-BFGS-B, sklearn _gpr.py:658-674) on the SAME
objective values: `DistGP.lml_grad` returns the log marginal likelihood and its gradient after all-reduces, so they are
bit-identical on every rank and the iterates never diverge -- no broadcast of theta is needed.  One evaluation =
Gram assembly + 2-D block-cyclic Cholesky + alpha + distributed TRTRI/LAUUM + tile-wise gradient (dist_chol.py,
dist_grad.py).  After the last iteration the factor is recomputed at theta* and kept for `predict`.

Same restrictions as the single-GPU estimator (kernel algebra `[C*]RBF[+White]`, scalar alpha); `normalize_y` and
optimiser restarts are supported the way sklearn does them (restarts run one after the other: at this size a single
evaluation already uses every GPU).
"""
from __future__ import annotations

import warnings
from operator import itemgetter

import numpy as np
import scipy.optimize
from sklearn.base import clone
from sklearn.utils import check_random_state

from .dist_chol import DistGP
from .kernel_lowering import assemble_gradient, lower


class _EvaluationBudgetSpent(Exception):
    pass


class DistributedGPR:
    """sklearn-surface subset (fit / predict / log_marginal_likelihood / kernel_ / log_marginal_likelihood_value_) over
    the distributed factor.  ``maxiter`` truncates each L-BFGS-B run (None: SciPy's default, as sklearn).  ``maxfun`` is
    a HARD cap on the objective evaluations of one run: SciPy's own ``maxfun`` is only checked between iterations (its
    line search went to four evaluations under ``maxfun=2`` in the N=200k run of profiles/r02i_*), so the objective
    itself stops the run after the cap and the best point evaluated so far is taken as that run's result."""

    def __init__(self, kernel, *, alpha=1e-10, optimizer="fmin_l_bfgs_b", n_restarts_optimizer=0, normalize_y=False,
                 random_state=None, maxiter=None, maxfun=None, on_eval=None, NB=2048, grid=None, ops=None, lookahead=True):
        self.kernel, self.alpha, self.optimizer = kernel, alpha, optimizer
        self.n_restarts_optimizer, self.normalize_y, self.random_state = n_restarts_optimizer, normalize_y, random_state
        self.maxiter, self.NB, self.grid, self.ops, self.lookahead = maxiter, NB, grid, ops, lookahead
        self.maxfun, self.on_eval = maxfun, on_eval      # cap on objective evaluations per run; callback(k, lml, s)

    def fit(self, X, y):
        X = np.ascontiguousarray(X, dtype=np.float64)
        y = np.asarray(y, dtype=np.float64)
        self._y_1d = y.ndim == 1
        y2d = y[:, None] if self._y_1d else y
        if self.normalize_y:                                    # _gpr.py:275-285
            self._y_train_mean = y2d.mean(0)
            std = y2d.std(0)
            self._y_train_std = np.where(std < 10 * np.finfo(std.dtype).eps, 1.0, std)
        else:
            self._y_train_mean, self._y_train_std = np.zeros(y2d.shape[1]), np.ones(y2d.shape[1])
        y2d = (y2d - self._y_train_mean) / self._y_train_std
        self.kernel_ = clone(self.kernel)
        self.n_features_in_ = X.shape[1]
        template = lower(self.kernel_, X.shape[1])
        self.X_train_, self.y_train_ = X, (y2d[:, 0] if self._y_1d else y2d)
        self._gp = DistGP(X, y2d, float(self.alpha), NB=self.NB, grid=self.grid, ops=self.ops, lookahead=self.lookahead)
        self.n_lml_evals_ = 0
        self.evaluation_seconds_ = []

        run = {"evals": 0, "best": None}          # per L-BFGS-B run: evaluations so far, (f, theta) of the best one

        def obj_func(theta):
            import time
            if self.maxfun is not None and run["evals"] >= int(self.maxfun):
                raise _EvaluationBudgetSpent()
            t0 = time.perf_counter()
            low = template.with_theta(theta)
            lml, g_c, g_ls, g_noise = self._gp.lml_grad(1.0 / low.length_scale, low.c, low.noise)
            self.n_lml_evals_ += 1
            self.evaluation_seconds_.append(time.perf_counter() - t0)
            if self.on_eval is not None:
                self.on_eval(self.n_lml_evals_, float(lml), self.evaluation_seconds_[-1])
            run["evals"] += 1
            if np.isfinite(lml) and (run["best"] is None or -lml < run["best"][0]):
                run["best"] = (-float(lml), np.array(theta, dtype=np.float64))
            if not np.isfinite(lml):
                return np.inf, np.zeros(len(low.theta_map))    # _gpr.py:592-593
            return -lml, -assemble_gradient(low, g_c, g_ls, g_noise)

        if self.optimizer is not None and self.kernel_.n_dims > 0:
            rng = check_random_state(self.random_state)
            bounds = self.kernel_.bounds
            starts = [self.kernel_.theta] + [rng.uniform(bounds[:, 0], bounds[:, 1])
                                             for _ in range(self.n_restarts_optimizer)]   # _gpr.py:312-333
            opts = {} if self.maxiter is None else {"maxiter": int(self.maxiter)}
            optima = []
            for t0 in starts:
                run["evals"], run["best"] = 0, None
                try:
                    with warnings.catch_warnings():
                        warnings.simplefilter("ignore")
                        r = scipy.optimize.minimize(obj_func, t0, method="L-BFGS-B", jac=True, bounds=bounds, options=opts)
                    optima.append((r.x, r.fun))
                except _EvaluationBudgetSpent:            # raised identically on every rank: same values everywhere
                    if run["best"] is None:
                        raise np.linalg.LinAlgError("no finite objective value within the evaluation budget")
                    optima.append((run["best"][1], run["best"][0]))
            best = int(np.argmin(list(map(itemgetter(1), optima))))
            self.kernel_.theta = optima[best][0]
            self.log_marginal_likelihood_value_ = -float(optima[best][1])
        self._gp.release_inverse()
        low = lower(self.kernel_, X.shape[1])
        info = self._gp.factor(1.0 / low.length_scale, low.c, low.noise)
        if info != 0:
            raise np.linalg.LinAlgError(
                f"The kernel, {self.kernel_}, is not returning a positive definite matrix. Try gradually "
                "increasing the 'alpha' parameter of your GaussianProcessRegressor estimator.",
                f"{info}-th leading minor of the array is not positive definite")
        self._gp.solve()
        if self.optimizer is None or self.kernel_.n_dims == 0:
            self.log_marginal_likelihood_value_ = self._gp.lml()
        return self

    def log_marginal_likelihood(self, theta=None, eval_gradient=False):
        if theta is None:
            return self.log_marginal_likelihood_value_
        low = lower(self.kernel_, self.n_features_in_).with_theta(theta)
        self._ensure_factored()
        gp = self._gp
        if eval_gradient:
            lml, g_c, g_ls, g_noise = gp.lml_grad(1.0 / low.length_scale, low.c, low.noise)
            out = (lml, assemble_gradient(low, g_c, g_ls, g_noise)) if np.isfinite(lml) else \
                (-np.inf, np.zeros(len(low.theta_map)))
        else:
            info = gp.factor(1.0 / low.length_scale, low.c, low.noise)
            if info == 0:
                gp.solve()
            out = gp.lml() if info == 0 else -np.inf
        cur = lower(self.kernel_, self.n_features_in_)           # leave the fitted factor in place
        if gp.factor(1.0 / cur.length_scale, cur.c, cur.noise) == 0:
            gp.solve()
        return out

    # -- pickling: host NumPy only (train_gp_offline.py:188-197 stores the model under ['gp_model']).  The factor is
    # NOT stored -- it is 160 GB at N=200k -- and is rebuilt from kernel_ on the first predict after loading, on
    # whatever process group the loading job runs in (set .ops / .grid before that to choose).
    def __getstate__(self):
        state = dict(self.__dict__)
        state.pop("_gp", None)
        state["ops"] = None              # communicators and device handles do not travel
        state["on_eval"] = None
        return state

    def __setstate__(self, state):
        self.__dict__.update(state)
        self._gp = None

    def _ensure_factored(self):
        if getattr(self, "_gp", None) is not None:
            return
        y2d = self.y_train_[:, None] if self._y_1d else self.y_train_
        self._gp = DistGP(self.X_train_, y2d, float(self.alpha), NB=self.NB, grid=self.grid, ops=self.ops,
                          lookahead=self.lookahead)
        low = lower(self.kernel_, self.n_features_in_)
        info = self._gp.factor(1.0 / low.length_scale, low.c, low.noise)
        if info != 0:
            raise np.linalg.LinAlgError(f"The kernel, {self.kernel_}, is not returning a positive definite matrix.")
        self._gp.solve()

    def predict(self, X, return_std=False):
        """mean [, std] with sklearn's output shapes (_gpr.py:446-500); the same rows on every rank"""
        self._ensure_factored()
        res = self._gp.predict(X, return_std=return_std)
        mean, std = (res if return_std else (res, None))
        mean = mean * self._y_train_std + self._y_train_mean
        if self._y_1d:
            mean = mean[:, 0]
        if not return_std:
            return mean
        std = std * self._y_train_std
        return mean, (std[:, 0] if self._y_1d or std.shape[1] == 1 else std)
