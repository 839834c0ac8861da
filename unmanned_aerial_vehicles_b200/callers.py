"""Host-side mirrors of the reference callers either side of the hot path, each keeping the
reference's per-row API and adding the batched form the GPU wants.

  SimpleQuadrotorGP        <- src/px4/simple_gp.py:24-223      (ring buffers, filters, train_gp, predict_residual)
  build_gp_residuals       <- src/px4/mpc.py:1475-1511          (25 sequential one-row predicts -> ONE batched call)
  PerOutputGP              <- src/px4/pretrained_gp.py:27-98, gp_trainer.py:207-227 (schema B: per-output models
                              + scalers), gp_enhanced_mpc.py:94-183 (schema C)
  PreTrainedGP             <- src/px4/pretrained_gp.py:12-118   (schema-B loader with the per-row API of the controller)
  EnhancedMPCResiduals     <- src/px4/gp_enhanced_mpc.py:94-183 (schema-C loader: pos_*/vel_* members, mean only)
  GPTrainer                <- src/px4/gp_trainer.py:27-245     (npz ingest, per-output scaler + ARD fit loop, test
                              metrics, schema-B pickle); fits shard over ranks, test predictions are ONE device call
  adopt_pickle             <- the three pickle schemas of SURVEY section 5
  evaluate_residual_metrics<- src/px4/evaluate_gp_offline.py:186-312 (MSE/RMSE/R2/improvement/frac_better)
  load_csv_dataset         <- src/px4/train_gp_offline.py:22-76 (vectorised: no per-row Python loop)
"""
from __future__ import annotations

import os
import pickle
from collections import deque

import numpy as np
from sklearn.gaussian_process.kernels import RBF, WhiteKernel

from . import _lib
from .gpr import GaussianProcessRegressor, predict_stacked

FEATURE_COLS = ["x", "y", "z", "vx", "vy", "vz", "ax", "ay", "az", "yaw_rate"]
RESIDUAL_COLS = ["res_dx", "res_dy", "res_dz", "res_dvx", "res_dvy", "res_dvz"]


def load_csv_dataset(path, max_residual_norm=5.0):
    """(X (n,10), Y (n,6)) with the reference's row filter: finite and ||res|| < 5 (train_gp_offline.py:59-65)."""
    import pandas as pd

    df = pd.read_csv(path)
    missing = [c for c in FEATURE_COLS + RESIDUAL_COLS if c not in df.columns]
    if missing:
        raise ValueError(f"Missing columns: {missing}")
    X = df[FEATURE_COLS].to_numpy(dtype=np.float64)
    Y = df[RESIDUAL_COLS].to_numpy(dtype=np.float64)
    ok = np.isfinite(X).all(1) & np.isfinite(Y).all(1) & (np.linalg.norm(Y, axis=1) < max_residual_norm)
    return X[ok], Y[ok]


class SimpleQuadrotorGP:
    """Same surface as the reference wrapper; the estimator underneath is the B200 one."""

    def __init__(self, max_data_points=1000, random_state=None):
        self.max_data_points = max_data_points
        self.X_train = deque(maxlen=max_data_points)
        self.Y_train = deque(maxlen=max_data_points)
        self.gp_model = None
        self.is_trained = False
        self.training_count = 0
        self.prediction_count = 0
        self.random_state = random_state

    # -- data ------------------------------------------------------------------------------------
    @staticmethod
    def _nominal_dynamics(state, control, dt):
        state = np.asarray(state, dtype=np.float64)
        control = np.asarray(control, dtype=np.float64)
        return state + dt * np.concatenate([state[..., 3:6], control[..., :3]], axis=-1)

    def add_training_data(self, state, control, state_next, dt=0.02):
        if len(state) < 6 or len(state_next) < 6:
            return
        state, control, state_next = (np.asarray(v, dtype=np.float64) for v in (state, control, state_next))
        if np.linalg.norm(state[3:6]) > 5.0 or np.linalg.norm(control[:3]) > 3.0:
            return
        residual = state_next - self._nominal_dynamics(state[:6], control, dt)
        if np.linalg.norm(residual) > 2.0:
            return
        self.X_train.append(np.concatenate([state[:6], control[:4]]))
        self.Y_train.append(residual.copy())

    def extend(self, X, Y):
        for xi, yi in zip(X, Y):
            self.X_train.append(np.asarray(xi, dtype=np.float64))
            self.Y_train.append(np.asarray(yi, dtype=np.float64))

    def _to_numpy(self):
        if len(self.X_train) == 0:
            return np.empty((0, 10)), np.empty((0, 6))
        return np.array(self.X_train), np.array(self.Y_train)

    def save_dataset(self, csv_path, include_header=True, overwrite=True):
        """simple_gp.py:75-115: the 16-column CSV (np.savetxt, '%.18e'), appended without header when
        ``overwrite=False`` and the file exists; nothing is written for an empty store."""
        X, Y = self._to_numpy()
        if X.shape[0] == 0:
            return
        data = np.hstack([X, Y])
        os.makedirs(os.path.dirname(csv_path) or ".", exist_ok=True)
        if overwrite or not os.path.exists(csv_path):
            np.savetxt(csv_path, data, delimiter=",", header=",".join(FEATURE_COLS + RESIDUAL_COLS) if include_header else "",
                       comments="")
        else:
            with open(csv_path, "ab") as f:
                np.savetxt(f, data, delimiter=",")

    # -- train / predict -------------------------------------------------------------------------
    def train_gp(self):
        if len(self.X_train) < 30:
            return
        try:
            X = np.array(list(self.X_train))
            Y = np.array(list(self.Y_train))
            kernel = RBF(length_scale=0.5) + WhiteKernel(noise_level=0.1)          # simple_gp.py:167
            self.gp_model = GaussianProcessRegressor(kernel=kernel, alpha=1e-4, normalize_y=True,
                                                     n_restarts_optimizer=1, random_state=self.random_state)
            self.gp_model.fit(X, Y)
            self.is_trained = True
            self.training_count += 1
        except (np.linalg.LinAlgError, ValueError) as e:  # the reference swallows and reports
            print(f"GP training failed: {e}")
            self.is_trained = False

    def predict_residual(self, state, control):
        """one row -> (mean (6,), variance (6,)), as simple_gp.py:187-201: (zeros, ones) when untrained or when the
        prediction raises (a missing CUDA library is not swallowed: there is no silent fallback)"""
        if not self.is_trained:
            return np.zeros(6), np.ones(6)
        try:
            x = np.concatenate([state, control]).reshape(1, -1)
            mean, std = self.gp_model.predict(x, return_std=True)
        except _lib.GpkUnavailable:
            raise
        except Exception as e:
            print(f"GP prediction failed: {e}")
            return np.zeros(6), np.ones(6)
        self.prediction_count += 1
        return mean.flatten(), std.flatten() ** 2

    def predict_residual_batch(self, states, controls, return_var=True):
        """(M,6),(M,4) -> mean (M,T) [, variance (M,T)] in ONE device call."""
        states = np.asarray(states, dtype=np.float64)
        if not self.is_trained:
            return np.zeros((len(states), 6)), np.ones((len(states), 6))
        x = np.hstack([states, np.asarray(controls, dtype=np.float64)])
        self.prediction_count += len(x)
        if not return_var:
            return np.atleast_2d(self.gp_model.predict(x)), None
        mean, std = self.gp_model.predict(x, return_std=True)
        return mean.reshape(len(x), -1), std.reshape(len(x), -1) ** 2

    def get_uncertainty(self, state, control):
        _, variance = self.predict_residual(state, control)
        return float(np.mean(np.sqrt(variance)))

    def predict_enhanced_dynamics(self, state, control, dt):
        mean, _ = self.predict_residual(state, control)
        return self._nominal_dynamics(np.asarray(state)[:6], control, dt) + mean

    def load_model(self, model_path):
        """simple_gp.py:50-73: schema-A pickle -> True; a missing file, a pickle without 'gp_model' or one that does
        not load -> False (reported, not raised)"""
        try:
            if not os.path.exists(model_path):
                return False
            with open(model_path, "rb") as f:
                data = pickle.load(f)
            model = adopt_pickle({"gp_model": data["gp_model"]})
            self.gp_model = model
            self.is_trained = True
            self.training_count = data.get("training_count", 0)
            return True
        except Exception as e:
            print(f"Failed to load model: {e}")
            return False

    def get_stats(self):
        return {"is_trained": self.is_trained, "data_points": len(self.X_train),
                "training_iterations": self.training_count, "predictions_made": self.prediction_count,
                "sklearn_available": True}


def build_gp_residuals(gp, X_guess, U_guess, dt, gain, n_states=6):
    """Batched QuadrotorMPC._build_gp_residuals (mpc.py:1475-1511): D (n_states, N) with rows 3:6 =
    gain * residual_mean[3:6] / dt for the N warm-start pairs (x_k, u_k), from ONE predict call."""
    N = U_guess.shape[1]
    D = np.zeros((n_states, N))
    if gp is None or not gp.is_trained:
        return D
    mean, _ = gp.predict_residual_batch(X_guess[:, :N].T, U_guess.T, return_var=False)
    if mean.shape[1] >= n_states:
        D[3:6, :] = gain * (mean[:, 3:6] / dt).T
    return D


def build_gp_residuals_scenarios(gp, X_guess, U_guess, dt, gain, n_states=6):
    """Scenario form: X_guess (S, 6, N+1), U_guess (S, 4, N) -> D (S, n_states, N), one device call."""
    S, _, N = U_guess.shape
    xs = np.transpose(X_guess[:, :, :N], (0, 2, 1)).reshape(S * N, -1)
    us = np.transpose(U_guess, (0, 2, 1)).reshape(S * N, -1)
    D = np.zeros((S, n_states, N))
    if gp is None or not gp.is_trained:
        return D
    mean, _ = gp.predict_residual_batch(xs, us, return_var=False)
    D[:, 3:6, :] = gain * np.transpose((mean[:, 3:6] / dt).reshape(S, N, 3), (0, 2, 1))
    return D


def _to_b200(model):
    if isinstance(model, GaussianProcessRegressor):
        return model
    if hasattr(model, "kernel_") and hasattr(model, "X_train_"):
        return GaussianProcessRegressor.from_sklearn(model)
    return model


class PerOutputGP:
    """Schema B/C: one single-output GP + input/output StandardScaler per residual component."""

    def __init__(self, models, scalers_X, scalers_y, names=None):
        self.names = list(names) if names is not None else list(models.keys())
        self.models = {k: _to_b200(v) for k, v in models.items()}
        self.scalers_X, self.scalers_y = scalers_X, scalers_y
        self.is_loaded = True

    @staticmethod
    def _plain_standard_scaler(sc):
        from sklearn.preprocessing import StandardScaler
        return type(sc) is StandardScaler and sc.with_mean and sc.with_std and getattr(sc, "mean_", None) is not None \
            and getattr(sc, "scale_", None) is not None

    @classmethod
    def _scale_in(cls, sc, X):
        """sc.transform(X); a plain fitted StandardScaler on a float64 matrix is done as its own two array operations
        ((X - mean_) / scale_, bit-identical) without sklearn's ~180 us of validation per call -- six scalers per
        control tick in pretrained_gp.py:75 / gp_enhanced_mpc.py:160."""
        if cls._plain_standard_scaler(sc) and X.shape[1] == sc.mean_.shape[0]:
            return (X - sc.mean_) / sc.scale_
        return sc.transform(X)

    @classmethod
    def _scale_out(cls, sc, m):
        """sc.inverse_transform(m.reshape(-1, 1)).ravel() (pretrained_gp.py:83-85), same shortcut"""
        m = np.asarray(m, dtype=np.float64).reshape(-1)
        if cls._plain_standard_scaler(sc) and sc.mean_.shape[0] == 1:
            return m * sc.scale_[0] + sc.mean_[0]
        return sc.inverse_transform(m.reshape(-1, 1)).ravel()

    def predict_batch(self, X, return_std=True):
        """(M, D) -> mean (M, n_outputs), std (M, n_outputs); missing outputs give (0, 1e6) like the reference.
        All members answer in ONE device call (gpr.predict_stacked), each through its own input scaler."""
        X = np.ascontiguousarray(X, dtype=np.float64)
        if X.ndim != 2 or not np.isfinite(X).all():
            from sklearn.utils.validation import check_array
            X = check_array(X, ensure_2d=True, dtype=np.float64)          # sklearn's own errors for bad input
        M = len(X)
        mean = np.zeros((M, len(self.names)))
        std = np.full((M, len(self.names)), 1e6)
        present = [(i, name) for i, name in enumerate(self.names) if name in self.models]
        if not present or M == 0:
            return mean, (std if return_std else None)
        Xs = [self._scale_in(self.scalers_X[name], X) for _, name in present]
        outs = predict_stacked([self.models[name] for _, name in present], Xs, return_std=return_std)
        for (i, name), out in zip(present, outs):
            sy = self.scalers_y[name]
            if return_std:
                m, s = out
                std[:, i] = np.abs(np.asarray(s).reshape(-1) * sy.scale_[0])                      # pretrained_gp.py:87
            else:
                m = out
            mean[:, i] = self._scale_out(sy, m)
        return mean, (std if return_std else None)

    def predict_residual(self, state, control):
        x = np.concatenate([np.asarray(state)[:6], np.asarray(control)[:4]]).reshape(1, -1)
        m, s = self.predict_batch(x)
        return m[0], s[0]


class PreTrainedGP:
    """pretrained_gp.py:12-118 with the same surface: loads a schema-B pickle at construction, one-row
    ``predict_residual`` -> (mean (6,), std (6,)), missing components (0, 1e6), everything (0, 1e6) when nothing is
    loaded.  The six members answer in one device call; ``predict_residual_batch`` is the horizon form."""

    def __init__(self, model_path):
        self.model_path = model_path
        self.gp_models, self.scalers_X, self.scalers_y, self.training_stats = {}, {}, {}, {}
        self.is_loaded = False
        self._members = None
        self.load_models()

    def load_models(self):
        if not os.path.exists(self.model_path):
            return False
        try:
            with open(self.model_path, "rb") as f:
                data = pickle.load(f)
            members = PerOutputGP(data["gp_models"], data["scalers_X"], data["scalers_y"], names=OUTPUT_NAMES)
            self.training_stats = data["training_stats"]
        except Exception as e:  # the reference reports and carries on without a model (pretrained_gp.py:48-50)
            print(f"Failed to load GP models: {e}")
            return False
        self._members = members
        self.gp_models, self.scalers_X, self.scalers_y = members.models, members.scalers_X, members.scalers_y
        self.is_loaded = True
        return True

    def predict_residual_batch(self, states, controls):
        """(M, >=6), (M, >=4) -> mean (M, 6), std (M, 6)"""
        states, controls = np.atleast_2d(np.asarray(states, dtype=np.float64)), np.atleast_2d(np.asarray(controls, dtype=np.float64))
        if not self.is_loaded:
            return np.zeros((len(states), 6)), np.full((len(states), 6), 1e6)
        try:
            return self._members.predict_batch(np.hstack([states[:, :6], controls[:, :4]]))
        except Exception as e:  # pretrained_gp.py:91-94: a failed prediction answers (0, 1e6)
            print(f"GP prediction failed: {e}")
            return np.zeros((len(states), 6)), np.full((len(states), 6), 1e6)

    def predict_residual(self, state, control):
        m, s = self.predict_residual_batch(np.asarray(state)[None, :], np.asarray(control)[None, :])
        return m[0], s[0]

    def get_uncertainty(self, state, control):
        return float(np.mean(self.predict_residual(state, control)[1]))

    def get_stats(self):
        return {"is_loaded": self.is_loaded, "model_path": self.model_path,
                "available_models": list(self.gp_models.keys()), "training_stats": self.training_stats}


class EnhancedMPCResiduals:
    """The GP part of gp_enhanced_mpc.py (:94-183) without the ROS node around it: schema-C pickle
    (``models`` / ``scalers_input`` / ``scalers_output`` keyed pos_x..vel_vz, all six required), 12-d MPC state of
    which [0:3] and [6:9] are position and velocity, mean only, ``(zeros(3), zeros(3))`` when unavailable.  Like the
    reference, every member sees the input through the scaler of ``pos_x`` (:160)."""

    REQUIRED = ["pos_x", "pos_y", "pos_z", "vel_vx", "vel_vy", "vel_vz"]

    def __init__(self, model_file=None):
        self.gp_available = False
        self.gp_predictions_enabled = False
        self.training_stats = {}
        self._members = None
        if model_file:
            self.setup_gp(model_file)

    def setup_gp(self, model_file):
        if not model_file or not os.path.exists(model_file):
            return False
        try:
            with open(model_file, "rb") as f:
                pkg = pickle.load(f)
            if not all(k in pkg["models"] for k in self.REQUIRED):
                raise ValueError("Incomplete GP model set")
            shared = {k: pkg["scalers_input"]["pos_x"] for k in self.REQUIRED}
            self._members = PerOutputGP({k: pkg["models"][k] for k in self.REQUIRED}, shared, pkg["scalers_output"],
                                        names=self.REQUIRED)
            self.training_stats = pkg.get("training_stats", {})
        except Exception as e:
            print(f"Failed to load GP models: {e}")
            return False
        self.gp_available = self.gp_predictions_enabled = True
        return True

    def predict_gp_residuals_batch(self, states, controls):
        """(M, >=9), (M, 4) -> position residuals (M, 3), velocity residuals (M, 3): one device call"""
        states, controls = np.atleast_2d(np.asarray(states, dtype=np.float64)), np.atleast_2d(np.asarray(controls, dtype=np.float64))
        if not self.gp_available or not self.gp_predictions_enabled:
            return np.zeros((len(states), 3)), np.zeros((len(states), 3))
        try:
            x = np.hstack([states[:, :3], states[:, 6:9], controls])
            mean, _ = self._members.predict_batch(x, return_std=False)
            return mean[:, :3], mean[:, 3:]
        except Exception as e:
            print(f"GP prediction failed: {e}")
            return np.zeros((len(states), 3)), np.zeros((len(states), 3))

    def predict_gp_residuals(self, state, control):
        p, v = self.predict_gp_residuals_batch(np.asarray(state)[None, :], np.asarray(control)[None, :])
        return p[0], v[0]


OUTPUT_NAMES = ["x_residual", "y_residual", "z_residual", "vx_residual", "vy_residual", "vz_residual"]


class GPTrainer:
    """The reference's offline trainer (gp_trainer.py:27-245) with the same surface and the same schema-B pickle:
    one StandardScaler pair + one ARD ``C(1, fixed) * RBF([1]*D, (0.1, 10)) + White(0.01, (1e-5, 10))`` model with
    3 restarts per residual component, fitted on an 80/20 split (seed 42), scored on the held-out rows.

    What differs underneath: the npz ingest is array arithmetic instead of a per-sample Python loop
    (gp_trainer.py:71-91); the component fits are independent, so under torch.distributed they are sharded over the
    ranks (parallel.run_sharded; every rank ends with every model); the held-out predictions of all components are
    one device call.  ``random_state`` seeds the restart draws (the reference leaves them on NumPy's global stream)."""

    def __init__(self, data_dir="data/training_data", model_dir="data/trained_models", random_state=None, group=None):
        self.data_dir, self.model_dir = data_dir, model_dir
        os.makedirs(model_dir, exist_ok=True)
        self.gp_models, self.scalers_X, self.scalers_y, self.training_stats = {}, {}, {}, {}
        self.random_state, self.group = random_state, group

    @staticmethod
    def _nominal_dynamics(state, control, dt):
        """double integrator (gp_trainer.py:105-119); accepts one sample or stacked samples with dt of shape (n,)"""
        state, control = np.asarray(state, dtype=np.float64), np.asarray(control, dtype=np.float64)
        dt = np.asarray(dt, dtype=np.float64)[..., None] if np.ndim(dt) else dt
        return np.concatenate([state[..., :3] + state[..., 3:6] * dt, state[..., 3:6] + control[..., :3] * dt], axis=-1)

    def load_training_data(self, max_samples=None):
        """every flight_data_*.npz under data_dir -> X (n, 10) = [state_prev, control], y (n, 6) = residual"""
        import glob

        files = glob.glob(os.path.join(self.data_dir, "flight_data_*.npz"))
        if not files:
            raise FileNotFoundError(f"No training data found in {self.data_dir}")
        Xs, ys = [], []
        for f in files:
            d = np.load(f)
            prev, ctrl = np.asarray(d["states_prev"], dtype=np.float64), np.asarray(d["controls"], dtype=np.float64)
            if len(prev) == 0:
                continue
            ys.append(np.asarray(d["states_next"], dtype=np.float64) - self._nominal_dynamics(prev, ctrl, d["dt_values"]))
            Xs.append(np.hstack([prev, ctrl]))
        X = np.vstack(Xs) if Xs else np.empty((0, 10))
        y = np.vstack(ys) if ys else np.empty((0, 6))
        if max_samples and len(X) > max_samples:
            idx = np.random.choice(len(X), max_samples, replace=False)            # gp_trainer.py:97-100
            X, y = X[idx], y[idx]
        return X, y

    def _make_estimator(self, n_features, t):
        from sklearn.gaussian_process.kernels import ConstantKernel

        kernel = ConstantKernel(1.0, constant_value_bounds="fixed") * RBF(
            length_scale=[1.0] * n_features, length_scale_bounds=(0.1, 10.0)) + WhiteKernel(
            noise_level=0.01, noise_level_bounds=(1e-5, 1e1))                      # gp_trainer.py:163-166
        return GaussianProcessRegressor(kernel=kernel, n_restarts_optimizer=3, alpha=1e-6, normalize_y=False,
                                        random_state=self.random_state)

    def train_gp_models(self, X, y, test_size=0.2):
        from sklearn.model_selection import train_test_split
        from sklearn.preprocessing import StandardScaler

        from .parallel import run_sharded

        X, y = np.asarray(X, dtype=np.float64), np.asarray(y, dtype=np.float64)
        X_train, X_test, y_train, y_test = train_test_split(X, y, test_size=test_size, random_state=42)
        todo = [(i, name) for i, name in enumerate(OUTPUT_NAMES[: y.shape[1]]) if np.std(y_train[:, i]) >= 1e-6]

        kept = {}

        def scaled(i):
            sX, sY = StandardScaler(), StandardScaler()
            return sX, sY, sX.fit_transform(X_train), sY.fit_transform(y_train[:, i].reshape(-1, 1)).flatten()

        def fit_one(item):
            i, name = item
            sX, sY, Xs, ys = scaled(i)
            kept[name] = (sX, sY, self._make_estimator(X.shape[1], i).fit(Xs, ys))
            gp = kept[name][2]
            return np.array(gp.kernel_.theta), float(gp.log_marginal_likelihood_value_)

        # only (theta*, lml*) crosses ranks; the other ranks rebuild each model from them (scalers are a function of
        # the split every rank holds) and factor lazily at the first predict -- no N x N matrix is pickled
        fitted = run_sharded(fit_one, todo, self.group)
        for (i, name), (theta, lml) in zip(todo, fitted):
            if name in kept:
                sX, sY, gp = kept[name]
            else:
                sX, sY, Xs, ys = scaled(i)
                gp = self._make_estimator(X.shape[1], i).fit_from_theta(Xs, ys, theta, lml)
            self.gp_models[name], self.scalers_X[name], self.scalers_y[name] = gp, sX, sY
        results = {}
        if todo:
            members = PerOutputGP({n: self.gp_models[n] for _, n in todo}, self.scalers_X, self.scalers_y,
                                  names=[n for _, n in todo])
            pred, _ = members.predict_batch(X_test, return_std=False)             # all components, one device call
            for k, (i, name) in enumerate(todo):
                gp, t = self.gp_models[name], y_test[:, i]
                mse = float(np.mean((t - pred[:, k]) ** 2))
                ss_tot = float(np.sum((t - t.mean()) ** 2))
                r2 = 1.0 - float(np.sum((t - pred[:, k]) ** 2)) / ss_tot if ss_tot > 0 else float("nan")
                results[name] = {"mse": mse, "rmse": float(np.sqrt(mse)), "r2": r2, "kernel": str(gp.kernel_),
                                 "log_marginal_likelihood": gp.log_marginal_likelihood()}
        self.training_stats = results
        return results

    def save_models(self, model_name=None):
        import time
        from datetime import datetime

        if model_name is None:
            model_name = f"gp_model_{datetime.now().strftime('%Y%m%d_%H%M%S')}"
        path = os.path.join(self.model_dir, f"{model_name}.pkl")
        with open(path, "wb") as f:                                               # schema B (gp_trainer.py:214-221)
            pickle.dump({"gp_models": self.gp_models, "scalers_X": self.scalers_X, "scalers_y": self.scalers_y,
                         "training_stats": self.training_stats, "model_name": model_name,
                         "creation_time": time.time()}, f)
        return path

    def load_models(self, model_path):
        with open(model_path, "rb") as f:
            data = pickle.load(f)
        self.gp_models = {k: _to_b200(v) for k, v in data["gp_models"].items()}
        self.scalers_X, self.scalers_y = data["scalers_X"], data["scalers_y"]
        self.training_stats = data["training_stats"]
        return self.training_stats


def adopt_pickle(obj):
    """Return a B200 estimator (schema A / bare estimator) or a PerOutputGP (schemas B, C)."""
    if isinstance(obj, dict):
        if "gp_model" in obj:                                                       # A: train_gp_offline.py:188
            return _to_b200(obj["gp_model"])
        if "gp_models" in obj:                                                      # B: gp_trainer.py:214
            return PerOutputGP(obj["gp_models"], obj["scalers_X"], obj["scalers_y"])
        if "models" in obj:                                                         # C: gp_enhanced_mpc.py:119
            return PerOutputGP(obj["models"], obj["scalers_input"], obj["scalers_output"])
        return obj
    return _to_b200(obj)


def evaluate_residual_metrics(Y_true, Y_gp, names=None):
    """evaluate_gp_offline.py:277-312: per component, nominal model = zero residual."""
    Y_true, Y_gp = np.asarray(Y_true, dtype=np.float64), np.asarray(Y_gp, dtype=np.float64)
    names = names or ["dx", "dy", "dz", "dvx", "dvy", "dvz"][: Y_true.shape[1]]
    rows = []
    for j, name in enumerate(names):
        t, g = Y_true[:, j], Y_gp[:, j]
        mse_nom, mse_gp = float(np.mean(t ** 2)), float(np.mean((t - g) ** 2))
        var = float(np.var(t))
        rows.append({"component": name, "mse_nom": mse_nom, "mse_gp": mse_gp, "rmse_nom": mse_nom ** 0.5,
                     "rmse_gp": mse_gp ** 0.5,
                     "r2_nom": 1.0 - mse_nom / var if var > 0 else float("nan"),
                     "r2_gp": 1.0 - mse_gp / var if var > 0 else float("nan"),
                     "improvement_%": (mse_nom - mse_gp) / max(mse_nom, 1e-12) * 100.0,
                     "frac_better": float(np.mean(t ** 2 > (t - g) ** 2))})
    return rows
