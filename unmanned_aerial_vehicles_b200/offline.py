"""Host-side mirrors of the reference's OFFLINE tools either side of the hot path:
the data formats that feed the fit and the evaluation that consumes predict.  Same function names, argument meaning,
return structures and on-disk formats as the reference; array arithmetic instead of per-sample Python loops, and the
GP underneath is the B200 estimator.

  GPDataCollector      <- src/px4/gp_data_collector.py:15-99   (flight_data_<session>.npz: the schema GPTrainer ingests)
  SimpleGPEnhancedMPC  <- src/px4/simple_gp.py:225-260         (uncertainty-gated use of the GP correction)
  f_nominal, load_gp_model, load_dataset, r2_score, evaluate_gp
                       <- src/px4/evaluate_gp_offline.py:49-365 (metrics dict + <prefix>_metrics.csv / .tex)
  GPModelEvaluator     <- src/px4/gp_evaluation.py:40-330      (pickle-format detection, seeded test-data generators,
                                                                  predict_on_test_data; the plotting half is out of scope)
  train_offline        <- src/px4/train_gp_offline.py:22-215    (CSV files -> SimpleQuadrotorGP.train_gp -> schema-A pickle
                                                                  + gp_model_latest.pkl link)
"""
from __future__ import annotations

import glob
import os
import pickle
import time
from datetime import datetime
from pathlib import Path

import numpy as np

from .callers import FEATURE_COLS, RESIDUAL_COLS, PerOutputGP, SimpleQuadrotorGP, _to_b200, adopt_pickle, load_csv_dataset

COMPONENT_NAMES = ["dx", "dy", "dz", "dvx", "dvy", "dvz"]


class GPDataCollector:
    """gp_data_collector.py:15-99: bounded sample store written as ``flight_data_<session>.npz`` with the keys
    ``states_prev, controls, states_next, timestamps, dt_values, flight_session, collection_time``."""

    def __init__(self, max_samples=10000, data_dir="data/training_data"):
        self.max_samples, self.data_dir = max_samples, data_dir
        os.makedirs(data_dir, exist_ok=True)
        self.states_prev, self.controls, self.states_next, self.timestamps, self.dt_values = [], [], [], [], []
        self.flight_session = datetime.now().strftime("%Y%m%d_%H%M%S")
        self.collection_active = True

    def add_sample(self, state_prev, control, state_next, dt):
        if not self.collection_active or len(self.states_prev) >= self.max_samples:
            return False
        self.states_prev.append(np.array(state_prev[:6], dtype=np.float64))
        self.controls.append(np.array(control[:4], dtype=np.float64))
        self.states_next.append(np.array(state_next[:6], dtype=np.float64))
        self.timestamps.append(time.time())
        self.dt_values.append(dt)
        return True

    def add_samples(self, states_prev, controls, states_next, dts):
        """batched form: rows beyond max_samples are dropped like the per-sample calls would drop them"""
        room = max(0, self.max_samples - len(self.states_prev)) if self.collection_active else 0
        n = min(room, len(states_prev))
        now = time.time()
        for a, b, c, d in zip(np.asarray(states_prev)[:n], np.asarray(controls)[:n], np.asarray(states_next)[:n], dts[:n]):
            self.states_prev.append(np.array(a[:6], dtype=np.float64))
            self.controls.append(np.array(b[:4], dtype=np.float64))
            self.states_next.append(np.array(c[:6], dtype=np.float64))
            self.timestamps.append(now)
            self.dt_values.append(float(d))
        return n

    def save_data(self, filename=None):
        if filename is None:
            filename = f"flight_data_{self.flight_session}.npz"
        path = os.path.join(self.data_dir, filename)
        np.savez_compressed(path, states_prev=np.array(self.states_prev), controls=np.array(self.controls),
                            states_next=np.array(self.states_next), timestamps=np.array(self.timestamps),
                            dt_values=np.array(self.dt_values), flight_session=self.flight_session,
                            collection_time=datetime.now().isoformat())
        return path

    def get_stats(self):
        return {"num_samples": len(self.states_prev), "max_samples": self.max_samples,
                "collection_active": self.collection_active, "flight_session": self.flight_session,
                "fill_percentage": (len(self.states_prev) / self.max_samples) * 100}

    def stop_collection(self):
        self.collection_active = False

    def clear_data(self):
        for buf in (self.states_prev, self.controls, self.states_next, self.timestamps, self.dt_values):
            buf.clear()


class SimpleGPEnhancedMPC:
    """simple_gp.py:225-260: use the GP-corrected step only where the mean predictive std is below a threshold."""

    def __init__(self, gp_model, confidence_threshold=0.5):
        self.gp_model, self.confidence_threshold = gp_model, confidence_threshold
        self.gp_usage = self.nominal_usage = 0

    def enhanced_dynamics_function(self, state, control, dt):
        gp = self.gp_model
        if not gp.is_trained:
            self.nominal_usage += 1
            return gp._nominal_dynamics(np.asarray(state)[:6], control, dt)
        # one predict answers both the gate and the correction (the reference calls predict twice: :240, :250)
        mean, var = gp.predict_residual(state, control)
        if float(np.mean(np.sqrt(var))) < self.confidence_threshold:
            self.gp_usage += 1
            return gp._nominal_dynamics(np.asarray(state)[:6], control, dt) + mean
        self.nominal_usage += 1
        return gp._nominal_dynamics(np.asarray(state)[:6], control, dt)

    def get_usage_ratio(self):
        total = self.gp_usage + self.nominal_usage
        return 0.0 if total == 0 else self.gp_usage / total * 100


# ---------------------------------------------------------------------------------------- evaluate_gp_offline.py
def f_nominal(x, u):
    """double-integrator derivative [v, a_cmd] (evaluate_gp_offline.py:49-65); one sample (6,),(4,) or stacked rows"""
    x, u = np.asarray(x, dtype=np.float64), np.asarray(u, dtype=np.float64)
    return np.concatenate([x[..., 3:6], u[..., 0:3]], axis=-1)


def load_gp_model(model_path):
    """evaluate_gp_offline.py:71-99: ``{'gp_model': m, ...}`` or a bare estimator; fitted sklearn estimators are adopted
    onto the B200 path.  Anything without ``.predict`` is an error."""
    with open(model_path, "rb") as f:
        obj = pickle.load(f)
    gp = obj["gp_model"] if isinstance(obj, dict) and "gp_model" in obj else obj
    if not hasattr(gp, "predict"):
        raise RuntimeError(f"Loaded object of type {type(gp)} has no .predict() method. "
                           "Check the contents of your pickle file.")
    return adopt_pickle(gp)


def load_dataset(csv_path):
    """evaluate_gp_offline.py:105-141 -> (df, X_feat (N,10), R_true (N,6), X_state (N,6), U_ctrl (N,4)); rows with NaN
    are dropped, a missing column raises ValueError naming it."""
    import pandas as pd

    df = pd.read_csv(csv_path).dropna().reset_index(drop=True)
    for c in FEATURE_COLS + RESIDUAL_COLS:
        if c not in df.columns:
            raise ValueError(f"Expected column '{c}' not found in CSV!")
    X_feat = df[FEATURE_COLS].to_numpy(dtype=float)
    R_true = df[RESIDUAL_COLS].to_numpy(dtype=float)
    return df, X_feat, R_true, X_feat[:, :6].copy(), X_feat[:, 6:10].copy()


def r2_score(y_true, y_pred):
    """evaluate_gp_offline.py:147-157: NaN when the target has no variance (ss_tot <= 1e-12)"""
    y_true, y_pred = np.asarray(y_true, dtype=float), np.asarray(y_pred, dtype=float)
    ss_tot = np.sum((y_true - np.mean(y_true)) ** 2)
    if ss_tot <= 1e-12:
        return np.nan
    return 1.0 - np.sum((y_true - y_pred) ** 2) / ss_tot


def _metrics_from_sums(sums, save_prefix=None):
    """the reference's result dict and metrics table (evaluate_gp_offline.py:277-365) from the device-side sums of
    GaussianProcessRegressor.predict_with_residual_metrics"""
    import pandas as pd

    n = float(sums["n"])

    def block(a, b):
        m_nom, m_gp = a / n, b / n
        return {"mse_nom": m_nom, "mse_gp": m_gp, "rmse_nom": np.sqrt(m_nom), "rmse_gp": np.sqrt(m_gp),
                "improvement_%": (m_nom - m_gp) / max(m_nom, 1e-12) * 100.0}

    rows = []
    for name, (s_t, s_t2, s_e2, n_better, s_dev2) in zip(COMPONENT_NAMES, sums["component"]):
        mse_nom, mse_gp = s_t2 / n, s_e2 / n
        r2 = (lambda ss_res: np.nan if s_dev2 <= 1e-12 else 1.0 - ss_res / s_dev2)
        rows.append({"component": name, "mse_nom": mse_nom, "mse_gp": mse_gp, "rmse_nom": np.sqrt(mse_nom),
                     "rmse_gp": np.sqrt(mse_gp), "improvement_%": (mse_nom - mse_gp) / max(mse_nom, 1e-12) * 100.0,
                     "r2_nom": r2(s_t2), "r2_gp": r2(s_e2), "frac_better": n_better / n})
    metrics_df = pd.DataFrame(rows)
    _save_metrics(metrics_df, save_prefix)
    se_nom, se_gp, n_b, n_w, acc_nom, acc_gp = sums["rows"]
    return {"global": block(se_nom, se_gp), "acc_only": block(acc_nom, acc_gp),
            "fractions": {"frac_better": n_b / n, "frac_worse": n_w / n, "frac_equal": 1.0 - n_b / n - n_w / n},
            "per_component": metrics_df}


def _save_metrics(metrics_df, save_prefix):
    if save_prefix is None:
        return
    base = Path(save_prefix)
    metrics_df.to_csv(base.parent / f"{base.name}_metrics.csv", index=False)
    try:
        metrics_df.to_latex(base.parent / f"{base.name}_metrics.tex", index=False, float_format="%.3e")
    except Exception as e:  # the reference reports and carries on (needs jinja2)
        print(f"Failed to save LaTeX table: {e}")


def evaluate_gp(gp, X_feat, R_true, X_state, U_ctrl, save_prefix=None):
    """evaluate_gp_offline.py:163-365: nominal vs GP-corrected derivative errors.  ONE batched mean-only predict over
    the whole file; returns the reference's dict (``global``, ``acc_only``, ``fractions``, ``per_component`` DataFrame)
    and, with ``save_prefix``, writes ``<prefix>_metrics.csv`` / ``.tex`` in the reference's column order."""
    import pandas as pd

    X_feat, R_true = np.asarray(X_feat, dtype=float), np.asarray(R_true, dtype=float)
    N = X_feat.shape[0]
    fused = getattr(gp, "predict_with_residual_metrics", None)
    if fused is not None and R_true.ndim == 2 and R_true.shape[1] == 6 and getattr(gp, "y_train_", np.empty(0)).ndim == 2 \
            and gp.y_train_.shape[1] == 6 and N > 0:
        # the B200 estimator with the reference's six residual components: the error sums are reduced on the device
        # behind the batched predict (the nominal model's error is the true residual itself, the corrected model's
        # is true minus predicted: the x_dot terms of evaluate_gp_offline.py:225-231 cancel)
        _, sums = fused(X_feat, R_true)
        return _metrics_from_sums(sums, save_prefix)
    xdot_nom = f_nominal(X_state, U_ctrl)
    xdot_true = xdot_nom + R_true
    R_pred = np.asarray(gp.predict(X_feat))
    if R_pred.ndim == 1:
        if R_pred.shape[0] == 6:
            R_pred = np.tile(R_pred, (N, 1))
        else:
            raise RuntimeError(f"GP predicted shape {R_pred.shape}, expected (N, 6) or (6,).")
    if R_pred.shape[1] > 6:
        R_pred = R_pred[:, :6]
    elif R_pred.shape[1] < 6:
        R_pred = np.hstack([R_pred, np.zeros((N, 6 - R_pred.shape[1]))])
    err_nom = xdot_true - xdot_nom
    err_gp = xdot_true - (xdot_nom + R_pred)
    se_nom, se_gp = np.sum(err_nom ** 2, axis=1), np.sum(err_gp ** 2, axis=1)

    def block(a, b):
        m_nom, m_gp = np.mean(a), np.mean(b)
        return {"mse_nom": m_nom, "mse_gp": m_gp, "rmse_nom": np.sqrt(m_nom), "rmse_gp": np.sqrt(m_gp),
                "improvement_%": (m_nom - m_gp) / max(m_nom, 1e-12) * 100.0}

    diff = se_nom - se_gp
    frac_better, frac_worse = np.mean(diff > 0.0), np.mean(diff < 0.0)
    rows = []
    for j, name in enumerate(COMPONENT_NAMES):
        e_nom, e_gp = err_nom[:, j], err_gp[:, j]
        mse_nom, mse_gp = np.mean(e_nom ** 2), np.mean(e_gp ** 2)
        rows.append({"component": name, "mse_nom": mse_nom, "mse_gp": mse_gp, "rmse_nom": np.sqrt(mse_nom),
                     "rmse_gp": np.sqrt(mse_gp), "improvement_%": (mse_nom - mse_gp) / max(mse_nom, 1e-12) * 100.0,
                     "r2_nom": r2_score(R_true[:, j], np.zeros(N)), "r2_gp": r2_score(R_true[:, j], R_pred[:, j]),
                     "frac_better": np.mean(e_nom ** 2 > e_gp ** 2)})
    metrics_df = pd.DataFrame(rows)
    _save_metrics(metrics_df, save_prefix)
    return {"global": block(se_nom, se_gp), "acc_only": block(np.sum(err_nom[:, 3:6] ** 2, axis=1),
                                                              np.sum(err_gp[:, 3:6] ** 2, axis=1)),
            "fractions": {"frac_better": frac_better, "frac_worse": frac_worse,
                          "frac_equal": 1.0 - frac_better - frac_worse},
            "per_component": metrics_df}


# ---------------------------------------------------------------------------------------- train_gp_offline.py
def train_offline(data_dir, output_dir, model_name=None, pattern="*.csv", max_data_points=10000, random_state=None):
    """train_gp_offline.py:79-215 as a function: every CSV matching ``pattern`` (schema and row filter of :22-76) ->
    ``SimpleQuadrotorGP(max_data_points).train_gp()`` -> schema-A pickle ``{'gp_model', 'training_count',
    'data_points_used', 'timestamp', 'is_trained'}`` plus a ``gp_model_latest.pkl`` link.  Returns the model path,
    or None when no usable data was found or training failed (the reference's exit code 1)."""
    files = sorted(glob.glob(os.path.join(os.path.expanduser(data_dir), pattern)))
    gp = SimpleQuadrotorGP(max_data_points=max_data_points, random_state=random_state)
    loaded = 0
    for f in files:
        try:
            X, Y = load_csv_dataset(f)
        except Exception as e:  # a malformed file is reported and skipped (:71-76)
            print(f"Failed to load {f}: {e}")
            continue
        gp.extend(X, Y)
        loaded += len(X)
    if loaded == 0:
        return None
    gp.train_gp()
    if not (gp.is_trained and gp.gp_model is not None):
        return None
    output_dir = os.path.expanduser(output_dir)
    os.makedirs(output_dir, exist_ok=True)
    name = model_name if model_name else f"gp_model_{datetime.now().strftime('%Y%m%d_%H%M%S')}"
    path = os.path.join(output_dir, f"{name}.pkl")
    with open(path, "wb") as f:
        pickle.dump({"gp_model": gp.gp_model, "training_count": gp.training_count, "data_points_used": len(gp.X_train),
                     "timestamp": datetime.now().isoformat(), "is_trained": gp.is_trained}, f)
    latest = os.path.join(output_dir, "gp_model_latest.pkl")
    try:
        if os.path.lexists(latest):
            os.remove(latest)
        os.symlink(os.path.basename(path), latest)
    except OSError as e:
        print(f"Could not create symlink: {e}")
    return path


# ---------------------------------------------------------------------------------------- gp_evaluation.py
class _IdentityScaler:
    """gp_evaluation.py:40-50: what a multi-output pickle without scalers gets"""

    def __init__(self):
        self.scale_ = np.array([1.0])

    def transform(self, X):
        return X

    def inverse_transform(self, X):
        return X


class GPModelEvaluator:
    """The model-facing half of gp_evaluation.py (:53-330): format detection (``mode`` "multi" for a dict with
    ``models`` [+ ``scalers_input`` / ``scalers_output``], "single" for a ``gp_model`` container), the seeded test-data
    generators and ``predict_on_test_data`` with its mean / std / upper / lower bands.  All members of a multi-output
    pickle answer in one device call.  Plots and the text report are not part of the path."""

    def __init__(self, model_path):
        self.model_path = model_path
        self.mode = ""
        self.gp_models, self.scalers_X, self.scalers_y = {}, {}, {}
        self.gp_model, self.n_features = None, 0
        self.training_stats, self.test_data = {}, {}
        self.load_model()

    def load_model(self):
        with open(self.model_path, "rb") as f:
            data = pickle.load(f)
        if "models" in data:
            self.mode = "multi"
            self.gp_models = {k: _to_b200(v) for k, v in data["models"].items()}
            self.scalers_X = data.get("scalers_input") or {k: _IdentityScaler() for k in self.gp_models}
            self.scalers_y = data.get("scalers_output") or {k: _IdentityScaler() for k in self.gp_models}
            self.training_stats = data.get("training_stats", {})
            return
        if "gp_model" in data:
            gp_obj = data["gp_model"]
            if not hasattr(gp_obj, "predict"):
                raise KeyError("Unsupported gp_model type inside pickle")
            self.mode = "single"
            self.gp_model = _to_b200(gp_obj)
            self.training_stats = {k: data.get(k) for k in ("training_count", "data_points_used", "timestamp", "is_trained")}
            if hasattr(gp_obj, "n_features_in_"):
                self.n_features = gp_obj.n_features_in_
            elif hasattr(gp_obj, "X_train_"):
                self.n_features = gp_obj.X_train_.shape[1]
            else:
                raise ValueError("gp_model has no 'n_features_in_' or 'X_train_'")
            return
        raise KeyError("Model file does not contain 'models' or 'gp_model'")

    def generate_physical_test_data(self, n_samples=1000):
        """gp_evaluation.py:150-207: n_samples uniform rows + 100 hover rows + a 200-row figure-8, drawn from NumPy's
        global stream after ``np.random.seed(42)`` in the reference's order, so the arrays are the reference's."""
        r = np.random
        r.seed(42)
        u = r.uniform
        base = {"x": u(-10, 10, n_samples), "y": u(-10, 10, n_samples), "z": u(-2, 15, n_samples),
                "vx": u(-5, 5, n_samples), "vy": u(-5, 5, n_samples), "vz": u(-3, 3, n_samples),
                "ax": u(-8, 8, n_samples), "ay": u(-8, 8, n_samples), "az": u(1, 18, n_samples),
                "yaw_rate": u(-1, 1, n_samples)}
        pos = u(-5, 5, (100, 3))
        hover = {"x": pos[:, 0], "y": pos[:, 1], "z": pos[:, 2] + 5.0}
        for k, (m, sd) in (("vx", (0, 0.5)), ("vy", (0, 0.5)), ("vz", (0, 0.2)), ("ax", (0, 2)), ("ay", (0, 2)),
                           ("az", (9.81, 1)), ("yaw_rate", (0, 0.3))):
            hover[k] = r.normal(m, sd, 100)
        t = np.linspace(0, 10, 200)
        traj = {"x": 3 * np.sin(0.5 * t), "y": 3 * np.sin(t), "z": 5 + 2 * np.sin(0.3 * t),
                "vx": 1.5 * np.cos(0.5 * t), "vy": 3.0 * np.cos(t), "vz": 0.6 * np.cos(0.3 * t)}
        traj["ax"] = -0.75 * np.sin(0.5 * t) + r.normal(0, 1, 200)
        traj["ay"] = -3.0 * np.sin(t) + r.normal(0, 1, 200)
        traj["az"] = -0.18 * np.sin(0.3 * t) + 9.81 + r.normal(0, 0.5, 200)
        traj["yaw_rate"] = r.normal(0, 0.2, 200)
        return {k: np.concatenate([base[k], hover[k], traj[k]]) for k in base}

    def generate_generic_test_data(self, n_samples=2000):
        np.random.seed(42)
        X = np.random.uniform(-1, 1, size=(n_samples, self.n_features))
        return {f"feature_{i}": X[:, i] for i in range(self.n_features)}

    @staticmethod
    def _bands(mean, std):
        std = np.abs(np.asarray(std).reshape(-1))
        mean = np.asarray(mean).reshape(-1)
        return {"mean": mean, "std": std, "upper": mean + 2 * std, "lower": mean - 2 * std}

    def predict_on_test_data(self, test_data):
        if self.mode == "multi":
            X = np.column_stack([test_data[k] for k in FEATURE_COLS])
            names = list(self.gp_models)
            members = PerOutputGP(self.gp_models, {k: self.scalers_X.get(k, _IdentityScaler()) for k in names},
                                  {k: self.scalers_y.get(k, _IdentityScaler()) for k in names}, names=names)
            mean, std = members.predict_batch(X)
            return {k: self._bands(mean[:, i], std[:, i]) for i, k in enumerate(names)}
        if self.mode == "single":
            if self.n_features == 10 and all(k in test_data for k in FEATURE_COLS):
                X = np.column_stack([test_data[k] for k in FEATURE_COLS])
            else:
                X = np.column_stack([test_data[k] for k in sorted(k for k in test_data if k.startswith("feature_"))])
            mean, std = self.gp_model.predict(X, return_std=True)
            return {"output": self._bands(mean, std)}
        return {}
