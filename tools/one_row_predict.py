"""One-row predict(return_std=True) at the reference's production shape, for an ncu launch list / host timing split."""
import os, sys, time, warnings
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from sklearn.gaussian_process.kernels import RBF, WhiteKernel
from unmanned_aerial_vehicles_b200 import GaussianProcessRegressor
g = np.load(os.path.join(ROOT, "tests", "golden", "simple_gp_csv.npz"))
X, Y, theta = g["X"], g["Y"], g["theta"]
kern = RBF(length_scale=float(np.exp(theta[0]))) + WhiteKernel(noise_level=float(np.exp(theta[1])))
with warnings.catch_warnings():
    warnings.simplefilter("ignore")
    m = GaussianProcessRegressor(kernel=kern, alpha=1e-4, normalize_y=True, optimizer=None).fit(X, Y)
x = X[:1] + 0.01
for _ in range(5):
    m.predict(x, return_std=True)
eng = m._engine
t0 = time.perf_counter()
for _ in range(200):
    m.predict(x, return_std=True)
t1 = time.perf_counter()
for _ in range(200):
    eng.predict(x, m._y_train_mean, m._y_train_std, True)
t2 = time.perf_counter()
for _ in range(200):
    eng.predict(x, m._y_train_mean, m._y_train_std, False)
t3 = time.perf_counter()
print(f"estimator.predict(std) {1e6*(t1-t0)/200:.1f} us; engine.predict(std) {1e6*(t2-t1)/200:.1f} us; engine.predict(mean) {1e6*(t3-t2)/200:.1f} us")
