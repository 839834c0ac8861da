"""BASELINE config 2 proper: N=20000, D=9, T=3, C(1,fixed)*RBF(ARD)+White, alpha=1e-6, n_restarts_optimizer=9
(10 L-BFGS-B runs, random_state=0), through the estimator's public fit().  Under torchrun the 10 runs are claimed
from a shared queue by whichever rank is free (parallel.fit_sharded_restarts: same starting points as the serial fit,
one all_gather_object of (theta*, -lml*), argmin everywhere).

  python tools/full_fit_run.py
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 tools/full_fit_run.py
Rank 0 prints one JSON line and appends it to gpurun_out/full_fit.jsonl."""
import argparse, json, os, sys, time, warnings
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sklearn.gaussian_process.kernels import RBF, ConstantKernel as CK, WhiteKernel
from unmanned_aerial_vehicles_b200 import GaussianProcessRegressor
from unmanned_aerial_vehicles_b200.parallel import fit_sharded_restarts

ap = argparse.ArgumentParser()
ap.add_argument("--size", dest="n", type=int, default=20000)
ap.add_argument("--restarts", type=int, default=9)
ap.add_argument("--pilot-rows", type=int, default=0, help="hand the runs out longest first, lengths guessed by a pilot fit on this many rows (0: plain queue)")
a = ap.parse_args()
world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
D, T = 9, 3
rng = np.random.default_rng(20251129)
X = rng.standard_normal((a.n, D)); w = rng.standard_normal((D, T)) / 3
Y = np.sin(X @ w) + 0.1 * rng.standard_normal((a.n, T))
k = CK(1.0, constant_value_bounds="fixed") * RBF([1.0] * D, (0.1, 10.0)) + WhiteKernel(0.01, (1e-5, 10.0))
g = GaussianProcessRegressor(kernel=k, alpha=1e-6, n_restarts_optimizer=a.restarts, random_state=0)
GaussianProcessRegressor(kernel=k, alpha=1e-6, optimizer=None).fit(X[:2048], Y[:2048])   # library / context warm-up
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
t0 = time.perf_counter()
with warnings.catch_warnings():
    warnings.simplefilter("ignore")
    if world > 1:
        fit_sharded_restarts(g, X, Y, pilot_rows=a.pilot_rows or None)
    else:
        g.fit(X, Y)
torch.cuda.synchronize()
dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
ev = torch.tensor([float(g.n_lml_evals_)], dtype=torch.float64, device="cuda")   # counted inside fit()'s objective
if world > 1:
    dist.all_reduce(dt, op=dist.ReduceOp.MAX); dist.all_reduce(ev)
if rank == 0:
    out = {"what": "full_fit", "config": f"config2: N={a.n} D={D} T={T}, {a.restarts + 1} L-BFGS-B runs, random_state=0",
           "n_gpus": world, "fit_seconds": float(dt[0]), "lml_evaluations_all_ranks": int(ev[0]),
           "lml": float(g.log_marginal_likelihood_value_), "kernel": str(g.kernel_),
           "pilot_rows": a.pilot_rows, "pilot_order": getattr(g, "pilot_order_", None), "pilot_seconds": getattr(g, "pilot_seconds_", None)}
    print(json.dumps(out), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    open("gpurun_out/full_fit.jsonl", "a").write(json.dumps(out) + "\n")
if world > 1:
    dist.barrier(); dist.destroy_process_group()
