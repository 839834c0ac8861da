"""BASELINE config 4: N=50000, D=9, three single-output ARD models x 32 optimiser runs (n_restarts_optimizer=31,
random_state=t), through parallel.fit_outputs_sharded_restarts: outputs one after the other, each model's runs
sharded over all ranks (96 runs, 12 per GPU on 8), device buffers released between outputs.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29514 tools/config4_run.py
  (--size / --restarts / --outputs shrink it for a dry run; estimate for the full job on 8xB200: ~22 min)
Rank 0 prints one JSON line and appends it to gpurun_out/config4.jsonl."""
import argparse, json, os, sys, time, warnings
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sklearn.gaussian_process.kernels import RBF, ConstantKernel as CK, WhiteKernel
from unmanned_aerial_vehicles_b200 import GaussianProcessRegressor, _lib
from unmanned_aerial_vehicles_b200.parallel import fit_outputs_sharded_restarts

ap = argparse.ArgumentParser()
ap.add_argument("--size", dest="n", type=int, default=50000)
ap.add_argument("--restarts", type=int, default=31)
ap.add_argument("--outputs", type=int, default=3)
ap.add_argument("--maxiter", type=int, default=0, help="truncate every L-BFGS-B run to this many iterations (0: SciPy's default)")
a = ap.parse_args()
world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
D = 9
rng = np.random.default_rng(20251129)
X = rng.standard_normal((a.n, D)); w = rng.standard_normal((D, 3)) / 3
Y = (np.sin(X @ w) + 0.1 * rng.standard_normal((a.n, 3)))[:, : a.outputs]
k = CK(1.0, constant_value_bounds="fixed") * RBF([1.0] * D, (0.1, 10.0)) + WhiteKernel(0.01, (1e-5, 10.0))
def truncated(obj_func, x0, bounds):
    import scipy.optimize
    r = scipy.optimize.minimize(obj_func, x0, method="L-BFGS-B", jac=True, bounds=bounds, options={"maxiter": a.maxiter})
    return r.x, r.fun
make = lambda t: GaussianProcessRegressor(kernel=k, alpha=1e-6, n_restarts_optimizer=a.restarts, random_state=t,
                                          optimizer=truncated if a.maxiter > 0 else "fmin_l_bfgs_b")
GaussianProcessRegressor(kernel=k, alpha=1e-6, optimizer=None).fit(X[:2048], Y[:2048, 0])   # library / context warm-up
lib = _lib.load()
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
l0, t0 = int(lib.gpk_launch_count()), time.perf_counter()
with warnings.catch_warnings():
    warnings.simplefilter("ignore")
    models = fit_outputs_sharded_restarts(make, X, Y)
torch.cuda.synchronize()
dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
mem = torch.tensor([torch.cuda.max_memory_allocated() / 2 ** 30], dtype=torch.float64, device="cuda")
evals = torch.tensor([float(sum(m.n_lml_evals_ for m in models))], dtype=torch.float64, device="cuda")
ev_max = evals.clone()
if world > 1:
    dist.all_reduce(dt, op=dist.ReduceOp.MAX); dist.all_reduce(mem, op=dist.ReduceOp.MAX)
    dist.all_reduce(evals); dist.all_reduce(ev_max, op=dist.ReduceOp.MAX)
if rank == 0:
    out = {"what": "config4", "config": f"N={a.n} D={D}, {a.outputs} single-output ARD models x {a.restarts + 1} L-BFGS-B runs",
           "n_gpus": world, "fit_seconds": float(dt[0]), "runs_total": a.outputs * (a.restarts + 1), "maxiter": a.maxiter,
           "lml_evals_all_ranks": int(evals[0]), "lml_evals_busiest_rank": int(ev_max[0]),
           "seconds_per_evaluation_busiest_rank": float(dt[0]) / max(float(ev_max[0]), 1.0),
           "kernel_launches_rank0": int(lib.gpk_launch_count()) - l0, "max_mem_gib_per_gpu": float(mem[0]),
           "lml": [float(m.log_marginal_likelihood_value_) for m in models], "kernels": [str(m.kernel_) for m in models]}
    print(json.dumps(out), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    open("gpurun_out/config4.jsonl", "a").write(json.dumps(out) + "\n")
if world > 1:
    dist.barrier(); dist.destroy_process_group()
