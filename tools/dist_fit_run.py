"""torchrun driver for a hyper-parameter FIT on the distributed factor (BASELINE config 5: "an N=200k fit completing"):
DistributedGPR = SciPy L-BFGS-B on every rank over DistGP.lml_grad (block-cyclic Cholesky + TRTRI + LAUUM + gradient).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 \
      tools/dist_fit_run.py --size 200000 --maxiter 2 [--grid 2x4] [--block 2048] [--check]
--check (small sizes): the single-GPU estimator's fit with the same optimiser settings on rank 0.
Rank 0 prints one JSON line and appends it to gpurun_out/dist_fit.jsonl."""
import argparse, json, os, sys, time, warnings
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sklearn.gaussian_process.kernels import RBF, ConstantKernel as CK, WhiteKernel
from unmanned_aerial_vehicles_b200.dist_fit import DistributedGPR

ap = argparse.ArgumentParser()
ap.add_argument("--size", dest="n", type=int, default=200000)
ap.add_argument("--block", dest="nb", type=int, default=2048)
ap.add_argument("--grid", default="")
ap.add_argument("--maxiter", type=int, default=2)
ap.add_argument("--maxfun", type=int, default=0, help="cap on objective evaluations of the run (0: SciPy's default)")
ap.add_argument("--no-predict", action="store_true")
ap.add_argument("--check", action="store_true")
a = ap.parse_args()
world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
D = 9
rng = np.random.default_rng(20251129)
X = rng.standard_normal((a.n, D)); w = rng.standard_normal((D, 1)) / 3
y = (np.sin(X @ w) + 0.1 * rng.standard_normal((a.n, 1)))[:, 0]
k = CK(1.0, constant_value_bounds="fixed") * RBF([1.0] * D, (0.1, 10.0)) + WhiteKernel(0.01, (1e-5, 10.0))
grid = tuple(int(v) for v in a.grid.split("x")) if a.grid else None
os.makedirs("gpurun_out", exist_ok=True)
def progress(kk, lml, sec):          # one line per objective evaluation, so a cut-off run still leaves its evidence
    if rank == 0:
        open("gpurun_out/dist_fit_progress.jsonl", "a").write(json.dumps(
            {"n": a.n, "n_gpus": world, "evaluation": kk, "lml": lml, "seconds": sec, "t": time.time()}) + "\n")
g = DistributedGPR(k, alpha=1e-6, maxiter=a.maxiter if a.maxiter > 0 else None, maxfun=a.maxfun if a.maxfun > 0 else None,
                   on_eval=progress, NB=a.nb, grid=grid)
torch.cuda.synchronize()
if world > 1: dist.barrier()
t0 = time.perf_counter()
g.fit(X, y)
torch.cuda.synchronize()
if world > 1: dist.barrier()
dt = time.perf_counter() - t0
mem = torch.tensor([torch.cuda.max_memory_allocated() / 2 ** 30], dtype=torch.float64, device="cuda")
if world > 1: dist.all_reduce(mem, op=dist.ReduceOp.MAX)
if rank == 0:
    open("gpurun_out/dist_fit_progress.jsonl", "a").write(json.dumps(
        {"n": a.n, "n_gpus": world, "fit_completed_s": dt, "evaluations": g.n_lml_evals_, "kernel": str(g.kernel_),
         "lml": float(g.log_marginal_likelihood_value_)}) + "\n")
Xq = X[:512] + 0.05
tp = time.perf_counter()
if a.no_predict:
    m = s = None
else:
    m, s = g.predict(Xq, return_std=True)
torch.cuda.synchronize()
tp = time.perf_counter() - tp
out = {"what": "dist_fit", "n": a.n, "nb": a.nb, "n_gpus": world, "grid": f"{g._gp.Pr}x{g._gp.Pc}", "maxiter": a.maxiter,
       "fit_seconds": dt, "lml_evaluations": g.n_lml_evals_, "seconds_per_evaluation": [round(v, 3) for v in g.evaluation_seconds_],
       "theta": [float(v) for v in g.kernel_.theta], "kernel": str(g.kernel_), "lml": float(g.log_marginal_likelihood_value_),
       "max_mem_gib_per_gpu": float(mem[0]), "maxfun": a.maxfun, "predict_512_rows_mean_std_s": None if a.no_predict else tp,
       "evaluation_tflops_per_gpu_on_N3": float(a.n) ** 3 / float(np.median(g.evaluation_seconds_)) / 1e12 / world}
if a.check and rank == 0:
    from unmanned_aerial_vehicles_b200 import GaussianProcessRegressor
    import scipy.optimize

    def opt(obj, x0, bounds):
        r = scipy.optimize.minimize(obj, x0, method="L-BFGS-B", jac=True, bounds=bounds,
                                    options={} if a.maxiter <= 0 else {"maxiter": a.maxiter})
        return r.x, r.fun
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        e = GaussianProcessRegressor(kernel=k, alpha=1e-6, optimizer=opt).fit(X, y)
    me, se = e.predict(Xq, return_std=True)
    out.update({"check_theta_maxabs": float(np.max(np.abs(e.kernel_.theta - g.kernel_.theta))),
                "check_lml_rel": abs(e.log_marginal_likelihood_value_ - g.log_marginal_likelihood_value_) / abs(e.log_marginal_likelihood_value_),
                "check_mean_maxabs": float(np.max(np.abs(me - m))), "check_std_maxabs": float(np.max(np.abs(se - s)))})
if rank == 0:
    print(json.dumps(out), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    open("gpurun_out/dist_fit.jsonl", "a").write(json.dumps(out) + "\n")
if world > 1:
    dist.barrier(); dist.destroy_process_group()
