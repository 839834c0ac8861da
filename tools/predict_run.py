"""BASELINE config 3: batched GP predict mean+std, 4096 trajectories x 20-step MPC horizon x reps >= 1e7 query
points against the N=20k, D=9, T=3 training set of config 2 (theta fixed at the optimum of the full fit,
profiles/r01b_bench_fullfit.json), query rows sharded over the ranks (no collective in the data path).

  python tools/predict_run.py [--reps 123]                                   # 1 GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node P --master-addr 127.0.0.1 --master-port 29512 \
      tools/predict_run.py [--reps 123]

Every rep is one (81920, 9) host array -> pinned H2D -> cross-Gram / mean / L^-1 contraction / std -> D2H of
mean (81920,3) and std (81920,3): the timed region is end to end.  Rank 0 prints one JSON line and appends it to
gpurun_out/predict_run.jsonl."""
import argparse, ctypes as C, json, os, sys, time
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unmanned_aerial_vehicles_b200 import _lib
from unmanned_aerial_vehicles_b200.engine import DeviceGP
from unmanned_aerial_vehicles_b200.parallel import row_block

ap = argparse.ArgumentParser()
ap.add_argument("--size", dest="n", type=int, default=20000)
ap.add_argument("--reps", type=int, default=123)
ap.add_argument("--traj", type=int, default=4096)
ap.add_argument("--horizon", type=int, default=20)
a = ap.parse_args()
world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
D, T = 9, 3
rng = np.random.default_rng(20251129)
X = rng.standard_normal((a.n, D)); w = rng.standard_normal((D, T)) / 3
Y = np.sin(X @ w) + 0.1 * rng.standard_normal((a.n, T))
ls = np.array([2.91, 2.21, 4.09, 4.53, 3.17, 4.25, 9.06, 3.86, 3.95]); noise, jit = 0.00996, 1e-6
lib = _lib.load()
eng = DeviceGP(X, Y, jit)
assert eng.factor(1.0 / ls, 1.0, noise) == 0          # every rank refactors from theta (no 3.2 GB broadcast)
rows_per_rep = a.traj * a.horizon
lo, hi = row_block(a.reps, rank, world)               # reps are the shard unit (contiguous row blocks of M)


def queries(rep):
    """smooth MPC-like roll-outs: random start N(0,1)^9 + random walk 0.05 N(0,1) per step (seed 7 + rep)"""
    r = np.random.default_rng(7 + rep)
    return (r.standard_normal((a.traj, 1, D)) + 0.05 * np.cumsum(r.standard_normal((a.traj, a.horizon, D)), 1)).reshape(-1, D)


eng.predict(queries(0)[:4096], 0.0, 1.0, True)        # warm-up
batches = [queries(r) for r in range(lo, hi)]         # host arrays exist before the timed region starts
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
l0 = lib.gpk_launch_count()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter(); e0.record()
chk = 0.0
for Xq in batches:
    mean, std, _, neg = eng.predict(Xq, 0.0, 1.0, True)
    chk += float(mean.sum()) + float(std.sum())
e1.record(); torch.cuda.synchronize()
wall = time.perf_counter() - t0
dev = e0.elapsed_time(e1) * 1e-3
launches = int(lib.gpk_launch_count() - l0)
tt = torch.tensor([dev, wall], dtype=torch.float64, device="cuda")
cs = torch.tensor([chk, float(launches)], dtype=torch.float64, device="cuda")
if world > 1:
    dist.all_reduce(tt, op=dist.ReduceOp.MAX); dist.all_reduce(cs)
dev, wall = float(tt[0]), float(tt[1])
Mtot = a.reps * rows_per_rep
if rank == 0:
    sink = torch.zeros(8, dtype=torch.float64, device="cuda"); pf = C.c_double(0); peak = 0.0
    for _ in range(3):
        e0.record()
        lib.gpk_dmma_probe(296, 20000, C.c_void_p(sink.data_ptr()), C.byref(pf), C.c_void_p(torch.cuda.current_stream().cuda_stream))
        e1.record(); torch.cuda.synchronize()
        peak = max(peak, pf.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    flops_pt = float(eng.NP) ** 2 + 2.0 * eng.NP * (D + 4 + T)       # L^-1 contraction (tri) + cross-Gram + mean
    out = {"what": "predict_mean_std", "config": f"config3: N={a.n} D={D} T={T}, M={Mtot} = {a.traj} traj x {a.horizon} steps x {a.reps} reps",
           "n_gpus": world, "sharding": "query rows (reps) over ranks, factor recomputed per rank, no collective",
           "pts_per_s_device_timed": Mtot / dev, "pts_per_s_e2e_wall": Mtot / wall, "seconds": dev,
           "h2d_bytes_per_rep": rows_per_rep * D * 8, "d2h_bytes_per_rep": rows_per_rep * 2 * T * 8,
           "fp64_tflops_total": Mtot * flops_pt / dev / 1e12, "fp64_tflops_per_gpu": Mtot * flops_pt / dev / 1e12 / world,
           "fp64_dmma_peak_probe": peak, "frac_of_fp64_peak_per_gpu": Mtot * flops_pt / dev / 1e12 / world / peak,
           "algorithmic_hbm_GBps": Mtot * 8 * (D + 2 * T) / dev / 1e9, "gpu_launches": int(cs[1]), "checksum": float(cs[0])}
    print(json.dumps(out), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    open("gpurun_out/predict_run.jsonl", "a").write(json.dumps(out) + "\n")
if world > 1:
    dist.barrier(); dist.destroy_process_group()
