"""torchrun driver for the distributed GP factorisation (BASELINE config 5).
  python -m torch.distributed.run --nnodes=1 --nproc-per-node P --master-addr 127.0.0.1 --master-port 29511 \
      tools/dist_chol_run.py --size 200000 --block 2048 [--grid 2x4] [--no-lookahead] [--check]
Rank 0 prints one JSON line and appends it to gpurun_out/dist_chol.jsonl."""
import argparse, json, os, sys, time
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unmanned_aerial_vehicles_b200.dist_chol import DistGP

ap = argparse.ArgumentParser()
ap.add_argument("--size", dest="n", type=int, default=200000); ap.add_argument("--block", dest="nb", type=int, default=2048)
ap.add_argument("--check", action="store_true")
ap.add_argument("--grid", default="", help="PrxPc, e.g. 2x4 (default: most square)")
ap.add_argument("--no-lookahead", action="store_true")
ap.add_argument("--trace", action="store_true", help="CUDA-event spans per phase and panel (rank 0's and the slowest rank's summary)")
ap.add_argument("--grad", action="store_true", help="also the distributed inverse (TRTRI + LAUUM) and the LML gradient: one full objective evaluation")
ap.add_argument("--predict-rows", type=int, default=0, help="time DistGP.predict(return_std=True) on this many query rows")
a = ap.parse_args()
world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
D, T = 9, (3 if a.n <= 50000 else 1)
rng = np.random.default_rng(20251129)
X = rng.standard_normal((a.n, D)); w = rng.standard_normal((D, T)) / 3
Y = np.sin(X @ w) + 0.1 * rng.standard_normal((a.n, T))
ls, noise, jit = np.ones(D), 0.01, 1e-6
t0 = time.perf_counter()
grid = tuple(int(v) for v in a.grid.split("x")) if a.grid else None
g = DistGP(X, Y, jit, NB=a.nb, grid=grid, lookahead=not a.no_lookahead)
torch.cuda.synchronize()
if world > 1: dist.barrier()
if a.trace:
    g.ops.trace = []
g.ops.bcast_bytes = g.ops.bcast_calls = 0
t1 = time.perf_counter()
info = g.factor(1.0 / ls, 1.0, noise)
torch.cuda.synchronize()
if world > 1: dist.barrier()
t2 = time.perf_counter()
g.solve(); lml = g.lml()
torch.cuda.synchronize()
t3 = time.perf_counter()
fl = torch.tensor([g.gemm_flops], dtype=torch.float64, device="cuda")
mem = torch.tensor([torch.cuda.max_memory_allocated() / 2 ** 30], dtype=torch.float64, device="cuda")
if world > 1:
    dist.all_reduce(fl); dist.all_reduce(mem, op=dist.ReduceOp.MAX)
out = {"what": "dist_chol", "n": a.n, "nb": a.nb, "n_gpus": world, "grid": f"{g.Pr}x{g.Pc}", "lookahead": g.lookahead, "T": T, "info": info, "alloc_s": t1 - t0,
       "assemble_factor_s": t2 - t1, "solve_lml_s": t3 - t2, "lml": lml, "potrf_alg_tflops": (a.n ** 3 / 3) / (t2 - t1) / 1e12,
       "executed_gemm_tflops": float(fl[0]) / (t2 - t1) / 1e12, "max_mem_gib_per_gpu": float(mem[0]),
       "local_gib": g.local_bytes() / 2 ** 30}
out["bcast_GB_rank0"] = g.ops.bcast_bytes / 1e9
out["bcast_calls_rank0"] = g.ops.bcast_calls
if a.trace:
    # per-phase seconds on this rank's streams: main = diag_potrf_inv, panel_gemm, update_*, wait_panel (stall for the
    # next panel's broadcasts); side = bcast_panel_rows / bcast_panel_cols; bcast_diag_inverse sits on the main stream
    summ = g.ops.trace_summary()
    g.ops.trace = None
    allr = [None] * world
    if world > 1:
        dist.all_gather_object(allr, summ)
    else:
        allr = [summ]
    out["phases_rank0"] = {k: round(v["seconds"], 4) for k, v in sorted(allr[0].items())}
    out["phases_max_over_ranks"] = {k: round(max(r.get(k, {"seconds": 0})["seconds"] for r in allr), 4) for k in sorted(allr[0])}
    out["idle_before_phase_rank0"] = {k: round(v["idle_before_s"], 4) for k, v in sorted(allr[0].items()) if v["idle_before_s"] > 0}
    out["idle_before_phase_max_over_ranks"] = {k: round(max(r.get(k, {"idle_before_s": 0})["idle_before_s"] for r in allr), 4)
                                               for k in sorted(allr[0])}
    out["phases_min_over_ranks"] = {k: round(min(r.get(k, {"seconds": 0})["seconds"] for r in allr), 4) for k in sorted(allr[0])}
    big = max(allr[0].get("bcast_panel_rows", {"seconds": 0})["seconds"], 1e-9)
    out["bcast_panel_rows_GBps_rank0"] = out["bcast_GB_rank0"] / big if "bcast_panel_rows" in allr[0] else None
if a.grad:
    if a.trace:
        g.ops.trace = []
    b0 = g.ops.bcast_bytes
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    tg = time.perf_counter()
    g.inverse()
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    t_inv = time.perf_counter() - tg
    tg = time.perf_counter()
    g_c, g_ls, g_noise = g.gradient()
    torch.cuda.synchronize()
    t_grad = time.perf_counter() - tg
    memg = torch.tensor([torch.cuda.max_memory_allocated() / 2 ** 30], dtype=torch.float64, device="cuda")
    if world > 1: dist.all_reduce(memg, op=dist.ReduceOp.MAX)
    out.update({"inverse_trtri_lauum_s": t_inv, "inverse_alg_tflops_on_2N3_over_3": (2 * a.n ** 3 / 3) / t_inv / 1e12,
                "gradient_contraction_s": t_grad, "objective_evaluation_with_gradient_s": (t2 - t1) + (t3 - t2) + t_inv + t_grad,
                "grad_ls": [float(v) for v in g_ls], "grad_noise": float(g_noise), "grad_c": float(g_c),
                "inverse_bcast_GB_rank0": (g.ops.bcast_bytes - b0) / 1e9, "max_mem_gib_per_gpu_with_inverse": float(memg[0])})
    if a.trace:
        summ = g.ops.trace_summary(); g.ops.trace = None
        out["inverse_phases_rank0"] = {k: round(v["seconds"], 4) for k, v in sorted(summ.items())}
        out["inverse_idle_before_phase_rank0"] = {k: round(v["idle_before_s"], 4) for k, v in sorted(summ.items()) if v["idle_before_s"] > 0}
    if a.check and rank == 0:
        from unmanned_aerial_vehicles_b200.engine import DeviceGP
        e3 = DeviceGP(X, Y, jit)
        _, gc_e, gls_e, gn_e = e3.lml(1.0 / ls, 1.0, noise, True)
        out["check_grad_ls_maxrel"] = float(np.max(np.abs(g_ls - gls_e)) / np.max(np.abs(gls_e)))
        out["check_grad_noise_rel"] = abs(g_noise - gn_e) / abs(gn_e)
        del e3
if a.predict_rows > 0:                                  # collective: every rank runs it on the same rows
    Xq = X[np.random.default_rng(7).integers(0, a.n, a.predict_rows)] + 0.05
    g.predict(Xq[:256], return_std=True)                # warm-up
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    tp = time.perf_counter()
    pm, ps = g.predict(Xq, return_std=True)
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    tp = time.perf_counter() - tp
    out.update({"predict_rows": a.predict_rows, "predict_mean_std_s": tp, "predict_pts_per_s": a.predict_rows / tp,
                "predict_fp64_tflops_aggregate": a.predict_rows * float(g.NP) ** 2 / tp / 1e12})
    if a.check and rank == 0:
        from unmanned_aerial_vehicles_b200.engine import DeviceGP
        e2 = DeviceGP(X, Y, jit)
        assert e2.factor(1.0 / ls, 1.0, noise) == 0
        m_e, s_e, _, _ = e2.predict(Xq, np.zeros(T), np.ones(T), True)
        out["check_predict_mean_maxabs"] = float(np.max(np.abs(pm - m_e)))
        out["check_predict_std_maxabs"] = float(np.max(np.abs(ps - s_e)))
        del e2
if a.check and rank == 0:
    from unmanned_aerial_vehicles_b200.engine import DeviceGP
    eng = DeviceGP(X, Y, jit)
    assert eng.factor(1.0 / ls, 1.0, noise) == 0
    ref = eng.lml(1.0 / ls, 1.0, noise, False)[0]
    eng.factor(1.0 / ls, 1.0, noise)
    out["check_lml_rel"] = abs(lml - ref) / abs(ref)
    out["check_alpha_maxrel"] = float(np.max(np.abs(g.alpha_host() - eng.alpha_host())) / np.max(np.abs(eng.alpha_host())))
if rank == 0:
    print(json.dumps(out), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    open("gpurun_out/dist_chol.jsonl", "a").write(json.dumps(out) + "\n")
if world > 1:
    dist.barrier(); dist.destroy_process_group()
