#!/usr/bin/env python
"""bench.py -- BASELINE.json config 2 ("Synthetic N=20k, D=9 exact GP fit: Gram + Cholesky + LML/grad").

A step = ONE log-marginal-likelihood + gradient evaluation (the unit L-BFGS-B repeats ~30x per restart):
Gram assembly -> blocked Cholesky fused with the triangular inverse -> alpha -> K^-1 -> gradient contraction.
`value` = evaluations/s over all ranks (restarts are sharded across GPUs: each rank evaluates its own theta
sequence on a replicated training set, no data-path collective -> weak scaling).

The `extra` block carries the other quantities of BASELINE.json's metric and the paths that DO communicate, at
every N (each under its own guard: a failure is recorded in place, the headline line still prints):
  full_fit        config 2 proper, 10 L-BFGS-B runs through GaussianProcessRegressor.fit -- "GP fit s @N=20k";
                  at N > 1 the runs are claimed from a shared queue (parallel.fit_sharded_restarts): strong scaling
  predict_sharded config 3 shape, 81 920 x N query rows through parallel.predict_sharded, gather inside the timed
                  region (all_gather_into_tensor over NCCL); mean+std and mean-only
  dist_chol       config 5 path, DistGP.factor + solve + lml (2-D block-cyclic Cholesky, NCCL panel broadcasts) and
                  the gradient from the distributed inverse, at N_train = 20k / 100k / 140k / 200k for 1 / 2 / 4 / 8
                  GPUs, checked against the single-GPU engine at 20k

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--skip full_fit,predict_sharded,...]
"""
import os
import sys

if any(a == "reference" or a.endswith("=reference") for a in sys.argv[1:]):
    # The CPU arm uses every host core.  torch.distributed.run exports OMP_NUM_THREADS=1 when nproc > 1, and the BLAS
    # libraries read it once, at load: lift it here, before NumPy / SciPy are imported.
    for _v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[_v] = str(os.cpu_count() or 1)

import argparse  # noqa: E402
import ctypes as C  # noqa: E402
import json  # noqa: E402
import subprocess  # noqa: E402
import tempfile  # noqa: E402
import threading  # noqa: E402
import time  # noqa: E402

import numpy as np  # noqa: E402

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

D, T = 9, 3
JITTER, NOISE0 = 1e-6, 0.01


def make_data(n):
    """SURVEY.md section 8(d) config 2 (config 4 with --n 50000 --targets 1: one single-output model per run)"""
    rng = np.random.default_rng(20251129)
    X = rng.standard_normal((n, D))
    w = rng.standard_normal((D, 3)) / 3
    Y = np.sin(X @ w) + 0.1 * rng.standard_normal((n, 3))
    return X, np.ascontiguousarray(Y[:, :T])


def workload_name(n):
    tag = "config2" if n == 20000 else "config4-shaped" if n == 50000 else "custom"
    return (f"{tag}: N={n} D={D} T={T} ARD-RBF+White LML+grad evaluation "
            "(Gram + Cholesky/inverse + K^-1 + gradient), one theta per step")


def bench_config(n):
    """the SAME dict in both arms (the driver compares them)"""
    npad = -(-n // 128) * 128
    return {"workload": workload_name(n),
            "sharding": "restarts x ranks, training set replicated, no collective in the data path",
            "l2": f"inputs (3 x {8e-9 * npad ** 2:.1f} GB factorisation buffers) larger than L2; no flush needed"}


def metric_name(n):
    return "gp_fit_lml_grad_evals_per_s_N%dk" % (n // 1000)


def theta_sequence(rank, count):
    """log-parameters [log l_1..l_9, log noise] an optimiser run would visit (seeded per rank/restart)"""
    rng = np.random.default_rng(1000 + rank)
    base = np.concatenate([np.zeros(D), [np.log(NOISE0)]])
    return [base + 0.2 * rng.standard_normal(D + 1) for _ in range(count)]


def ard_kernel():
    """gp_trainer.py:163-166"""
    from sklearn.gaussian_process.kernels import RBF, ConstantKernel as CK, WhiteKernel
    return CK(1.0, constant_value_bounds="fixed") * RBF([1.0] * D, (0.1, 10.0)) + WhiteKernel(NOISE0, (1e-5, 10.0))


# ------------------------------------------------------------------------------------------------
# CPU side: the oracle port of sklearn's path (and stock sklearn beside it), all host cores
def host_threads():
    """lift whatever thread cap the launcher left on the BLAS pools; report what is in force"""
    import threadpoolctl

    ncpu = os.cpu_count() or 1
    limiter = threadpoolctl.threadpool_limits(limits=ncpu)        # stays in force for the rest of the process
    info = [{"api": d.get("internal_api"), "threads": d.get("num_threads"), "lib": os.path.basename(d.get("filepath", ""))}
            for d in threadpoolctl.threadpool_info()]
    return limiter, ncpu, info


def cpu_full_evaluations(n, budget_s, max_evals):
    """Full-size LML+grad evaluations of the oracle port (sklearn _gpr.py:583-656 restated without the
    (N, N, n_theta) tensor): at least one, then more while the time budget lasts.  Measured, never extrapolated."""
    from oracle import gp_oracle as orc

    X, Y = make_data(n)
    Xw, Yw = X[:1500], Y[:1500]
    orc.lml_and_grad(Xw, Yw, np.ones(D), 1.0, NOISE0, JITTER)       # loads the BLAS / LAPACK paths once, untimed
    ts = []
    for th in theta_sequence(0, max_evals):
        t0 = time.perf_counter()
        lml, _ = orc.lml_and_grad(X, Y, np.exp(th[:D]), 1.0, float(np.exp(th[D])), JITTER)
        ts.append(time.perf_counter() - t0)
        assert np.isfinite(lml)
        if sum(ts) + ts[-1] > budget_s:
            break
    return ts


def stock_sklearn_sample(n_sample, n_full):
    """The reference's own arithmetic, unrestated: scikit-learn's GaussianProcessRegressor.log_marginal_likelihood(theta,
    eval_gradient=True) (_gpr.py:541-656; it materialises the (N, N, n_theta) gradient tensor, so it is run at a size
    that fits: 10 x 8 N^2 bytes) -- one evaluation, with its cubic part scaled by r^3 and the rest by r^2 to n_full.
    Reported beside the measured port figure; never the denominator."""
    import warnings

    from scipy.linalg import cho_solve, cholesky
    from sklearn.gaussian_process import GaussianProcessRegressor as SkGPR

    X, Y = make_data(n_sample)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        g = SkGPR(kernel=ard_kernel(), alpha=JITTER, optimizer=None).fit(X, Y)
        th = theta_sequence(0, 1)[0]
        t0 = time.perf_counter()
        g.log_marginal_likelihood(th, eval_gradient=True)
        t_all = time.perf_counter() - t0
    K = g.kernel_(X)
    K[np.diag_indices_from(K)] += JITTER
    t0 = time.perf_counter()
    L = cholesky(K, lower=True, check_finite=False)
    cho_solve((L, True), np.eye(n_sample), check_finite=False)
    t_cubic = min(time.perf_counter() - t0, t_all)
    r = n_full / n_sample
    return {"n": n_sample, "eval_s": t_all, "cubic_s": t_cubic,
            "scaled_eval_s_extrapolated": t_cubic * r ** 3 + (t_all - t_cubic) * r ** 2,
            "what": "sklearn GaussianProcessRegressor.log_marginal_likelihood(theta, eval_gradient=True), one evaluation"}


def cpu_baseline_block(n, budget_s, max_evals, with_stock):
    import scipy.linalg  # noqa: F401 -- SciPy ships its own OpenBLAS: load it before the pools are resized/listed
    from oracle import gp_oracle  # noqa: F401

    limiter, ncpu, pools = host_threads()
    ts = cpu_full_evaluations(n, budget_s, max_evals)
    eval_s = float(np.median(ts))
    block = {"value": 1.0 / eval_s, "unit": "evals/s", "cores": ncpu, "kind": "port",
             "sample": (f"{len(ts)} full N={n} LML+grad evaluation(s) of the oracle port, measured: "
                        + ", ".join(f"{t:.1f}" for t in ts) + " s (median taken); no extrapolation"),
             "eval_s": eval_s, "blas_threads": pools,
             "launcher_omp_num_threads": os.environ.get("OMP_NUM_THREADS")}
    if with_stock:
        block["stock_sklearn"] = stock_sklearn_sample(min(3000, n), n)
    return block, ts


def reference_arm(args):
    """The reference's CPU implementation of the path (oracle port of sklearn on SciPy LAPACK / OpenBLAS) on every
    host core.  One step = one full N=20k evaluation (~50 s on 16 cores), so the timed steps are capped by a time
    budget whatever --steps says; the line reports the number actually timed."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    block, ts = cpu_baseline_block(args.n, args.cpu_budget_s, max(1, args.steps), not args.no_stock_sklearn)
    eval_s = block["eval_s"]
    line = {
        "impl": "reference", "metric": metric_name(args.n), "value": 1.0 / eval_s, "unit": "evals/s",
        "n_gpus": args.gpus, "steps": len(ts), "warmup": 0, "steps_requested": args.steps,
        "warmup_requested": args.warmup, "ms_per_step": eval_s * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": bench_config(args.n),
        "reference_engine": "oracle port of sklearn 1.9.0 _gpr.py:583-656 on SciPy LAPACK/OpenBLAS, all host cores; "
                            "rank 0 only (the CPU path has no multi-process form: value does not grow with n_gpus)",
        "cpu_baseline": block,
        "e2e": {"value": 1.0 / eval_s, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-i", str(gpu_index), "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = [ln.split(",") for ln in open(self.f.name).read().strip().splitlines() if ln.count(",") >= 8]
        os.unlink(self.f.name)
        sm, reasons = [], set()
        for r in rows:
            try:
                sm.append(float(r[1])); smax = float(r[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.strip().lower() == "active":
                    reasons.add(name)
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": smax, "reasons": sorted(reasons),
                   "samples": len(sm)}
        return out


class Ctx:
    """what every measurement of the GPU arm shares"""

    def __init__(self):
        import torch
        import torch.distributed as dist

        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, seconds):
        t = self.torch.tensor([seconds], dtype=self.torch.float64, device="cuda")
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t[0])

    def gather_obj(self, obj):
        if self.world == 1:
            return [obj]
        out = [None] * self.world
        self.dist.all_gather_object(out, obj)
        return out


def fp64_peaks(ctx, lib):
    """FP64 tensor-pipe peak: the register-resident DMMA probe (the roofline's denominator) and, as a library
    cross-check in the same run, torch.matmul (cuBLAS dgemm) at 8192^3."""
    torch = ctx.torch
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sink = torch.zeros(8, dtype=torch.float64, device="cuda")
    pf, best = C.c_double(0), 0.0
    for _ in range(3):
        e0.record()
        lib.gpk_dmma_probe(296, 20000, C.c_void_p(sink.data_ptr()), C.byref(pf),
                           C.c_void_p(torch.cuda.current_stream().cuda_stream))
        e1.record(); torch.cuda.synchronize()
        best = max(best, pf.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    a = torch.randn(8192, 8192, dtype=torch.float64, device="cuda")
    b = torch.randn(8192, 8192, dtype=torch.float64, device="cuda")
    c = torch.empty_like(a)
    torch.matmul(a, b, out=c)
    lib_best = 0.0
    for _ in range(3):
        e0.record(); torch.matmul(a, b, out=c); e1.record(); torch.cuda.synchronize()
        lib_best = max(lib_best, 2 * 8192.0 ** 3 / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    del a, b, c
    torch.cuda.empty_cache()
    return best, lib_best


# ------------------------------------------------------------------------------------------------
# extras: the metric's other quantities and the communicating paths (every rank calls each of them)
def extra_full_fit(ctx, X, Y):
    """BASELINE config 2 proper -- the metric's "GP fit s @N=20k": 10 L-BFGS-B runs (n_restarts_optimizer=9,
    random_state=0) through GaussianProcessRegressor.fit; at N > 1 the runs are claimed from a shared queue by
    whichever rank is free (parallel.fit_sharded_restarts), total work fixed -> strong scaling."""
    import warnings

    from unmanned_aerial_vehicles_b200 import GaussianProcessRegressor, parallel

    g = GaussianProcessRegressor(kernel=ard_kernel(), alpha=JITTER, n_restarts_optimizer=9, random_state=0)
    ctx.barrier()
    t0 = time.perf_counter()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        # N > 1: runs handed out longest first, lengths guessed by a 2000-row pilot fit on rank 0 (about a second,
        # inside the timed region); the result is the serial fit's whatever the order
        parallel.fit_sharded_restarts(g, X, Y, pilot_rows=2000 if ctx.world > 1 else None)
    ctx.torch.cuda.synchronize()
    dt = ctx.max_over_ranks(time.perf_counter() - t0)
    st = parallel.last_stats or {}
    per_rank = ctx.gather_obj({"rank": ctx.rank, "runs": st.get("items"), "lml_evals": int(g.n_lml_evals_),
                               "busy_s": st.get("busy_s"), "idle_s": st.get("idle_s")})
    out = {"full_fit_s_10_restarts": dt, "lml_evals_all_ranks": int(sum(p["lml_evals"] for p in per_rank)),
           "max_lml_evals_on_one_rank": int(max(p["lml_evals"] for p in per_rank)),
           "lml": float(g.log_marginal_likelihood_value_), "kernel": str(g.kernel_),
           "schedule": ("dynamic queue over ranks, claim order = longest first by a 2000-row pilot fit (timed)"
                        if ctx.world > 1 else "serial"), "scaling": "strong",
           "pilot_order": getattr(g, "pilot_order_", None), "pilot_seconds": getattr(g, "pilot_seconds_", None),
           "per_rank": per_rank}
    g._engine = None
    ctx.torch.cuda.empty_cache()
    return out


def config3_queries(m):
    """trajectories of 20 steps, random start + random-walk steps, seed 7"""
    rq = np.random.default_rng(7)
    return (rq.standard_normal((m // 20, 1, D)) + 0.05 * np.cumsum(rq.standard_normal((m // 20, 20, D)), 1)).reshape(-1, D)


def extra_predict_sharded(ctx, est, rows_per_gpu, peak):
    """Config 3's shape through parallel.predict_sharded: 81 920 x N query rows split into contiguous row blocks,
    every rank holding the factor; host arrays in, host arrays out, the gather of the outputs
    (all_gather_into_tensor) INSIDE the timed region."""
    from unmanned_aerial_vehicles_b200 import parallel

    Xq = config3_queries(rows_per_gpu * ctx.world)
    parallel.predict_sharded(est, Xq[: 2048 * ctx.world], return_std=True)         # warm-up: scratch, NCCL buffers
    parallel.predict_sharded(est, Xq[: 2048 * ctx.world], return_std=False)
    out = {"rows": int(Xq.shape[0]), "gather": "all_gather_into_tensor inside the timed region"}
    NP = est._engine.NP
    for name, std in (("mean_std", True), ("mean_only", False)):
        ctx.barrier()
        t0 = time.perf_counter()
        res = parallel.predict_sharded(est, Xq, return_std=std)
        ctx.torch.cuda.synchronize()
        dt = ctx.max_over_ranks(time.perf_counter() - t0)
        got = res[0] if std else res
        assert got.shape == (Xq.shape[0], T) and np.isfinite(got).all()
        out[f"{name}_pts_per_s"] = Xq.shape[0] / dt
        out[f"{name}_s"] = dt
        if std:   # the variance contraction: N^2 flops per query row, the bound of this branch
            out["mean_std_fp64_tflops_per_gpu"] = Xq.shape[0] * float(NP) ** 2 / dt / 1e12 / ctx.world
            out["mean_std_frac_fp64_peak_per_gpu"] = out["mean_std_fp64_tflops_per_gpu"] / peak
            out["algorithmic_hbm_bytes_per_s"] = 8.0 * (D + 2 * T) * Xq.shape[0] / dt
    return out


def extra_dist_chol(ctx, est, peak):
    """Config 5's path at this N: 2-D block-cyclic Cholesky over NCCL (DistGP.factor + solve + lml).  First at
    N_train = 20 000 against the single-GPU engine every rank holds (check_lml_rel), then at the size that matches
    the GPU count (100k / 140k / 200k for 2 / 4 / 8; total flops N^3/3 grow, per-GPU memory ~20-28 GB)."""
    from unmanned_aerial_vehicles_b200.dist_chol import DistGP, default_grid

    torch = ctx.torch
    th = theta_sequence(0, 1)[0]
    inv_ls, noise = np.exp(-th[:D]), float(np.exp(th[D]))
    grid = default_grid(ctx.world)
    if os.environ.get("GPK_DIST_GRID"):                       # e.g. GPK_DIST_GRID=1x8 (tools/dist_chol_run.py sweeps it)
        grid = tuple(int(v) for v in os.environ["GPK_DIST_GRID"].split("x"))
    out = {"grid": "x".join(map(str, grid)), "NB": 2048}
    sizes = [20000] + ([{2: 100000, 4: 140000, 8: 200000}.get(ctx.world, 20000 * ctx.world)] if ctx.world > 1 else [])
    for n in sizes:
        X, Y = make_data(n)
        gp = DistGP(X, Y, JITTER, NB=2048, grid=grid)
        ops = gp.ops
        ctx.barrier()
        ops.bcast_bytes = ops.bcast_calls = 0
        t0 = time.perf_counter()
        info = gp.factor(inv_ls, 1.0, noise)            # reads the pivot status back: the factorisation has finished
        torch.cuda.synchronize()
        t_f = ctx.max_over_ranks(time.perf_counter() - t0)
        t0 = time.perf_counter()
        gp.solve()
        lml = gp.lml()
        torch.cuda.synchronize()
        t_s = ctx.max_over_ranks(time.perf_counter() - t0)
        fl = float(n) ** 3 / 3
        rec = {"N": n, "info": int(info), "factor_s": t_f, "solve_lml_s": t_s, "lml": lml,
               "tflops_per_gpu_on_N3_over_3": fl / t_f / 1e12 / ctx.world,
               "frac_fp64_peak_per_gpu": fl / t_f / 1e12 / ctx.world / peak,
               "bcast_GB_this_rank": ops.bcast_bytes / 1e9, "bcast_calls_this_rank": ops.bcast_calls,
               "local_GB": gp.local_bytes() / 1e9}
        # the gradient at this size: distributed TRTRI + LAUUM (2 N^3 / 3) and the tile-wise contraction (dist_grad.py)
        ctx.barrier()
        b0 = ops.bcast_bytes
        t0 = time.perf_counter()
        gp.inverse()
        g_c, g_ls, g_noise = gp.gradient()
        torch.cuda.synchronize()
        t_g = ctx.max_over_ranks(time.perf_counter() - t0)
        rec.update({"inverse_plus_gradient_s": t_g, "objective_evaluation_with_gradient_s": t_f + t_s + t_g,
                    "evaluation_tflops_per_gpu_on_N3": float(n) ** 3 / (t_f + t_s + t_g) / 1e12 / ctx.world,
                    "evaluation_frac_fp64_peak_per_gpu": float(n) ** 3 / (t_f + t_s + t_g) / 1e12 / ctx.world / peak,
                    "inverse_bcast_GB_this_rank": (ops.bcast_bytes - b0) / 1e9,
                    "grad_log_noise": g_noise, "grad_log_length_scale_0": float(g_ls[0])})
        if n == 20000:
            ref, _, gls_ref, gn_ref = est._engine.lml(inv_ls, 1.0, noise, True)
            rec["check_lml_rel_vs_single_gpu_engine"] = abs(lml - ref) / abs(ref)
            rec["check_grad_maxrel_vs_single_gpu_engine"] = float(max(np.max(np.abs(g_ls - gls_ref)) / np.max(np.abs(gls_ref)),
                                                                     abs(g_noise - gn_ref) / abs(gn_ref)))
        out[f"N{n // 1000}k"] = rec
        gp.release_inverse()
        del gp, ops
        torch.cuda.empty_cache()
    return out


def b200_arm(args):
    import warnings

    ctx = Ctx()
    torch = ctx.torch
    from unmanned_aerial_vehicles_b200 import GaussianProcessRegressor, _lib

    lib = _lib.load()
    N = args.n
    X, Y = make_data(N)
    # the estimator a user holds (fixed theta: optimizer=None); its engine keeps the training set and the three
    # NP x NP factorisation buffers resident in HBM
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        est = GaussianProcessRegressor(kernel=ard_kernel(), alpha=JITTER, optimizer=None).fit(X, Y)
    eng = est._engine
    K, W = args.steps, max(args.warmup, 3)
    ths = theta_sequence(ctx.rank, W + 2 * K + 2)

    def dev_step(th):                                   # inputs resident in HBM
        return eng.lml(np.exp(-th[:D]), 1.0, float(np.exp(th[D])), True)

    Xp = torch.from_numpy(X).pin_memory()
    Yp = torch.from_numpy(Y).pin_memory()

    def e2e_step(th):                                   # the call a user makes, host buffers in, host results out
        eng.upload(Xp, Yp)                              # host -> device: this step's inputs (pinned)
        return est.log_marginal_likelihood(th, eval_gradient=True)      # device -> host: lml + gradient

    for th in ths[:W]:
        out = dev_step(th)
    assert np.isfinite(out[0]), "warm-up evaluation failed"
    # ---- timed region: device-resident inputs -------------------------------------------------------
    sampler = ClockSampler(ctx.local) if ctx.rank == 0 else None
    ctx.barrier()
    l0 = lib.gpk_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for th in ths[W:W + K]:
        dev_step(th)
    e1.record()
    ctx.barrier()
    launches = int(lib.gpk_launch_count() - l0)
    dt = ctx.max_over_ranks(e0.elapsed_time(e1) * 1e-3)
    # ---- e2e: GaussianProcessRegressor.log_marginal_likelihood, host buffers, copies inside the timed region ----
    e2e_step(ths[0])
    ctx.barrier()
    t0 = time.perf_counter()
    for th in ths[W + K:W + 2 * K]:
        lml_e2e, grad_e2e = e2e_step(th)
    torch.cuda.synchronize()
    dt_e2e = ctx.max_over_ranks(time.perf_counter() - t0)
    ctx.barrier()
    clocks = sampler.stop() if sampler else None
    assert np.isfinite(lml_e2e) and grad_e2e.shape == (D + 1,)

    line = None
    if ctx.rank == 0:
        # ---- roofline of the dominant kernel (gpk_gemm_nt_kernel): CUDA events around every launch --
        # (per-launch timing serialises the factorisation: the side streams of the Cholesky recursion are
        #  bypassed while it is on, so the share is taken against THIS step's own duration)
        lib.gpk_profile_begin()
        e0.record(); dev_step(ths[W + 2 * K]); e1.record()
        sec, fl, nl = C.c_double(0), C.c_double(0), C.c_int(0)
        lib.gpk_profile_end(C.byref(sec), C.byref(fl), C.byref(nl))
        torch.cuda.synchronize()
        step_s = e0.elapsed_time(e1) * 1e-3
        peak, lib_peak = fp64_peaks(ctx, lib)
        alg_flops = float(N) ** 3   # potrf N^3/3 + trtri N^3/3 + lauum N^3/3
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            traffic = json.load(open(tpath)).get("gpk_gemm_nt_kernel_dram_bytes_per_launch")
        roofline = {"bound": "tensor", "kernel": "gpk_gemm_nt_kernel (FP64 DMMA m8n8k4)",
                    "achieved": alg_flops / sec.value / 1e12, "peak": peak, "unit": "TFLOP/s",
                    "frac": alg_flops / sec.value / 1e12 / peak, "traffic": traffic,
                    "peak_source": "FP64 DMMA register-loop probe measured live in this run (MEASURED_PEAKS.json has "
                                   "no FP64 entry; the B200_PROFILING.md fallback has none either)",
                    "peak_cross_check_cublas_dgemm_8192_tflops": lib_peak,
                    "launches_per_step": nl.value, "gemm_seconds_per_step": sec.value,
                    "gemm_share_of_step": sec.value / step_s, "profiled_step_ms": step_s * 1e3,
                    "executed_tile_tflops": fl.value / sec.value / 1e12,
                    "algorithmic_flops_per_step": alg_flops,
                    "step_frac_of_peak": alg_flops / (dt / K) / 1e12 / peak}
        extra = {}
        torch.cuda.synchronize(); e0.record(); eng.factor(np.ones(D), 1.0, NOISE0); e1.record(); torch.cuda.synchronize()
        chol_s = e0.elapsed_time(e1) * 1e-3
        extra["cholesky_plus_inverse_s"] = chol_s
        extra["cholesky_plus_inverse_frac_fp64_peak"] = (2 * float(N) ** 3 / 3) / chol_s / 1e12 / peak
        line = {
            "metric": metric_name(N), "value": ctx.world * K / dt, "unit": "evals/s",
            "n_gpus": ctx.world, "steps": K, "warmup": W, "ms_per_step": dt / K * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": bench_config(N),
            "e2e": {"value": ctx.world * K / dt_e2e, "unit": "evals/s", "h2d_bytes_per_step": int(X.nbytes + Y.nbytes),
                    "d2h_bytes_per_step": int(eng.out_len * 8 + 4),
                    "api": "GaussianProcessRegressor.log_marginal_likelihood(theta, eval_gradient=True)"},
            "gpu_launches": launches, "roofline": roofline, "cpu_baseline": None, "clocks": clocks, "extra": extra,
        }
    # peak for the extras' fractions: rank 0's live probe, shared
    peak_all = ctx.gather_obj(line["roofline"]["peak"] if line else None)[0]

    # ---- extras, all ranks; a watchdog prints the headline line if one of them hangs -----------------
    skip = set(filter(None, args.skip.split(",")))
    done = threading.Event()

    def watchdog():
        if not done.wait(args.extras_deadline_s):
            if line is not None:
                line["extra"]["watchdog"] = f"extras exceeded {args.extras_deadline_s} s; line printed without the rest"
                print(json.dumps(line), flush=True)
            os._exit(0)
    threading.Thread(target=watchdog, daemon=True).start()

    def run_extra(name, fn):
        if name in skip:
            return
        try:
            res = fn()
        except Exception as e:  # noqa: BLE001 -- recorded in place; the headline line still prints
            res = {"error": f"{type(e).__name__}: {e}"[:400]}
        if line is not None:
            line["extra"][name] = res

    # the single-GPU predict figure of the metric (config 3's chunk: 4096 trajectories x 20 steps)
    # (the estimator's engine stays alive: the distributed extra checks itself against it; the riskiest extra -- the
    #  one with the most collectives -- runs last, so a hang there costs nothing that was already measured)
    run_extra("predict_sharded", lambda: extra_predict_sharded(ctx, est, args.predict_rows, peak_all))
    run_extra("full_fit", lambda: extra_full_fit(ctx, X, Y))
    run_extra("dist_chol", lambda: extra_dist_chol(ctx, est, peak_all))
    eng = None
    est._engine = None
    torch.cuda.empty_cache()
    if line is not None:
        ps = line["extra"].get("predict_sharded") or {}
        if "mean_std_pts_per_s" in ps:       # the metric's three quantities, side by side
            line["extra"]["predict_mean_std_pts_per_s"] = ps["mean_std_pts_per_s"]
            line["extra"]["predict_mean_only_pts_per_s"] = ps["mean_only_pts_per_s"]
        ff = line["extra"].get("full_fit") or {}
        if "full_fit_s_10_restarts" in ff:
            line["extra"]["gp_fit_s_N%dk_10_restarts" % (N // 1000)] = ff["full_fit_s_10_restarts"]
        if ctx.world == 1 and "cpu_baseline" not in skip and not args.no_cpu_baseline:
            try:
                line["cpu_baseline"], _ = cpu_baseline_block(N, 0.0, 1, False)     # ONE full evaluation, measured
            except Exception as e:  # noqa: BLE001
                line["cpu_baseline"] = {"error": f"{type(e).__name__}: {e}"[:400]}
        done.set()
        print(json.dumps(line), flush=True)
    done.set()
    if ctx.world > 1:
        ctx.dist.barrier()
        ctx.dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--size", "--n", dest="n", type=int, default=20000)  # torchrun's parser chokes on --n
    ap.add_argument("--targets", type=int, default=3, choices=[1, 2, 3])
    ap.add_argument("--predict-rows", type=int, default=81920, help="query rows per GPU of the predict extra")
    ap.add_argument("--skip", default="", help="comma list of extras to leave out: predict_sharded, dist_chol, "
                                               "full_fit, cpu_baseline")
    ap.add_argument("--extras-deadline-s", type=float, default=900.0)
    ap.add_argument("--cpu-budget-s", type=float, default=100.0,
                    help="reference arm: full evaluations are repeated while their total stays under this")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-stock-sklearn", action="store_true")
    args = ap.parse_args()
    global T
    T = args.targets
    if args.impl == "reference":
        reference_arm(args)
    else:
        b200_arm(args)


if __name__ == "__main__":
    main()
