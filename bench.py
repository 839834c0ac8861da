#!/usr/bin/env python
"""bench.py -- BASELINE.json config 2 ("Synthetic N=20k, D=9 exact GP fit: Gram + Cholesky + LML/grad").

A step = ONE log-marginal-likelihood + gradient evaluation (the unit L-BFGS-B repeats ~30x per restart):
Gram assembly -> blocked Cholesky fused with the triangular inverse -> alpha -> K^-1 -> gradient contraction.
`value` = evaluations/s over all ranks (restarts are sharded across GPUs: each rank evaluates its own theta
sequence on a replicated training set, no data-path collective -> weak scaling).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--full-fit]
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

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


def metric_name(n):
    return "gp_fit_lml_grad_evals_per_s_N%dk" % (n // 1000)


def theta_sequence(rank, count):
    """log-parameters [log l_1..l_9, log noise] an optimiser run would visit (seeded per rank/restart)"""
    rng = np.random.default_rng(1000 + rank)
    base = np.concatenate([np.zeros(D), [np.log(NOISE0)]])
    return [base + 0.2 * rng.standard_normal(D + 1) for _ in range(count)]


# ------------------------------------------------------------------------------------------------
def cpu_eval_sample(n_sample, n_full, reps):
    """Oracle (sklearn restatement) LML+grad at n_sample on all host cores, scaled to n_full:
    cubic part (dpotrf + N-rhs dpotrs) by r^3, the rest (kernel, gradient passes) by r^2."""
    from oracle import gp_oracle as orc
    from scipy.linalg import cho_solve, cholesky

    X, Y = make_data(n_sample)
    ths = theta_sequence(0, reps)
    r = n_full / n_sample
    est = []
    for th in ths:
        ls, noise = np.exp(th[:D]), float(np.exp(th[D]))
        t0 = time.perf_counter()
        orc.lml_and_grad(X, Y, ls, 1.0, noise, JITTER)
        t_all = time.perf_counter() - t0
        K = orc.gram(X, ls, 1.0, noise + JITTER)
        t0 = time.perf_counter()
        L = cholesky(K, lower=True, check_finite=False)
        cho_solve((L, True), np.eye(n_sample), check_finite=False)
        t_cubic = time.perf_counter() - t0
        t_cubic = min(t_cubic, t_all)
        est.append((t_all, t_cubic, t_cubic * r ** 3 + (t_all - t_cubic) * r ** 2))
    t_all, t_cubic, t_scaled = map(float, np.median(np.array(est), axis=0))
    return {"sample_eval_s": t_all, "sample_cubic_s": t_cubic, "scaled_eval_s": t_scaled}


def stock_sklearn_sample(n_sample, n_full):
    """The reference's own arithmetic, unrestated: scikit-learn's GaussianProcessRegressor.log_marginal_likelihood(theta,
    eval_gradient=True) (_gpr.py:541-656; it materialises the (N, N, n_theta) gradient tensor, so it is run at a size
    that fits: 10 x 8 N^2 bytes) -- one evaluation, scaled to n_full like the port's sample."""
    import warnings

    from scipy.linalg import cho_solve, cholesky
    from sklearn.gaussian_process import GaussianProcessRegressor as SkGPR
    from sklearn.gaussian_process.kernels import RBF, ConstantKernel as CK, WhiteKernel

    X, Y = make_data(n_sample)
    k = CK(1.0, constant_value_bounds="fixed") * RBF([1.0] * D, (0.1, 10.0)) + WhiteKernel(NOISE0, (1e-5, 10.0))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        g = SkGPR(kernel=k, alpha=JITTER, optimizer=None).fit(X, Y)
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
    return {"n": n_sample, "eval_s": t_all, "cubic_s": t_cubic, "scaled_eval_s": t_cubic * r ** 3 + (t_all - t_cubic) * r ** 2,
            "what": "sklearn GaussianProcessRegressor.log_marginal_likelihood(theta, eval_gradient=True), one evaluation"}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_s = args.cpu_sample_n or 5000
    if args.cpu_full:
        from oracle import gp_oracle as orc
        X, Y = make_data(args.n)
        ts = []
        for th in theta_sequence(0, max(1, args.steps)):
            t0 = time.perf_counter()
            orc.lml_and_grad(X, Y, np.exp(th[:D]), 1.0, float(np.exp(th[D])), JITTER)
            ts.append(time.perf_counter() - t0)
        eval_s, sample = float(np.median(ts)), f"full N={args.n} LML+grad, {len(ts)} evaluations"
    else:
        for _ in range(max(0, min(args.warmup, 1))):
            cpu_eval_sample(n_s, args.n, 1)
        r = cpu_eval_sample(n_s, args.n, max(1, args.steps))
        eval_s = r["scaled_eval_s"]
        sample = (f"LML+grad at N={n_s} ({r['sample_eval_s']:.3f} s, cubic part {r['sample_cubic_s']:.3f} s), "
                  f"scaled to N={args.n}: cubic x r^3 + rest x r^2")
    value = 1.0 / eval_s
    cores = os.cpu_count()
    stock = None
    if not args.cpu_full and not args.no_stock_sklearn:
        stock = stock_sklearn_sample(min(3000, args.n), args.n)     # reported beside the lean port, which is faster
    line = {
        "impl": "reference", "metric": metric_name(args.n), "value": value, "unit": "evals/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": eval_s * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args.n),
                   "engine": "oracle port of sklearn 1.9.0 _gpr.py:583-656 on SciPy LAPACK/OpenBLAS"},
        "cpu_baseline": {"value": value, "unit": "evals/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if stock is not None:
        line["cpu_baseline"]["stock_sklearn"] = stock
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


def b200_arm(args):
    import torch
    import torch.distributed as dist

    from unmanned_aerial_vehicles_b200 import _lib
    from unmanned_aerial_vehicles_b200.engine import DeviceGP

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = _lib.load()
    N = args.n
    X, Y = make_data(N)
    eng = DeviceGP(X, Y, JITTER)
    K, W = args.steps, max(args.warmup, 3)
    ths = theta_sequence(rank, W + 2 * K + 2)

    def dev_step(th):
        return eng.lml(np.exp(-th[:D]), 1.0, float(np.exp(th[D])), True)

    Xp = torch.from_numpy(X).pin_memory()
    Yp = torch.from_numpy(Y).pin_memory()

    def e2e_step(th):
        eng.X_d.copy_(Xp, non_blocking=True)          # host -> device: this step's inputs
        eng.y_d[:N].copy_(Yp, non_blocking=True)
        return eng.lml(np.exp(-th[:D]), 1.0, float(np.exp(th[D])), True)   # device -> host: lml + gradient

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for th in ths[:W]:
        out = dev_step(th)
    assert np.isfinite(out[0]), "warm-up evaluation failed"
    # ---- timed region: device-resident inputs -------------------------------------------------------
    sampler = ClockSampler(local) if rank == 0 else None
    barrier()
    l0 = lib.gpk_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for th in ths[W:W + K]:
        dev_step(th)
    e1.record()
    barrier()
    launches = int(lib.gpk_launch_count() - l0)
    dt = torch.tensor([e0.elapsed_time(e1) * 1e-3], dtype=torch.float64, device="cuda")
    # ---- e2e: host buffers, copies inside the timed region ------------------------------------------
    barrier()
    t0 = time.perf_counter()
    for th in ths[W + K:W + 2 * K]:
        e2e_step(th)
    torch.cuda.synchronize()
    dt_e2e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    barrier()
    clocks = sampler.stop() if sampler else None
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        dist.all_reduce(dt_e2e, op=dist.ReduceOp.MAX)
    dt, dt_e2e = float(dt[0]), float(dt_e2e[0])

    if rank == 0:
        # ---- roofline of the dominant kernel (gpk_gemm_nt_kernel): CUDA events around every launch --
        # (per-launch timing serialises the factorisation: the side streams of the Cholesky recursion are
        #  bypassed while it is on, so the share is taken against THIS step's own duration)
        lib.gpk_profile_begin()
        e0.record(); dev_step(ths[W + 2 * K]); e1.record()
        sec, fl, nl = C.c_double(0), C.c_double(0), C.c_int(0)
        lib.gpk_profile_end(C.byref(sec), C.byref(fl), C.byref(nl))
        torch.cuda.synchronize()
        step_s = e0.elapsed_time(e1) * 1e-3
        sink = torch.zeros(8, dtype=torch.float64, device="cuda")
        pf = C.c_double(0)
        best = 0.0
        for _ in range(3):
            e0.record()
            lib.gpk_dmma_probe(296, 20000, C.c_void_p(sink.data_ptr()), C.byref(pf),
                               C.c_void_p(torch.cuda.current_stream().cuda_stream))
            e1.record(); torch.cuda.synchronize()
            best = max(best, pf.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
        alg_flops = float(N) ** 3   # potrf N^3/3 + trtri N^3/3 + lauum N^3/3
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            traffic = json.load(open(tpath)).get("gpk_gemm_nt_kernel_dram_bytes_per_launch")
        roofline = {"bound": "tensor", "kernel": "gpk_gemm_nt_kernel (FP64 DMMA m8n8k4)",
                    "achieved": alg_flops / sec.value / 1e12, "peak": best, "unit": "TFLOP/s",
                    "frac": alg_flops / sec.value / 1e12 / best, "traffic": traffic,
                    "peak_source": "FP64 DMMA register-loop probe measured live in this run (MEASURED_PEAKS.json has "
                                   "no FP64 entry; the B200_PROFILING.md fallback has none either)",
                    "launches_per_step": nl.value, "gemm_seconds_per_step": sec.value,
                    "gemm_share_of_step": sec.value / step_s, "profiled_step_ms": step_s * 1e3, "executed_tile_tflops": fl.value / sec.value / 1e12,
                    "algorithmic_flops_per_step": alg_flops}
        # ---- secondary figures of the metric ---------------------------------------------------------
        extra = {}
        inv = np.ones(D)
        torch.cuda.synchronize(); e0.record(); info = eng.factor(inv, 1.0, NOISE0); e1.record(); torch.cuda.synchronize()
        chol_s = e0.elapsed_time(e1) * 1e-3
        extra["cholesky_plus_inverse_s"] = chol_s
        extra["cholesky_plus_inverse_frac_fp64_peak"] = (2 * float(N) ** 3 / 3) / chol_s / 1e12 / best
        Mq = args.predict_rows
        rq = np.random.default_rng(7)
        Xq = (rq.standard_normal((Mq // 20, 1, D)) + 0.05 * np.cumsum(rq.standard_normal((Mq // 20, 20, D)), 1)).reshape(-1, D)
        eng.predict(Xq[:2048], 0.0, 1.0, True)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        eng.predict(Xq, 0.0, 1.0, True)
        torch.cuda.synchronize(); tp = time.perf_counter() - t0
        extra["predict_mean_std_pts_per_s"] = Xq.shape[0] / tp
        extra["predict_sample_rows"] = int(Xq.shape[0])
        extra["predict_fp64_tflops"] = Xq.shape[0] * float(eng.NP) ** 2 / tp / 1e12
        if args.full_fit:
            extra.update(full_fit(X, Y))
        cpu = None
        if not args.no_cpu_baseline:
            n_s = args.cpu_sample_n or 4000
            r = cpu_eval_sample(n_s, N, 2)
            cpu = {"value": 1.0 / r["scaled_eval_s"], "unit": "evals/s", "cores": os.cpu_count(), "kind": "port",
                   "sample": (f"oracle LML+grad at N={n_s}: {r['sample_eval_s']:.3f} s (cubic part "
                              f"{r['sample_cubic_s']:.3f} s), scaled to N={N}: cubic x r^3 + rest x r^2")}
        h2d = int(X.nbytes + Y.nbytes)
        d2h = int(eng.out_len * 8 + 4)
        line = {
            "metric": metric_name(N), "value": world * K / dt, "unit": "evals/s",
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": dt / K * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(N),
                       "sharding": "restarts x ranks, training set replicated, no collective in the data path",
                       "l2": f"inputs (3 x {8e-9 * eng.NP ** 2:.1f} GB factorisation buffers) larger than L2; no flush needed"},
            "e2e": {"value": world * K / dt_e2e, "unit": "evals/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h},
            "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks, "extra": extra,
        }
        if getattr(eng, "i8_mode", 0):   # opt-in experiment: a line measured through it says so (never the default)
            line["dtype"] = "f64 with int8-sliced emulation on parts of the path"
            line["config"]["int8_emulation"] = {"GPK_I8": eng.i8_mode, "slices": int(os.environ.get("GPK_I8_SLICES", "7")),
                                                "note": "bit 1 LAUUM, bit 2 predict variance, bit 4 right-most-path "
                                                        "inverse products; roofline block still times the DMMA kernel only"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def full_fit(X, Y):
    """BASELINE config 2 proper: 10 L-BFGS-B runs (n_restarts_optimizer=9), random_state=0."""
    import warnings

    from sklearn.gaussian_process.kernels import RBF, ConstantKernel as CK, WhiteKernel

    from unmanned_aerial_vehicles_b200 import GaussianProcessRegressor
    from unmanned_aerial_vehicles_b200 import _lib
    k = CK(1.0, constant_value_bounds="fixed") * RBF([1.0] * D, (0.1, 10.0)) + WhiteKernel(NOISE0, (1e-5, 10.0))
    g = GaussianProcessRegressor(kernel=k, alpha=JITTER, n_restarts_optimizer=9, random_state=0)
    calls = {"n": 0}
    orig = g.log_marginal_likelihood

    def counted(*a, **kw):
        calls["n"] += 1
        return orig(*a, **kw)
    g.log_marginal_likelihood = counted
    t0 = time.perf_counter()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        g.fit(X, Y)
    return {"full_fit_s_10_restarts": time.perf_counter() - t0, "full_fit_lml_evals": calls["n"],
            "full_fit_lml": float(g.log_marginal_likelihood_value_), "full_fit_kernel": str(g.kernel_)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--size", "--n", dest="n", type=int, default=20000)  # torchrun's parser chokes on --n
    ap.add_argument("--targets", type=int, default=3, choices=[1, 2, 3])
    ap.add_argument("--cpu-sample-n", type=int, default=0)
    ap.add_argument("--cpu-full", action="store_true")
    ap.add_argument("--predict-rows", type=int, default=81920)
    ap.add_argument("--full-fit", action="store_true")
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
