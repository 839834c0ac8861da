"""bench.py contract checks that need no GPU: the reference arm (the oracle port timed on the host cores) prints one
JSON line with the keys the driver reads, on the same metric/config as the B200 arm."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_contract_line():
    """launched the way torchrun launches it at N > 1 (OMP_NUM_THREADS=1 exported): the arm must lift the cap, time
    FULL evaluations (no extrapolation) and say how many threads and steps it really used"""
    env = dict(os.environ, OMP_NUM_THREADS="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "20",
                          "--warmup", "5", "--size", "2000", "--cpu-budget-s", "1.5"],
                         capture_output=True, text=True, timeout=300, cwd=ROOT, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "evals/s" and line["higher_is_better"] is True
    assert line["metric"] == "gp_fit_lml_grad_evals_per_s_N2k" and line["value"] > 0
    assert line["dtype"] == "f64" and line["vs_baseline"] is None and line["gpu_launches"] == 0
    cb = line["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] == os.cpu_count() and cb["value"] == line["value"]
    assert "no extrapolation" in cb["sample"] and "full N=2000" in cb["sample"]
    assert cb["blas_threads"] and all(p["threads"] == os.cpu_count() for p in cb["blas_threads"])   # cap lifted
    assert cb["launcher_omp_num_threads"] == str(os.cpu_count())
    assert 1 <= line["steps"] <= 20 and line["steps_requested"] == 20          # timed steps are capped by the budget
    assert abs(line["ms_per_step"] * 1e-3 - cb["eval_s"]) < 1e-9
    assert line["e2e"] == {"value": line["value"], "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "N=2000 D=9 T=3" in line["config"]["workload"] and set(line["config"]) == {"workload", "sharding", "l2"}
    stock = cb["stock_sklearn"]                              # the reference's own arithmetic beside the lean port
    assert stock["n"] == 2000 and stock["eval_s"] > 0 and stock["scaled_eval_s_extrapolated"] >= stock["eval_s"] * 0.999


def test_both_arms_name_the_same_config():
    sys.path.insert(0, ROOT)
    import bench
    assert bench.bench_config(20000) == bench.bench_config(20000)
    src = open(os.path.join(ROOT, "bench.py")).read()
    assert src.count('"config": bench_config(') == 2           # one per arm, nothing arm-specific inside


def test_non_zero_ranks_of_the_reference_arm_do_nothing():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                         capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""
