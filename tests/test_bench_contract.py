"""bench.py contract checks that need no GPU: the reference arm (the oracle port timed on the host cores) prints one
JSON line with the keys the driver reads, on the same metric/config as the B200 arm."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_contract_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0", "--size", "2000", "--cpu-sample-n", "400"],
                         capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "evals/s" and line["higher_is_better"] is True
    assert line["metric"] == "gp_fit_lml_grad_evals_per_s_N2k" and line["value"] > 0
    assert line["dtype"] == "f64" and line["vs_baseline"] is None and line["gpu_launches"] == 0
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"] == {"value": line["value"], "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "N=2000 D=9 T=3" in line["config"]["workload"]
    stock = line["cpu_baseline"]["stock_sklearn"]          # the reference's own arithmetic beside the lean port
    assert stock["n"] == 2000 and stock["eval_s"] > 0 and stock["scaled_eval_s"] >= stock["eval_s"] * 0.999


def test_non_zero_ranks_of_the_reference_arm_do_nothing():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                         capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""
