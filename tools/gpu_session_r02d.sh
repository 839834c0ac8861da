#!/usr/bin/env bash
# Round-2 one-GPU validation after the distributed-gradient fix:   gpurun --timeout 900 -- 'bash tools/gpu_session_r02d.sh'
set -u
mkdir -p gpurun_out
step() {  # name, timeout seconds, command...
  local name=$1 lim=$2; shift 2
  local t0=$(date +%s)
  timeout "$lim" "$@" > "gpurun_out/r02d_${name}.out" 2> "gpurun_out/r02d_${name}.err"
  echo "{\"step\": \"${name}\", \"rc\": $?, \"s\": $(( $(date +%s) - t0 ))}" | tee -a gpurun_out/r02d_session.jsonl
}
: > gpurun_out/r02d_session.jsonl
step pytest_gpu 900 python -m pytest tests -m gpu -x -q
step per_output_latency 300 python tools/per_output_latency.py
step bench_quick 600 python bench.py --steps 5 --warmup 3 --skip full_fit,cpu_baseline
cat gpurun_out/r02d_session.jsonl
tail -c 1200 gpurun_out/r02d_pytest_gpu.out
