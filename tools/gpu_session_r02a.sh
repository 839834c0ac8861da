#!/usr/bin/env bash
# Round-2 first GPU session, ONE gpurun call (1 GPU):   gpurun --timeout 1500 -- 'bash tools/gpu_session_r02a.sh'
# Every step runs under its own timeout and writes to gpurun_out/r02a_*; later steps still run when one fails.
set -u
mkdir -p gpurun_out
step() {  # name, timeout seconds, command...
  local name=$1 lim=$2; shift 2
  local t0=$(date +%s)
  timeout "$lim" "$@" > "gpurun_out/r02a_${name}.out" 2> "gpurun_out/r02a_${name}.err"
  echo "{\"step\": \"${name}\", \"rc\": $?, \"s\": $(( $(date +%s) - t0 ))}" | tee -a gpurun_out/r02a_session.jsonl
}
: > gpurun_out/r02a_session.jsonl
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/r02a_gpu.csv 2>&1
nproc > gpurun_out/r02a_nproc.txt
step smoke 300 python __graft_entry__.py smoke
step pytest_gpu 1500 python -m pytest tests -m gpu -x -q
step bench 900 python bench.py --steps 10 --warmup 3
step mean_predict 300 python tools/mean_predict_run.py
step gemm_breakdown 300 python tools/gemm_breakdown.py
step ncu_launches 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv \
  --log-file gpurun_out/r02a_launches.csv python bench.py --steps 1 --warmup 1 --skip full_fit,predict_sharded,dist_chol,cpu_baseline
step ncu_mean_fused 600 ncu --set full --clock-control none -k regex:gpk_predict_mean_fused -c 1 --csv --page raw \
  --log-file gpurun_out/r02a_mean_fused_ncu.csv python tools/mean_predict_run.py --rows 16384
step tick_latency 300 python tools/tick_latency.py
cat gpurun_out/r02a_session.jsonl
tail -c 3000 gpurun_out/r02a_pytest_gpu.out
