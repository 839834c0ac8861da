#!/usr/bin/env bash
# Round-2 second GPU session, TWO GPUs, one call:   gpurun --gpus 2 --timeout 1500 -- 'bash tools/gpu_session_r02b.sh'
# NCCL evidence for every path that communicates, at the cheapest GPU count; results under gpurun_out/r02b_*.
set -u
mkdir -p gpurun_out
step() {  # name, timeout seconds, command...
  local name=$1 lim=$2; shift 2
  local t0=$(date +%s)
  timeout "$lim" "$@" > "gpurun_out/r02b_${name}.out" 2> "gpurun_out/r02b_${name}.err"
  echo "{\"step\": \"${name}\", \"rc\": $?, \"s\": $(( $(date +%s) - t0 ))}" | tee -a gpurun_out/r02b_session.jsonl
}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
: > gpurun_out/r02b_session.jsonl
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/r02b_gpu.csv 2>&1
step pytest_gpu 1500 python -m pytest tests -m gpu -x -q
step ref_arm_threads 300 $TR --master-port 29521 bench.py --impl reference --gpus 2 --size 4000 --cpu-budget-s 1 --steps 3
step bench2 1200 $TR --master-port 29522 bench.py --gpus 2 --steps 5 --warmup 3
step dist40k_check 600 $TR --master-port 29523 tools/dist_chol_run.py --size 40000 --check --trace --predict-rows 2048
step dist100k_1x2 600 $TR --master-port 29524 tools/dist_chol_run.py --size 100000 --trace --grid 1x2
step dist100k_2x1 600 $TR --master-port 29525 tools/dist_chol_run.py --size 100000 --trace --grid 2x1
step dist100k_nb1024 600 $TR --master-port 29526 tools/dist_chol_run.py --size 100000 --trace --grid 1x2 --block 1024
step config4_small 600 $TR --master-port 29527 tools/config4_run.py --size 20000 --restarts 3 --outputs 2 --maxiter 2
step ncu_mean_fused 600 ncu --set full --clock-control none -k regex:gpk_predict_mean_fused -c 1 --csv --page raw \
  --log-file gpurun_out/r02b_mean_fused_ncu.csv python tools/mean_predict_run.py --rows 16384
cat gpurun_out/r02b_session.jsonl
tail -c 1500 gpurun_out/r02b_pytest_gpu.out
