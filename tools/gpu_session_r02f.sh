#!/usr/bin/env bash
# Round-2 FOUR-GPU (fallback when no 8-GPU slot is free) session, one call:   gpurun --gpus 4 --timeout 1500 -- 'bash tools/gpu_session_r02e.sh'
# Everything that needs 8 GPUs, most important first; every step under its own timeout; results gpurun_out/r02f_*.
set -u
mkdir -p gpurun_out
step() {  # name, timeout seconds, command...
  local name=$1 lim=$2; shift 2
  local t0=$(date +%s)
  timeout "$lim" "$@" > "gpurun_out/r02f_${name}.out" 2> "gpurun_out/r02f_${name}.err"
  echo "{\"step\": \"${name}\", \"rc\": $?, \"s\": $(( $(date +%s) - t0 ))}" | tee -a gpurun_out/r02f_session.jsonl
}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
: > gpurun_out/r02f_session.jsonl
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/r02f_gpu.csv 2>&1
# 0. the GPU suite (one GPU; the NCCL test takes two): the distributed-gradient fix and the double-buffered mean kernel
step pytest_gpu 900 python -m pytest tests -m gpu -x -q
# 1. the bench line at 8 GPUs: sharded 10-restart fit, row-sharded predict, N=200k factor + gradient
step bench4 900 $TR --master-port 29541 bench.py --gpus 4 --steps 5 --warmup 3
# 2. config 5 variants with the per-phase trace (grid / block size), the first with the distributed gradient + predict
step dist140k_2x2_grad 600 $TR --master-port 29542 tools/dist_chol_run.py --size 140000 --grid 2x2 --trace --grad --predict-rows 4096
step dist140k_1x4 300 $TR --master-port 29543 tools/dist_chol_run.py --size 140000 --grid 1x4 --trace
step dist140k_2x2_nb1024 300 $TR --master-port 29544 tools/dist_chol_run.py --size 140000 --grid 2x2 --block 1024 --trace
step dist140k_4x1 300 $TR --master-port 29545 tools/dist_chol_run.py --size 140000 --grid 4x1 --trace
# 3. config 4, truncated: one output x 8 runs x 3 L-BFGS-B iterations at N=50k (60 GB per model per GPU)
step config4_trunc 600 $TR --master-port 29546 tools/config4_run.py --size 50000 --restarts 3 --outputs 1 --maxiter 3
# 4. the same at two outputs, to see the buffers released between outputs
step config4_trunc2 600 $TR --master-port 29547 tools/config4_run.py --size 50000 --restarts 3 --outputs 2 --maxiter 1
# 5. an N=200k hyper-parameter FIT (truncated to one L-BFGS-B iteration: 2-3 objective evaluations with gradient)
step dist_fit140k 900 $TR --master-port 29548 tools/dist_fit_run.py --size 140000 --maxiter 1 --grid 2x2
cat gpurun_out/r02f_session.jsonl
