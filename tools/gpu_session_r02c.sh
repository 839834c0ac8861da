#!/usr/bin/env bash
# Round-2 third GPU session, TWO GPUs:   gpurun --gpus 2 --timeout 1500 -- 'bash tools/gpu_session_r02c.sh'
# The distributed gradient (TRTRI + LAUUM over NCCL) and the double-buffered mean-only kernel; results gpurun_out/r02c_*.
set -u
mkdir -p gpurun_out
step() {  # name, timeout seconds, command...
  local name=$1 lim=$2; shift 2
  local t0=$(date +%s)
  timeout "$lim" "$@" > "gpurun_out/r02c_${name}.out" 2> "gpurun_out/r02c_${name}.err"
  echo "{\"step\": \"${name}\", \"rc\": $?, \"s\": $(( $(date +%s) - t0 ))}" | tee -a gpurun_out/r02c_session.jsonl
}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
: > gpurun_out/r02c_session.jsonl
step pytest_gpu 1500 python -m pytest tests -m gpu -x -q
step mean_predict 300 python tools/mean_predict_run.py
step dist40k_grad_check 600 $TR --master-port 29531 tools/dist_chol_run.py --size 40000 --check --grad --trace
step dist100k_grad_1x2 900 $TR --master-port 29532 tools/dist_chol_run.py --size 100000 --grad --trace --grid 1x2
step dist100k_grad_2x1 900 $TR --master-port 29533 tools/dist_chol_run.py --size 100000 --grad --trace --grid 2x1
step bench2 1200 $TR --master-port 29534 bench.py --gpus 2 --steps 5 --warmup 3 --skip full_fit
cat gpurun_out/r02c_session.jsonl
tail -c 1500 gpurun_out/r02c_pytest_gpu.out
