#!/usr/bin/env bash
# Round-2 bounded multi-GPU bench:   gpurun --gpus N --timeout T -- 'bash tools/gpu_session_r02h.sh N'
# ONE bench.py line at N GPUs (sharded fit, row-sharded predict, block-cyclic factor + gradient at 140k / 200k), then --
# only if the box time used so far allows -- one traced distributed run.  Results gpurun_out/r02h_*.
set -u
N=${1:-4}
mkdir -p gpurun_out
T0=$(date +%s)
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 330 $TR --master-port 29561 bench.py --gpus $N --steps 5 --warmup 3 --extras-deadline-s 240 \
  > gpurun_out/r02h_bench${N}.out 2> gpurun_out/r02h_bench${N}.err
echo "{\"step\": \"bench${N}\", \"rc\": $?, \"s\": $(( $(date +%s) - T0 ))}" | tee gpurun_out/r02h_session${N}.jsonl
if [ $(( $(date +%s) - T0 )) -lt ${2:-200} ]; then
  if [ "$N" = "8" ]; then SZ=200000; GR=2x4; else SZ=140000; GR=2x2; fi
  timeout 200 $TR --master-port 29562 tools/dist_chol_run.py --size $SZ --grid $GR --trace --grad \
    > gpurun_out/r02h_dist${N}.out 2> gpurun_out/r02h_dist${N}.err
  echo "{\"step\": \"dist${N}\", \"rc\": $?, \"s\": $(( $(date +%s) - T0 ))}" | tee -a gpurun_out/r02h_session${N}.jsonl
fi
grep -c "^{" gpurun_out/r02h_bench${N}.out
