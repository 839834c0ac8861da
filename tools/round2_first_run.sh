#!/usr/bin/env bash
# First GPU session of round 2 in ONE gpurun call (everything written after round 1's GPU minutes were spent):
#   gpurun --timeout 2400 -- 'bash tools/round2_first_run.sh'
# Every step runs under its own timeout and writes to gpurun_out/r2_*; later steps still run when one fails.
set -u
mkdir -p gpurun_out
step() {  # name, timeout seconds, command...
  local name=$1 lim=$2; shift 2
  timeout "$lim" "$@" > "gpurun_out/r2_${name}.out" 2> "gpurun_out/r2_${name}.err"
  echo "{\"step\": \"${name}\", \"rc\": $?}" | tee -a gpurun_out/r2_first_run.jsonl
}
: > gpurun_out/r2_first_run.jsonl
# 1. the default suite (includes tests/test_zz_late_additions_gpu.py: stacked predict, GPTrainer mirror, offline flow)
step pytest_gpu 1200 python -m pytest tests -m gpu -x -q
# 2. the bench line and its ncu launch list (shares must agree with profiles/r01g_launch_summary.csv)
step bench 600 python bench.py
step ncu_launches 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv \
  --log-file gpurun_out/r2_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline
# 3. per-output loaders: member loop vs gpk_predict_stacked; the control tick of the single multi-output model
step per_output_latency 300 python tools/per_output_latency.py
step tick_latency 300 python tools/tick_latency.py
# 4. GEMM raster group height: time, then DRAM bytes of the same launches (one process per G; env is read once)
for G in 4 8 16 24; do
  GPK_GEMM_RASTER=$G step "raster_time_G${G}" 300 python tools/gemm_raster_sweep.py
  GPK_GEMM_RASTER=$G step "raster_ncu_G${G}" 600 ncu --set full --clock-control none -k regex:gpk_gemm_nt -c 1 \
    --csv --page raw --log-file "gpurun_out/r2_raster_ncu_G${G}.csv" python tools/gemm_raster_sweep.py --small
done
# 5. the int8 (tcgen05) experiment's own first session
step i8 1500 bash tools/i8_first_run.sh
cat gpurun_out/r2_first_run.jsonl
