#!/usr/bin/env bash
# First GPU session of the experimental int8 (tcgen05) contraction, in the order DESIGN.md section 8 gives.
# One gpurun call:   gpurun --timeout 1500 -- 'bash tools/i8_first_run.sh'
# Every step runs under its own timeout (a wrong mbarrier protocol hangs instead of failing) and writes to
# gpurun_out/; later steps still run when an earlier one fails, so one call answers as much as it can.
set -u
mkdir -p gpurun_out
step() {  # name, timeout seconds, command...
  local name=$1 lim=$2; shift 2
  timeout "$lim" "$@" > "gpurun_out/i8_${name}.out" 2> "gpurun_out/i8_${name}.err"
  echo "{\"step\": \"${name}\", \"rc\": $?}" | tee -a gpurun_out/i8_first_run.jsonl
}
: > gpurun_out/i8_first_run.jsonl
# 1. correctness ladder (smallest case first), then LAUUM- and predict-shaped timing; both stage widths
step check_kb64 300 python tools/i8_gemm_check.py
GPK_I8_KB=32 step check_kb32 300 python tools/i8_gemm_check.py
# 2. the gated tests: slices and products bit-identical to the NumPy statement, estimator-level modes 3 and 7
GPK_I8_EXPERIMENT=1 step pytest 900 python -m pytest tests/test_i8_gemm.py -m gpu -x -q
# 3. the default suite must not notice any of it
step pytest_default 900 python -m pytest tests -m gpu -x -q
# 4. the config-2 step: FP64 only, LAUUM, LAUUM + tail, everything (bench.py reads GPK_I8 through engine.py)
for mode in 0 1 5 7; do
  GPK_I8=$mode step "bench_mode${mode}" 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline
done
# 5. config-3 predict with the variance contraction on the int8 pipe
GPK_I8=2 step predict_mode2 600 python tools/predict_run.py --reps 4
cat gpurun_out/i8_first_run.jsonl
