#!/bin/bash
# First GPU call of the next round: validate the opt-in persistent tcgen05 GEMM on the whole GPU suite and
# measure what it buys on the wide configuration, next to the default policy.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
echo "== default suite"; timeout 300 python -m pytest tests -m gpu -x -q --timeout 200 > gpurun_out/r2_suite_default.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/r2_suite_default.log
echo "== suite with FTG_GEMM_PERSISTENT=1"; FTG_GEMM_PERSISTENT=1 timeout 300 python -m pytest tests -m gpu -x -q --timeout 200 > gpurun_out/r2_suite_persist.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/r2_suite_persist.log
for mode in 0 1; do
  for cfg in adult wide; do
    FTG_GEMM_PERSISTENT=$mode timeout 120 python tools/profile_step.py tf32 $cfg > gpurun_out/r2_profile_p${mode}_${cfg}.log 2>&1
    echo "-- persistent=$mode $cfg rc=$?"; head -1 gpurun_out/r2_profile_p${mode}_${cfg}.log | cut -c1-400
  done
done
echo "== stage / split-K sweep"; timeout 120 python tools/gemm_stage_sweep.py > gpurun_out/r2_sweep.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/r2_sweep.log
echo "== bench persistent=1"; FTG_GEMM_PERSISTENT=1 timeout 200 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_persist.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/r2_bench_persist.log | cut -c1-300
