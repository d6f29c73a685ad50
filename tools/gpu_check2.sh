#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
echo "== native tcgen05"; timeout 600 python -m pytest tests/test_native_gpu.py -q -k "tf32 or splitk or philox" --timeout 180 > gpurun_out/t_native_tc.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/t_native_tc.log
echo "== engine"; timeout 900 python -m pytest tests/test_engine_gpu.py -q --timeout 400 > gpurun_out/t_engine.log 2>&1; echo "rc=$?"; tail -25 gpurun_out/t_engine.log
echo "== profile tf32"; timeout 600 python tools/profile_step.py tf32 > gpurun_out/prof_tf32.log 2>&1; echo "rc=$?"; tail -45 gpurun_out/prof_tf32.log
echo "== profile fp32"; timeout 600 python tools/profile_step.py fp32 > gpurun_out/prof_fp32.log 2>&1; echo "rc=$?"; head -3 gpurun_out/prof_fp32.log | cut -c1-600
