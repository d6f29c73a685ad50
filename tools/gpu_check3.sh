#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
echo "== pytest gpu"; timeout 900 python -m pytest tests -m gpu -q --timeout 400 > gpurun_out/t_gpu.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/t_gpu.log
echo "== profile tf32"; timeout 600 python tools/profile_step.py tf32 > gpurun_out/prof_tf32.log 2>&1; echo "rc=$?"; head -1 gpurun_out/prof_tf32.log
echo "== ncu list"; python tools/ncu_iter.py tf32 > gpurun_out/ncu_plain.log 2>&1 && ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_tf32.csv python tools/ncu_iter.py tf32 > gpurun_out/ncu_run.log 2>&1; tail -2 gpurun_out/ncu_run.log
echo "== bench"; timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_tf32.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench_tf32.log | cut -c1-700
