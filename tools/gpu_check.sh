#!/bin/bash
# Staged GPU validation: SIMT/elementwise first, tcgen05 in its own process (a trap there must
# not take the other results down), then the engine tests and a short bench.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
echo "== smoke fp32"; FEDTGAN_PRECISION=fp32 timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke_fp32.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/smoke_fp32.log
echo "== native (no tcgen05)"; timeout 600 python -m pytest tests/test_native_gpu.py -q -k "not tf32 and not splitk" --timeout 180 > gpurun_out/t_native_simt.log 2>&1; echo "rc=$?"; tail -25 gpurun_out/t_native_simt.log
echo "== native tcgen05"; timeout 600 python -m pytest tests/test_native_gpu.py -q -k "tf32 or splitk" --timeout 180 > gpurun_out/t_native_tc.log 2>&1; echo "rc=$?"; tail -40 gpurun_out/t_native_tc.log
echo "== engine"; timeout 900 python -m pytest tests/test_engine_gpu.py -q --timeout 400 > gpurun_out/t_engine.log 2>&1; echo "rc=$?"; tail -40 gpurun_out/t_engine.log
echo "== bench fp32"; timeout 600 python bench.py --steps 3 --warmup 3 --precision fp32 > gpurun_out/bench_fp32.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/bench_fp32.log
echo "== bench tf32"; timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_tf32.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/bench_tf32.log
