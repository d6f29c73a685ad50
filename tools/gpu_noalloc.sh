#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
echo "== no-alloc alone (fresh process)"; timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu --timeout 120 > gpurun_out/t_noalloc_alone.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/t_noalloc_alone.log
echo "== whole gpu suite (no-alloc collected last)"; timeout 900 python -m pytest tests -m gpu -x -q --timeout 400 > gpurun_out/t_gpu.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/t_gpu.log
echo "== reference arm N=1"; timeout 1500 python bench.py --impl reference --steps 3 --warmup 2 > gpurun_out/bench_ref1.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench_ref1.log | cut -c1-1200
