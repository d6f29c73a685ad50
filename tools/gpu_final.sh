#!/bin/bash
# Round-end validation on one GPU: full GPU test suite, smoke, headline bench, then (time permitting)
# the per-op profile of the 1024-wide configuration.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 420 python -m pytest tests -m gpu -x -q --timeout 200 > gpurun_out/final_pytest.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/final_pytest.log
echo "== smoke"; timeout 120 python __graft_entry__.py smoke > gpurun_out/final_smoke.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/final_smoke.log
echo "== bench"; timeout 200 python bench.py --steps 5 --warmup 3 > gpurun_out/final_bench.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/final_bench.log
echo "== done"
