#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for C in intrusion credit covertype; do
  echo "== $C"; timeout 900 python bench.py --config $C --steps 5 --warmup 3 --no-e2e > gpurun_out/bench_$C.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench_$C.log | cut -c1-330; grep -E "Error|error" gpurun_out/bench_$C.log | head -3
done
echo "== wide (200k rows)"; timeout 1200 python bench.py --config wide --rows 200000 --steps 3 --warmup 3 --no-e2e > gpurun_out/bench_wide.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench_wide.log | cut -c1-330; grep -E "Error|error" gpurun_out/bench_wide.log | head -3
