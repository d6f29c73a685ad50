#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
echo "== covertype 2 clients skewed"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29641 bench.py --gpus 2 --config covertype --steps 4 --warmup 3 --no-e2e > gpurun_out/bench_cov2.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench_cov2.log | cut -c1-1500
echo "== wide 2 clients (200k rows)"; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29642 bench.py --gpus 2 --config wide --rows 200000 --steps 3 --warmup 3 --no-e2e > gpurun_out/bench_wide2.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench_wide2.log | cut -c1-1500
