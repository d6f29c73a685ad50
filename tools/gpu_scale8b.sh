#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
echo "== N=8 rotate"; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29608 bench.py --gpus 8 --steps 16 --warmup 3 --no-e2e > gpurun_out/bench_n8_rot.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench_n8_rot.log | cut -c1-330
echo "== N=8 rank0"; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29609 bench.py --gpus 8 --steps 16 --warmup 3 --no-e2e --no-rotate > gpurun_out/bench_n8_r0.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench_n8_r0.log | cut -c1-330
