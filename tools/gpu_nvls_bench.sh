#!/bin/bash
# 2-GPU federated bench with the aggregation routed through the NVLS kernel.
mkdir -p gpurun_out
FEDTGAN_AGGREGATION=nvls timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 \
  --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/nvls_bench2.log 2>&1
echo "exit $?" >> gpurun_out/nvls_bench2.log
grep -E "metric|exit|rror|NVLS|warn" gpurun_out/nvls_bench2.log | tail -8
