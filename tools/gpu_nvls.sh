#!/bin/bash
# 2-GPU (or N-GPU) check of the NVLS aggregation variant next to the P2P two-shot kernel.
N=${1:-2}
mkdir -p gpurun_out
FEDAVG_MODES=nvls,twoshot timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N \
  --master-addr 127.0.0.1 --master-port 29611 tools/fedavg_mgpu_check.py > gpurun_out/nvls_$N.log 2>&1
echo "exit $?" >> gpurun_out/nvls_$N.log
tail -30 gpurun_out/nvls_$N.log
