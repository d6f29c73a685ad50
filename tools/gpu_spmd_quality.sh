#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
timeout 280 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29641 \
  tools/spmd_quality.py 6 > gpurun_out/spmd_quality_$N.log 2>&1
echo "exit $?" >> gpurun_out/spmd_quality_$N.log
grep -E "QUALITY|exit|rror|avg_jsd" gpurun_out/spmd_quality_$N.log | tail -6
