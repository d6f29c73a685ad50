#!/bin/bash
# 8-GPU session used for the round-2 numbers: headline bench exactly as the driver launches it (N = 8, 4), then optional extras
# selected by arguments: fedavg | configs
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29521"
timeout 300 $TR --nproc-per-node 8 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/s8_adult.log 2> gpurun_out/s8_adult.err; echo "== adult8 rc=$?"; grep '^{' gpurun_out/s8_adult.log | cut -c1-400
timeout 200 $TR --nproc-per-node 4 bench.py --gpus 4 --steps 20 --warmup 5 --no-alt > gpurun_out/s4_adult.log 2> gpurun_out/s4_adult.err; echo "== adult4 rc=$?"; grep '^{' gpurun_out/s4_adult.log | cut -c1-300
for extra in "$@"; do
  case "$extra" in
    fedavg)
      FEDAVG_MODES=twoshot,nvls timeout 300 $TR --nproc-per-node 8 tools/fedavg_mgpu_check.py > gpurun_out/s8_fedavg.log 2>&1; echo "== fedavg8 rc=$?"; grep '^{' gpurun_out/s8_fedavg.log | cut -c1-400
      FEDAVG_MODES=nvls FEDAVG_UNIFORM=1 timeout 300 $TR --nproc-per-node 8 tools/fedavg_mgpu_check.py > gpurun_out/s8_fedavg_u.log 2>&1; echo "== fedavg8 uniform rc=$?"; grep '^{' gpurun_out/s8_fedavg_u.log | cut -c1-400;;
    configs)
      for c in covertype credit; do
        timeout 400 $TR --nproc-per-node 8 bench.py --gpus 8 --config $c --steps 5 --warmup 3 --no-alt --no-e2e > gpurun_out/s8_$c.log 2> gpurun_out/s8_$c.err; echo "== $c 8 rc=$?"; grep '^{' gpurun_out/s8_$c.log | cut -c1-300
      done
      timeout 300 $TR --nproc-per-node 4 bench.py --gpus 4 --config credit --steps 5 --warmup 3 --no-alt --no-e2e > gpurun_out/s4_credit.log 2> gpurun_out/s4_credit.err; echo "== credit 4 rc=$?"; grep '^{' gpurun_out/s4_credit.log | cut -c1-300
      timeout 500 $TR --nproc-per-node 8 bench.py --gpus 8 --config wide --rows 200000 --steps 5 --warmup 3 --no-alt --no-e2e > gpurun_out/s8_wide.log 2> gpurun_out/s8_wide.err; echo "== wide 8 rc=$?"; grep '^{' gpurun_out/s8_wide.log | cut -c1-300
      FEDTGAN_AGGREGATION=nvls timeout 500 $TR --nproc-per-node 8 bench.py --gpus 8 --config wide --rows 200000 --steps 5 --warmup 3 --no-alt --no-e2e > gpurun_out/s8_wide_nvls.log 2> gpurun_out/s8_wide_nvls.err; echo "== wide 8 nvls rc=$?"; grep '^{' gpurun_out/s8_wide_nvls.log | cut -c1-300;;
  esac
done
