#!/bin/bash
# RPC star on GPUs (federator + 2 clients, symmetric-memory aggregation) + fedavg bandwidth check
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
W=$(mktemp -d); mkdir -p $W/data/raw
python - <<PY
import sys; sys.path.insert(0, "$PWD")
from fed_tgan_b200.data.schemas import get_schema, make_table
s = get_schema("adult", rows=12000); make_table(s, seed=1).to_csv("$W/data/raw/Adult_train.csv", index=False)
open("$W/cols.txt","w").write(",".join(s.columns)+"\n"+",".join(s.categorical)+"\n"+s.target)
PY
COLS=$(sed -n 1p $W/cols.txt); CATS=$(sed -n 2p $W/cols.txt); TGT=$(sed -n 3p $W/cols.txt)
cd $W
PYTHONPATH=$OLDPWD timeout 600 python -m dtds.distributed -world_size 3 -epoch 3 -port 27788 -datapath data/raw/Adult_train.csv \
  -selected_variables "$COLS" -categorical_list "$CATS" -nonnegative_list "" -target_column "$TGT" --vgm_backend vb --sample_rows 5000 > $OLDPWD/gpurun_out/rpc_gpu.log 2>&1
echo "rpc rc=$?"; ls Adult_result | head; wc -l Adult_result/Adult_synthesis_epoch_2.csv; cat timestamp_experiment.csv; grep -E "final aggregation|Error|error" $OLDPWD/gpurun_out/rpc_gpu.log | head -5
cd $OLDPWD
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 tools/fedavg_mgpu_check.py > gpurun_out/mgpu2.log 2>&1; echo "fedavg rc=$?"; grep -E "^\{|ALL OK|FAILED" gpurun_out/mgpu2.log | cut -c1-250 | head -5
