#!/bin/bash
# A/B of the tcgen05 GEMM launch policy on the Adult and 1024-wide shapes (CUDA-event per-op profile).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
echo "== tests"; timeout 300 python -m pytest tests/test_native_gpu.py tests/test_engine_gpu.py tests/test_zz_no_alloc.py -x -q --timeout 200 > gpurun_out/ab_tests.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/ab_tests.log
run() {  # name, env...
  name=$1; shift
  for cfg in adult wide; do
    env "$@" timeout 120 python tools/profile_step.py tf32 $cfg > gpurun_out/ab_${name}_${cfg}.log 2>&1
    echo "-- $name $cfg rc=$?"; head -1 gpurun_out/ab_${name}_${cfg}.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print({k:d[k] for k in ('eager_sum_us','graph_iter_ms','generate_40000_ms')})"
    cp gpurun_out/profile_step_tf32_${cfg}.json gpurun_out/ab_${name}_${cfg}.json
  done
}
run new FTG_X=0
run promo256 FTG_TMA_L2_PROMO=256
run stages6 FTG_GEMM_STAGES=6
run stages6_promo256 FTG_GEMM_STAGES=6 FTG_TMA_L2_PROMO=256
