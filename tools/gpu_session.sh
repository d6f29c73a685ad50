#!/bin/bash
# One parameterised GPU-session script (replaces the round-1 one-off tools/gpu_*.sh).
#   tools/gpu_session.sh <step> [<step> ...]       steps run in order; logs go to gpurun_out/
# steps:  suite | profile[:prec[:cfg]] | bench[:N[:extra flags]] | ref[:N] | timeline:N | fedavg:N[:nvls] |
#         quality:N[:epochs] | persist | sweep | ncu_iter | ncu_fedavg | parity
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29511"
for step in "$@"; do
  IFS=: read -r name a1 a2 <<< "$step"
  tag="gpurun_out/s_${step//[: ]/_}"
  case "$name" in
    suite)    timeout 600 python -m pytest tests -m gpu -x -q > $tag.log 2>&1; echo "== suite rc=$? $(tail -1 $tag.log)";;
    profile)  timeout 300 python tools/profile_step.py ${a1:-fp32x3} ${a2:-adult} > $tag.log 2>&1; echo "== profile ${a1:-fp32x3} ${a2:-adult} rc=$?"; head -1 $tag.log | cut -c1-500;;
    bench)    n=${a1:-1}
              if [ "$n" = 1 ]; then timeout 900 python bench.py --gpus 1 --steps 10 --warmup 3 $a2 > $tag.log 2> $tag.err
              else timeout 900 $TR --nproc-per-node $n bench.py --gpus $n --steps 10 --warmup 3 $a2 > $tag.log 2> $tag.err; fi
              echo "== bench N=$n rc=$?"; grep '^{' $tag.log | cut -c1-420;;
    ref)      n=${a1:-1}
              if [ "$n" = 1 ]; then timeout 1200 python bench.py --impl reference --gpus 1 --steps ${a2:-3} --warmup 1 > $tag.log 2> $tag.err
              else timeout 1200 $TR --nproc-per-node $n bench.py --impl reference --gpus $n --steps ${a2:-3} --warmup 1 > $tag.log 2> $tag.err; fi
              echo "== reference N=$n rc=$?"; grep '^{' $tag.log | cut -c1-420;;
    timeline) timeout 600 $TR --nproc-per-node $a1 tools/round_timeline.py 20 > $tag.log 2>&1; echo "== timeline N=$a1 rc=$?"; tail -3 $tag.log | cut -c1-600;;
    fedavg)   FEDAVG_MODES=${a2:-twoshot} timeout 600 $TR --nproc-per-node $a1 tools/fedavg_mgpu_check.py > $tag.log 2>&1; echo "== fedavg N=$a1 ${a2:-twoshot} rc=$?"; grep '^{' $tag.log | cut -c1-420 | tail -12;;
    quality)  timeout 900 $TR --nproc-per-node $a1 tools/spmd_quality.py ${a2:-6} > $tag.log 2>&1; echo "== quality N=$a1 rc=$?"; tail -2 $tag.log | cut -c1-900;;
    rpc)      timeout 1500 python tools/rpc_cli_bench.py ${a1:-2} ${a2:-12} > $tag.log 2>&1; echo "== rpc star clients=${a1:-2} rc=$?"; tail -1 $tag.log | cut -c1-700;;
    persist)  timeout 300 python tools/gemm_persist_check.py > $tag.log 2>&1; echo "== persist rc=$?"; tail -5 $tag.log;;
    sweep)    timeout 300 python tools/gemm_stage_sweep.py > $tag.log 2>&1; echo "== sweep rc=$?"; tail -6 $tag.log;;
    parity)   timeout 3000 python tools/parity_run.py ${a1:-20} adult ${a2:-3} > $tag.log 2>&1; echo "== parity rc=$?"; tail -8 $tag.log | cut -c1-300;;
    *)        echo "unknown step $step";;
  esac
done
