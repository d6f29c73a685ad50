#!/bin/bash
# ncu evidence of one GPU session (1 GPU; every ncu run follows a plain run of the same command that exited 0).
# The .ncu-rep files are large (an --import-source --set full capture of 24 kernels is ~80 MB) and gpurun brings back
# at most 64 MiB, so each report is turned into its raw / source CSV pages ON THE BOX and then deleted.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
page() {  # page <rep> <name>
  ncu -i gpurun_out/$1.ncu-rep --page raw --csv > gpurun_out/$2_raw.csv 2>/dev/null
  rm -f gpurun_out/$1.ncu-rep
}
for prec in fp32x3 tf32; do
  python tools/ncu_iter.py $prec 2 > gpurun_out/plain_iter_$prec.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
      --log-file gpurun_out/launches_$prec.csv python tools/ncu_iter.py $prec 2 > gpurun_out/ncu_l_$prec.log 2>&1
done
python tools/ncu_iter.py tf32 1 > gpurun_out/plain_iter1.log 2>&1 && \
ncu --set full --clock-control none --profile-from-start off -k regex:k_gemm_tc -c 26 -f -o gpurun_out/prof_gemm_tc \
    python tools/ncu_iter.py tf32 1 > gpurun_out/ncu_g1.log 2>&1 && page prof_gemm_tc prof_gemm_tc
python tools/ncu_iter.py fp32x3 1 > gpurun_out/plain_iter1x.log 2>&1 && \
ncu --set full --clock-control none --profile-from-start off -k regex:k_gemm_tc -c 26 -f -o gpurun_out/prof_gemm_x3 \
    python tools/ncu_iter.py fp32x3 1 > gpurun_out/ncu_g2.log 2>&1 && page prof_gemm_x3 prof_gemm_x3
# one GEMM with source correlation (kept small: 2 launches), source page extracted here
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:k_gemm_tc -s 2 -c 2 -f \
    -o gpurun_out/prof_gemm_src python tools/ncu_iter.py fp32x3 1 > gpurun_out/ncu_g3.log 2>&1 && \
  ncu -i gpurun_out/prof_gemm_src.ncu-rep --page source --csv > gpurun_out/prof_gemm_src_source.csv 2>/dev/null; rm -f gpurun_out/prof_gemm_src.ncu-rep
python tools/ncu_misc.py > gpurun_out/plain_misc.log 2>&1 && \
ncu --set full --clock-control none --profile-from-start off -c 40 -f -o gpurun_out/prof_misc \
    python tools/ncu_misc.py > gpurun_out/ncu_m.log 2>&1 && page prof_misc prof_misc
ncu --set full --clock-control none --profile-from-start off \
    -k regex:'k_(activate|bn_relu|adam|d_head|gp_penalty|prep|colsum|cond|vgm)' -c 24 -f -o gpurun_out/prof_small \
    python tools/ncu_iter.py fp32x3 1 > gpurun_out/ncu_s.log 2>&1 && page prof_small prof_small
ls -la gpurun_out/ | tail -20; du -sh gpurun_out; tail -2 gpurun_out/plain_misc.log
