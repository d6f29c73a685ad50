cd /root/repo; mkdir -p gpurun_out
python tools/ncu_misc.py > gpurun_out/plain_misc.log 2>&1 && \
ncu --set full --clock-control none --profile-from-start off -c 44 -f -o gpurun_out/prof_misc python tools/ncu_misc.py > gpurun_out/ncu_m.log 2>&1 && \
ncu -i gpurun_out/prof_misc.ncu-rep --page raw --csv > gpurun_out/prof_misc_raw.csv 2>/dev/null; rm -f gpurun_out/prof_misc.ncu-rep
tail -1 gpurun_out/plain_misc.log; ls -la gpurun_out/prof_misc_raw.csv
