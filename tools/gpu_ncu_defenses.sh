#!/bin/bash
# ncu --set full of one launch of each defense kernel at N=256 (run only after the plain run exits 0)
mkdir -p gpurun_out
timeout 90 python tools/defense_pass.py 256 > gpurun_out/defense_pass.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/defense_pass.log; exit 1; }
timeout 200 ncu --set full --clock-control none --profile-from-start off --import-source on -f -o gpurun_out/defense_kernels \
  python tools/defense_pass.py 256 > gpurun_out/ncu_defenses.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_defenses.log
ncu -i gpurun_out/defense_kernels.ncu-rep --page raw --csv > gpurun_out/defense_raw.csv 2>/dev/null
ls -la gpurun_out/defense_kernels.ncu-rep gpurun_out/defense_raw.csv
