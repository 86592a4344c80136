#!/bin/bash
# ncu --set full with source for selected conv launches of one warm pass: IDX="5 97" (0-based among conv launches)
mkdir -p gpurun_out
for i in ${IDX:-5}; do
  timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:${KREGEX:-conv_gemm} -s $i -c 1 \
    -f -o gpurun_out/src_${TAG:-}$i python tools/one_pass.py 256 grad > gpurun_out/ncu_src_${TAG:-}$i.log 2>&1; echo "ncu $i rc=$?"
  ncu -i gpurun_out/src_${TAG:-}$i.ncu-rep --page source --csv > gpurun_out/src_${TAG:-}$i.csv 2>/dev/null
  ncu -i gpurun_out/src_${TAG:-}$i.ncu-rep --page raw --csv > gpurun_out/srcraw_${TAG:-}$i.csv 2>/dev/null
done
