#!/bin/bash
# ncu --set full with source for selected conv launches of the second pass: IDX="6 27 103"
mkdir -p gpurun_out
for i in ${IDX:-6}; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_gemm -s $((116 + i)) -c 1 \
    -f -o gpurun_out/src_$i python tools/one_pass.py 256 2 > gpurun_out/ncu_src_$i.log 2>&1; echo "ncu $i rc=$?"
  ncu -i gpurun_out/src_$i.ncu-rep --page source --csv > gpurun_out/src_$i.csv 2>/dev/null
done
# launch list of one whole attack (all kernels)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 2000 -c 700 --csv \
  --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 3 --batch 256 --no-cpu-baseline \
  > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
ls -la gpurun_out/*.ncu-rep
