#!/bin/bash
# Profile pass for profiles/: launch list of a 2-step PGD attack, ncu --set full of every conv launch of one
# forward + input-gradient pass (raw CSV; the .ncu-rep stays on the box), full capture with source of the top
# non-conv kernels, per-layer profile, bench line.
mkdir -p gpurun_out
timeout 600 python bench.py --steps 8 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json
timeout 300 python tools/layer_profile.py 256 3 > gpurun_out/layers_256.csv 2> gpurun_out/layers.err; tail -n 1 gpurun_out/layers_256.csv
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
  --log-file gpurun_out/launches_pgd2.csv python tools/one_pass.py 256 pgd > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 1500 ncu --set full --clock-control none --profile-from-start off -k 'regex:conv_gemm|fused_block' \
  -f -o /tmp/conv_raw python tools/one_pass.py 256 grad > gpurun_out/ncu_raw.log 2>&1; echo "ncu raw rc=$?"
ncu -i /tmp/conv_raw.ncu-rep --page raw --csv > gpurun_out/conv_raw.csv 2>/dev/null
timeout 900 ncu --set full --clock-control none --profile-from-start off --import-source on \
  -k regex:'maxpool|prep_hcol|linf_step|fc_nt' -f -o gpurun_out/small_kernels python tools/one_pass.py 256 pgd > gpurun_out/ncu_small.log 2>&1; echo "ncu small rc=$?"
ncu -i gpurun_out/small_kernels.ncu-rep --page raw --csv > gpurun_out/small_raw.csv 2>/dev/null
ls -la gpurun_out/*.ncu-rep gpurun_out/*.csv
