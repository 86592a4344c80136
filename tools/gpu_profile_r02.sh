#!/bin/bash
# Round-2 profile pass for profiles/: the bench line, the ncu launch list of THE SAME bench command (timed steps only),
# ncu --set full of every conv launch of one forward + input-gradient pass, the per-layer table.
mkdir -p gpurun_out
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_final.json 2> gpurun_out/r2_bench_final.err; echo "bench rc=$?"; tail -c 600 gpurun_out/r2_bench_final.json
timeout 300 python tools/layer_profile.py 256 3 > gpurun_out/r2_layers_256.csv 2> gpurun_out/r2_layers.err; tail -n 1 gpurun_out/r2_layers_256.csv
ADVB200_BENCH_PROFILE_RANGE=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
  --log-file gpurun_out/r2_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_ncu_list.log 2>&1; echo "ncu list rc=$?"
wc -l gpurun_out/r2_launches_bench.csv
timeout 1500 ncu --set full --clock-control none --profile-from-start off -k 'regex:conv_gemm|fused_block' \
  -f -o /tmp/conv_raw python tools/one_pass.py 256 grad > gpurun_out/r2_ncu_raw.log 2>&1; echo "ncu raw rc=$?"
ncu -i /tmp/conv_raw.ncu-rep --page raw --csv > gpurun_out/r2_conv_raw.csv 2>/dev/null
ls -la gpurun_out/r2_conv_raw.csv gpurun_out/r2_launches_bench.csv
