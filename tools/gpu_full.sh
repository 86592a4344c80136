#!/bin/bash
# Full GPU-box pass: parity tests, per-layer profile, bench, ncu launch list, ncu --set full of the conv launches of one pass.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python tools/layer_profile.py 256 3 > gpurun_out/layers_256.csv 2> gpurun_out/layers.err; tail -2 gpurun_out/layers.err; tail -n 1 gpurun_out/layers_256.csv
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json
if [ "${NCU:-1}" = "1" ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv \
    --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 3 --batch 256 --no-cpu-baseline \
    > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
fi
if [ "${NCUFULL:-1}" = "1" ]; then
  # one warm pass (skip the first pass' launches), capture every conv launch of the second pass
  timeout 1500 ncu --set full --clock-control none --import-source on -k regex:conv_gemm -s 130 -c 125 \
    -f -o gpurun_out/conv_full python tools/one_pass.py 256 2 > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
  ls -la gpurun_out/*.ncu-rep
fi
