#!/bin/bash
# quick GPU check: parity tests + per-layer profile (+ optional bench)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 300 python tools/layer_profile.py 256 3 > gpurun_out/layers_256.csv 2> gpurun_out/layers.err; tail -2 gpurun_out/layers.err; tail -n 1 gpurun_out/layers_256.csv
if [ "${BENCH:-0}" = "1" ]; then timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; cat gpurun_out/bench.json; fi
