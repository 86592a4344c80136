#!/bin/bash
# fused-block check: parity tests, per-layer profile lines of the fused launches, and an interleaved A/B bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python tools/layer_profile.py 256 > gpurun_out/layers.csv 2> gpurun_out/layers.err; grep -E "^#|c2\+c" gpurun_out/layers.csv | head -8
for r in 1 2; do
for v in X=1 ADVB200_NO_FUSEBLK=1; do
  env $v timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$v.json 2> gpurun_out/bench_$v.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_$v.json").read().strip().splitlines()[-1])
    print("$v", round(d["value"], 1), "img/s  e2e", round(d["e2e"]["value"], 1), d["clocks"]["sm_mhz"])
except Exception as e:
    print("$v failed", e); print(open("gpurun_out/bench_$v.err").read()[-1500:])
PY
done
done
