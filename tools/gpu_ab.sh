#!/bin/bash
# A/B pass: parity tests, per-layer profile, bench with tuning knobs on/off, ncu raw metrics of one pass (CSV only).
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python tools/layer_profile.py 256 3 > gpurun_out/layers_256.csv 2> gpurun_out/layers.err; tail -2 gpurun_out/layers.err; tail -n 1 gpurun_out/layers_256.csv
for v in ${VARIANTS:-base}; do
  case $v in
    base) envs="" ;;
    *) envs="$v" ;;
  esac
  echo "== variant: $envs"
  env $envs timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$v.json 2> gpurun_out/bench_$v.err; echo "bench rc=$?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_$v.json").read().strip().splitlines()[-1])
    print("$v", round(d["value"], 1), "img/s  e2e", round(d["e2e"]["value"], 1), "frac", round(d["roofline"]["frac"], 4), d["clocks"])
except Exception as e:
    print("$v failed", e)
PY
done
if [ "${NCURAW:-0}" = "1" ]; then
  timeout 1500 ncu --set full --clock-control none -k regex:conv_gemm -s 116 -c 116 \
    -f -o /tmp/conv_raw python tools/one_pass.py 256 2 > gpurun_out/ncu_raw.log 2>&1; echo "ncu raw rc=$?"
  ncu -i /tmp/conv_raw.ncu-rep --page raw --csv > gpurun_out/conv_raw.csv 2>/dev/null
  ls -la /tmp/conv_raw.ncu-rep gpurun_out/conv_raw.csv
fi
