#!/bin/bash
# parity tests + bench (+ optional launch list of one forward/backward pass)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
for v in ${VARIANTS:-base}; do
  case $v in base) envs="" ;; *) envs="$v" ;; esac
  env $envs timeout 600 python bench.py --steps ${STEPS:-4} --warmup 3 --no-cpu-baseline > gpurun_out/bench_$v.json 2> gpurun_out/bench_$v.err; echo "bench rc=$?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_$v.json").read().strip().splitlines()[-1])
    print("$v", round(d["value"], 1), "img/s  e2e", round(d["e2e"]["value"], 1), "frac", round(d["roofline"]["frac"], 4), "whole", round(d["roofline"]["whole_step_frac"], 4), d["clocks"])
except Exception as e:
    print("$v failed", e); print(open("gpurun_out/bench_$v.err").read()[-2000:])
PY
done
if [ "${LIST:-0}" = "1" ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/pass_launches.csv \
    python tools/one_pass.py 256 grad > gpurun_out/ncu_pass.log 2>&1; echo "ncu list rc=$?"
  python - <<'PY'
import csv, collections
rows = list(csv.reader(open("gpurun_out/pass_launches.csv")))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
h = rows[hi]; kn = h.index("Kernel Name"); mv = h.index("Metric Value")
seq = [(r[kn].split("(")[0][-48:], float(r[mv].replace(",", ""))) for r in rows[hi + 1:] if len(r) > mv]
agg = collections.OrderedDict()
for k, v in seq:
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
print("one pass: kernels", len(seq), "total us", sum(v for _, v in seq) / 1000)
for k, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:50s} n={c:4d} sum_us={v/1000:9.1f}")
PY
fi
