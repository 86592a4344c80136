#!/bin/bash
# HBM-bound kernels: bit-exactness tests, GB/s at N = 256 (CUDA events), then one ncu --set full launch of each
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_jpeg.py tests/test_gpu_cw_kernels.py -m gpu -q -x -k "defense or median or divide or edge or jpeg or cw or random_start or linf or misaligned" 2>&1 | tail -4
ADVB200_WORKLOADS_KERNELS_ONLY=1 python tools/workloads_bench.py 256 > gpurun_out/r2_kernels_b256.json 2> gpurun_out/r2_kernels_b256.err; echo "kernels rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2_kernels_b256.json"))
for k, v in d["kernels"].items():
    print(f"{k:22s} {v['ms']*1e3:7.1f} us {v['gbs']:7.0f} GB/s  {v['frac_of_hbm_peak']:.2f} of HBM peak ({v['bytes_per_elem']} B/elem)")
PY
timeout 90 python tools/defense_pass.py 256 > gpurun_out/defense_pass.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/defense_pass.log; exit 1; }
timeout 300 ncu --set full --clock-control none --profile-from-start off --import-source on -f -o gpurun_out/r2_defense_kernels \
  python tools/defense_pass.py 256 > gpurun_out/ncu_defenses.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_defenses.log
ncu -i gpurun_out/r2_defense_kernels.ncu-rep --page raw --csv > gpurun_out/r2_defense_raw.csv 2>/dev/null
ls -la gpurun_out/r2_defense_raw.csv
