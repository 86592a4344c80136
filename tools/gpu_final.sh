#!/bin/bash
# short closing check: median/defense parity, HBM-bound kernel rates, smoke, default bench
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_parity.py -x -q -k "median or defenses_bit_exact" > gpurun_out/final_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/final_pytest.log
ADVB200_WORKLOADS_KERNELS_ONLY=1 timeout 60 python tools/workloads_bench.py 256 > gpurun_out/final_kernels.json 2> gpurun_out/final_kernels.err; echo "kernels rc=$?"; cat gpurun_out/final_kernels.json
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/final_smoke.log
timeout 200 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench rc=$?"; cat gpurun_out/final_bench.json
