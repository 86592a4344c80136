#!/bin/bash
# closing check 2: HBM-bound kernel rates first (seconds), then the whole GPU suite
mkdir -p gpurun_out
ADVB200_WORKLOADS_KERNELS_ONLY=1 timeout 60 python tools/workloads_bench.py 256 > gpurun_out/final2_kernels.json 2> gpurun_out/final2_kernels.err; echo "kernels rc=$?"; cat gpurun_out/final2_kernels.json
timeout 250 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/final2_pytest.log 2>&1; echo "pytest rc=$?"; tail -16 gpurun_out/final2_pytest.log
