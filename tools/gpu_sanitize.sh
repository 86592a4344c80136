#!/bin/bash
# compute-sanitizer over the native kernels: memcheck, racecheck, synccheck and initcheck.
#   pass 1: tools/sanitize_pass.py — every kernel family, whole-network plan, heuristic tiles (autotuning off: the tuner
#           would only repeat the same kernels a few thousand times under the tool)
#   pass 2: the conv layer tests with every forced tile configuration (memcheck only; the candidates the tuner picks from)
# Logs: gpurun_out/r2_sanitizer_<tool>.log; copy the summaries to profiles/.
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
export ADVB200_NO_AUTOTUNE=1
for tool in memcheck racecheck synccheck initcheck; do
  timeout 1500 $CS --tool $tool --print-limit 20 --error-exitcode 9 python tools/sanitize_pass.py 3 > gpurun_out/r2_sanitizer_$tool.log 2>&1
  echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize pass|Error|error" gpurun_out/r2_sanitizer_$tool.log | head -8
done
unset ADVB200_NO_AUTOTUNE
timeout 1500 $CS --tool memcheck --print-limit 20 --error-exitcode 9 python -m pytest tests/test_gpu_conv_layers.py tests/test_gpu_cw_kernels.py -m gpu -q -x \
  -k "not injected_noise" > gpurun_out/r2_sanitizer_memcheck_layers.log 2>&1
echo "memcheck layers rc=$?"; grep -E "ERROR SUMMARY|passed|failed" gpurun_out/r2_sanitizer_memcheck_layers.log | head
