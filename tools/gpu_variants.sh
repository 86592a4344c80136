#!/bin/bash
# per-layer profile under tuning-knob variants: VARS="A=1,B=2 C=3" (comma = several env vars in one variant)
mkdir -p gpurun_out
i=0
for v in "" ${VARS:-}; do
  envs=$(echo "$v" | tr ',' ' ')
  tag=$(echo "${v:-default}" | tr -c 'A-Za-z0-9_=\n' '_')
  env $envs timeout 300 python tools/layer_profile.py ${N:-256} 3 > gpurun_out/layers_$tag.csv 2> gpurun_out/layers_$tag.err
  echo "$tag: $(tail -n 1 gpurun_out/layers_$tag.csv)"
done
