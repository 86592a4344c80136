#!/bin/bash
# Builds libadvb200.so (sm_100a only) in-tree.  Usage: ./build.sh [extra nvcc flags]
set -euo pipefail
ROOT="$(cd "$(dirname "$0")" && pwd)"
PKG="$ROOT/image-recognition-adversarial-example-attack_b200"
SRC="$PKG/csrc"
OUT="$PKG/lib"
mkdir -p "$OUT" "$OUT/obj"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS=(-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xptxas -v "$@")
pids=()
for f in elementwise jpeg preprocess layers transformer simt_conv tc_conv engine; do
  ( "$NVCC" "${FLAGS[@]}" -c "$SRC/$f.cu" -o "$OUT/obj/$f.o" > "$OUT/obj/$f.log" 2>&1 || { cat "$OUT/obj/$f.log"; exit 1; } ) &
  pids+=($!)
done
for p in "${pids[@]}"; do wait "$p"; done
"$NVCC" -shared -o "$OUT/libadvb200.so" "$OUT"/obj/*.o -lcudart_static -lpthread -ldl -lrt
echo "built $OUT/libadvb200.so"
