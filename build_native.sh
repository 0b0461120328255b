#!/bin/bash
# Build the native sm_100a library in-tree (cross-compiles without a GPU).
set -e
cd "$(dirname "$0")"
OUT=commefficient_b200/lib
mkdir -p "$OUT" build
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden --expt-relaxed-constexpr"
objs=""
pids=""
for src in csrc/*.cu; do
  obj=build/$(basename "${src%.cu}").o
  objs="$objs $obj"
  if [ ! -f "$obj" ] || [ "$src" -nt "$obj" ] || [ -n "$(find csrc include -newer "$obj" \( -name '*.cuh' -o -name '*.h' \) | head -1)" ]; then
    $NVCC $FLAGS ${PTXAS_V:+-Xptxas -v} -c "$src" -o "$obj" &
    pids="$pids $!"
  fi
done
for p in $pids; do wait $p; done
$NVCC -shared -o "$OUT/libceb200.so" $objs -gencode arch=compute_100a,code=sm_100a -lcuda -lcudart
echo "built $OUT/libceb200.so"
