#!/bin/bash
# Builds libpsb200.so in-tree for sm_100a (cross-compiles without a GPU).
set -e
cd "$(dirname "$0")"
OUT="${PSB_OUT:-parameter-server-architecture-on-apache-kafka_b200/ops/libpsb200.so}"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -O3 ${PSB_DEFS}"
BUILD="${PSB_BUILD_DIR:-build}"
mkdir -p "$BUILD"
OBJS=""
for src in csrc/*.cu csrc/*.cpp; do
  obj="$BUILD/$(basename "${src%.*}").o"
  if [ ! -f "$obj" ] || [ "$src" -nt "$obj" ] || [ -n "$(find csrc include -newer "$obj" \( -name '*.cuh' -o -name '*.h' \) | head -1)" ]; then
    rm -f "$obj"
    $NVCC $FLAGS ${PSB_PTXAS_V:+-Xptxas -v} -x cu -c "$src" -o "$obj" &
  fi
  OBJS="$OBJS $obj"
done
wait
for obj in $OBJS; do [ -f "$obj" ] || { echo "build failed: $obj missing"; exit 1; }; done
$NVCC -shared -gencode arch=compute_100a,code=sm_100a -o "$OUT" $OBJS -lcudart
echo "built $OUT"
