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
  if [ ! -f "$obj" ] || [ "$src" -nt "$obj" ] || [ -n "$(find csrc include -newer "$obj" \( -name '*.cuh' -o -name '*.h' -o -name '*.hpp' \) | head -1)" ]; then
    rm -f "$obj"
    $NVCC $FLAGS ${PSB_PTXAS_V:+-Xptxas -v} -x cu -c "$src" -o "$obj" &
  fi
  OBJS="$OBJS $obj"
done
wait
for obj in $OBJS; do [ -f "$obj" ] || { echo "build failed: $obj missing"; exit 1; }; done
stale=0
[ -f "$OUT" ] || stale=1
for obj in $OBJS; do [ "$obj" -nt "$OUT" ] && stale=1; done
if [ "$stale" = 1 ]; then
  # linked under another name and moved into place: a process that has the old library loaded keeps its own copy
  $NVCC -shared -gencode arch=compute_100a,code=sm_100a -o "$OUT.new" $OBJS -lcudart && mv -f "$OUT.new" "$OUT"
  echo "built $OUT"
else
  echo "up to date: $OUT"
fi
