#!/bin/bash
# Builds libcacophony_b200.so in-tree for sm_100a (cross-compiles without a GPU).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -Xcompiler -fPIC $EXTRA"
OUT=../libcacophony_b200.so
newest=$(ls -t *.cu *.cuh ../../include/*.h build.sh | head -1)
if [ -f "$OUT" ] && [ "$OUT" -nt "$newest" ] && [ -z "$FORCE" ]; then exit 0; fi
for f in *.cu; do
  $NVCC $FLAGS ${PTXAS_V:+-Xptxas -v} -c "$f" -o "${f%.cu}.o" &
done
wait
$NVCC -shared -Wno-deprecated-gpu-targets -o "$OUT" *.o -lcudart
echo "built $OUT"
