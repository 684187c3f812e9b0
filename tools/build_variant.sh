#!/bin/bash
# Builds a variant of the library with extra nvcc flags into build_ab/lib<name>.so (the default build stays in place).
# usage: tools/build_variant.sh <name> "<flags>"       e.g. tools/build_variant.sh B "-DCAC_STFT_SKEW=1"
set -e
name=$1; flags=$2
root=$(cd "$(dirname "$0")/.." && pwd)
src=$root/audio-analysis_b200/csrc
out=$root/build_ab
mkdir -p $out/obj_$name
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
for f in $src/*.cu; do
  b=$(basename ${f%.cu})
  $NVCC -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -Xcompiler -fPIC $flags -c $f -o $out/obj_$name/$b.o &
done
wait
$NVCC -shared -Wno-deprecated-gpu-targets -o $out/lib$name.so $out/obj_$name/*.o -lcudart
echo "built $out/lib$name.so ($flags)"
