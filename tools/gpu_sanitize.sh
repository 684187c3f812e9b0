#!/bin/bash
# compute-sanitizer over the GPU parity tests (VERDICT r1 #8a).  usage (under gpurun): bash tools/gpu_sanitize.sh <tag> [-k expr]
# PYTORCH_NO_CUDA_MEMORY_CACHING=1: every tensor is its own cudaMalloc, so an out-of-bounds access inside what would be
# torch's pool is seen by memcheck.
tag=${1:-rX}
sel=${2:-not fullsize}
mkdir -p gpurun_out
export PYTORCH_NO_CUDA_MEMORY_CACHING=1
for tool in memcheck racecheck; do
  start=$(date +%s)
  timeout ${SAN_TIMEOUT:-1500} compute-sanitizer --tool $tool --print-limit 40 --log-file gpurun_out/${tag}_${tool}.log \
      python -m pytest tests -m gpu -q -x -k "$sel" > gpurun_out/${tag}_${tool}_pytest.log 2>&1
  echo "$tool exit $? after $(( $(date +%s) - start )) s" | tee -a gpurun_out/${tag}_${tool}_pytest.log
  tail -3 gpurun_out/${tag}_${tool}_pytest.log
  tail -5 gpurun_out/${tag}_${tool}.log
done
