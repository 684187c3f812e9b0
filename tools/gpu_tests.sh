#!/bin/bash
# GPU parity tests only (no bench): usage (under gpurun): bash tools/gpu_tests.sh <tag> [pytest -k expression]
tag=${1:-rX}
mkdir -p gpurun_out
if [ -n "$2" ]; then
  timeout 1500 python -m pytest tests -m gpu -q -s -k "$2" > gpurun_out/${tag}_pytest.log 2>&1
else
  timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/${tag}_pytest.log 2>&1
fi
echo "pytest exit $?" >> gpurun_out/${tag}_pytest.log
tail -40 gpurun_out/${tag}_pytest.log
