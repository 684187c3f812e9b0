#!/bin/bash
# signal_noise chain timing: default library, then each variant library given (build_ab/lib<name>.so); run under gpurun.
timeout 600 python -m pytest tests/test_signal_noise_gpu.py tests/test_pipeline_gpu.py tests/test_guard_gpu.py -x -q -m gpu 2>&1 | tail -2
for rep in 1 2; do timeout 120 python tools/prof_signal_noise.py 48 5 2>&1 | grep signal_noise; done
cp audio-analysis_b200/libcacophony_b200.so /tmp/lib_default.so
for v in "$@"; do
  cp build_ab/lib$v.so audio-analysis_b200/libcacophony_b200.so
  echo "variant $v"; for rep in 1 2; do timeout 120 python tools/prof_signal_noise.py 48 5 2>&1 | grep signal_noise; done
done
cp /tmp/lib_default.so audio-analysis_b200/libcacophony_b200.so
