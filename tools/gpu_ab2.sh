#!/bin/bash
# Parity tests + A/B timing of the variant libraries in build_ab/ (tools/build_variant.sh).
# usage (under gpurun): VARIANTS="A B" bash tools/gpu_ab2.sh <tag> [pytest -k expression]
tag=${1:-rX}; sel=${2:-features or fullsize or pipeline}
mkdir -p gpurun_out
cp audio-analysis_b200/libcacophony_b200.so /tmp/lib_default.so
for v in ${VARIANTS:-A B}; do
  cp build_ab/lib$v.so audio-analysis_b200/libcacophony_b200.so
  timeout 900 python -m pytest tests -m gpu -q -x -k "$sel" > gpurun_out/${tag}_pytest_$v.log 2>&1; echo "variant $v pytest exit $?" | tee -a gpurun_out/${tag}_pytest_$v.log
  tail -3 gpurun_out/${tag}_pytest_$v.log
done
for rep in 1 2; do for v in ${VARIANTS:-A B}; do
cp build_ab/lib$v.so audio-analysis_b200/libcacophony_b200.so
CAC_BENCH_RECORDINGS=${AB_RECORDINGS:-256} CAC_BENCH_POOL=16 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-paths 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('$v', 'value %.2f' % d['value'], 'ms/step %.3f' % d['ms_per_step'], 'kernel_ms %.3f' % d['roofline']['kernel_ms'], 'checksum', d['checksum_first_recording'])"
done; done 2>&1 | tee gpurun_out/${tag}_ab.log
cp /tmp/lib_default.so audio-analysis_b200/libcacophony_b200.so
