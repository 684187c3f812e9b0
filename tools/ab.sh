#!/bin/bash
# A/B timing of two builds of the library (build_ab/libA.so, libB.so): 256-recording device-resident bench, twice each
for rep in 1 2; do for v in ${VARIANTS:-A B}; do
cp build_ab/lib$v.so audio-analysis_b200/libcacophony_b200.so
CAC_BENCH_RECORDINGS=256 CAC_BENCH_POOL=16 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('$v', 'value %.2f' % d['value'], 'ms/step %.3f' % d['ms_per_step'], 'kernel_ms %.3f' % d['roofline']['kernel_ms'], 'checksum', d['checksum_first_recording'])"
done; done
