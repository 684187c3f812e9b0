#!/bin/bash
# A/B of the two fused kernels in ONE library: CAC_FEAT_MODE=cta (CTA-wide stages) vs the default (warp groups).
# 256-recording device-resident bench, twice each; the checksums must agree bit for bit.
for rep in 1 2; do for v in cta groups; do
CAC_FEAT_MODE=$v CAC_BENCH_RECORDINGS=${CAC_BENCH_RECORDINGS:-256} CAC_BENCH_POOL=16 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('$v', 'value %.2f' % d['value'], 'ms/step %.3f' % d['ms_per_step'], 'kernel_ms %.3f' % d['roofline']['kernel_ms'], 'checksum', d['checksum_first_recording'])"
done; done
