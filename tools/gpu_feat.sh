#!/bin/bash
# feature-path check: parity tests of the fused kernel + a short device-only bench (128 recordings)
tag=${1:-f}
python -m pytest tests/test_features_gpu.py -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${tag}_pytest.log
tail -8 gpurun_out/${tag}_pytest.log
CAC_BENCH_RECORDINGS=256 CAC_BENCH_POOL=16 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2> gpurun_out/${tag}_bench.err | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('value', d['value'], 'ms/step', d['ms_per_step'], 'kernel_ms', d['roofline']['kernel_ms'], 'clocks', d['clocks'])"
tail -3 gpurun_out/${tag}_bench.err
