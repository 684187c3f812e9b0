#!/bin/bash
# A/B of the row-median kernels (CAC_SN_ROWS_VARIANT=0: scalar, 1: vector) + parity tests of the signal_noise chain;
# run under gpurun.
timeout 600 python -m pytest tests/test_signal_noise_gpu.py tests/test_pipeline_gpu.py tests/test_guard_gpu.py -x -q -m gpu 2>&1 | tail -8
for v in 0 1; do
  CAC_SN_ROWS_VARIANT=$v timeout 120 python tools/prof_signal_noise.py 48 5 2>&1 | grep signal_noise
done
CAC_SN_ROWS_VARIANT=1 timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/sn_rows_launches.csv python tools/prof_signal_noise.py 24 1 > /dev/null 2>&1
python tools/ncu_launches.py gpurun_out/sn_rows_launches.csv 2>&1 | grep -E "sn_|total"
