#!/bin/bash
# one ncu --set full capture: usage gpu_ncu_one.sh <tag> <kernel regex> [n recordings]
tag=${1:-rX}; rx=${2:-sn_rowmed}; n=${3:-24}
mkdir -p gpurun_out
timeout 300 python tools/prof_signal_noise.py $n 1 > gpurun_out/${tag}_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$rx -s 2 -c 1 -f -o gpurun_out/${tag}_$rx python tools/prof_signal_noise.py $n 1 > gpurun_out/${tag}_ncu.log 2>&1
ls -la gpurun_out/${tag}_*
