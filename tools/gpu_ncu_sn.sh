#!/bin/bash
# ncu --set full captures of the two big kernels of the batched signal_noise chain (24 one-minute recordings).
tag=${1:-rX}
mkdir -p gpurun_out
timeout 300 python tools/prof_signal_noise.py 24 1 > gpurun_out/${tag}_plain.log 2>&1 || exit 1
cat gpurun_out/${tag}_plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sn_stft -s 2 -c 1 -f -o gpurun_out/${tag}_sn_stft python tools/prof_signal_noise.py 24 1 > gpurun_out/${tag}_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sn_rowmed -s 2 -c 1 -f -o gpurun_out/${tag}_sn_rowmed python tools/prof_signal_noise.py 24 1 > gpurun_out/${tag}_ncu2.log 2>&1
ls -la gpurun_out | tail -5
