#!/bin/bash
# Launch list of one wave of the signal_noise chain + ncu --set full of its two big kernels; run under gpurun.
tag=${1:-rX}
mkdir -p gpurun_out
timeout 300 python tools/prof_signal_noise.py 48 5 > gpurun_out/${tag}_plain.log 2>&1 || exit 1
cat gpurun_out/${tag}_plain.log
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/${tag}_sn_launches.csv python tools/prof_signal_noise.py 24 1 > /dev/null 2>&1
python tools/ncu_launches.py gpurun_out/${tag}_sn_launches.csv 3
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sn_stft -s 2 -c 1 -f -o gpurun_out/${tag}_sn_stft python tools/prof_signal_noise.py 24 1 > gpurun_out/${tag}_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sn_rowmed -s 2 -c 1 -f -o gpurun_out/${tag}_sn_rowmed python tools/prof_signal_noise.py 24 1 > gpurun_out/${tag}_ncu2.log 2>&1
