#!/bin/bash
# ncu --set full of the morphology / component kernels of one wave of the signal_noise chain; run under gpurun.
tag=${1:-rX}
mkdir -p gpurun_out
timeout 300 python tools/prof_signal_noise.py 24 1 > gpurun_out/${tag}_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"morph_|runs_" -s 12 -c 12 -f -o gpurun_out/${tag}_morph python tools/prof_signal_noise.py 24 1 > gpurun_out/${tag}_ncu.log 2>&1
ncu -i gpurun_out/${tag}_morph.ncu-rep --page raw --csv --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,launch__registers_per_thread 2>/dev/null | cut -d, -f5,12- | head -20
