#!/bin/bash
# morepork path (256 recordings): plain run, ncu launch list with DRAM bytes, one --set full capture of band_gemm_kernel.
# usage (under gpurun): bash tools/gpu_prof_morepork.sh <tag>
tag=${1:-rX}
mkdir -p gpurun_out
timeout 120 python tools/prof_paths.py morepork > gpurun_out/${tag}_mp_plain.log 2>&1 || { echo "plain failed"; tail -5 gpurun_out/${tag}_mp_plain.log; exit 1; }
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/${tag}_mp_launches.csv python tools/prof_paths.py morepork > gpurun_out/${tag}_mp_ncu1.log 2>&1
python tools/ncu_launches.py gpurun_out/${tag}_mp_launches.csv 2>/dev/null | tail -12
timeout 300 ncu --set full --clock-control none --import-source on -k regex:band_gemm -s 2 -c 1 -f -o gpurun_out/${tag}_band_gemm python tools/prof_paths.py morepork > gpurun_out/${tag}_mp_ncu2.log 2>&1
python tools/ncu_summary.py gpurun_out/${tag}_band_gemm.ncu-rep | tee gpurun_out/${tag}_band_gemm_summary.txt
