#!/bin/bash
# GPU tests with the default library, the signal_noise launch list, then A/B of build_ab/lib{A,B}.so on the headline
# step (256 recordings) and on the signal_noise chain.   usage (under gpurun): bash tools/gpu_ab.sh <tag>
tag=${1:-rX}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${tag}_pytest.log
tail -6 gpurun_out/${tag}_pytest.log
timeout 300 python tools/prof_signal_noise.py 48 3 2>&1 | tee gpurun_out/${tag}_sn_plain.log
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/${tag}_sn_launches.csv python tools/prof_signal_noise.py 24 1 > gpurun_out/${tag}_sn_ncu.log 2>&1
cp audio-analysis_b200/libcacophony_b200.so /tmp/lib_default.so
VARIANTS="${VARIANTS:-A B}" bash tools/ab.sh 2>&1 | tee gpurun_out/${tag}_ab.log
for v in ${VARIANTS:-A B}; do cp build_ab/lib$v.so audio-analysis_b200/libcacophony_b200.so; echo "variant $v"; timeout 300 python tools/prof_signal_noise.py 48 3; done 2>&1 | tee -a gpurun_out/${tag}_ab.log
cp /tmp/lib_default.so audio-analysis_b200/libcacophony_b200.so
