#!/bin/bash
# signal_noise experiments: parity tests of the chain (default library), then timings of build_ab/lib{A,B}.so with the
# row-kernel variants, launch list of the default library.
tag=${1:-rX}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -k "signal_noise or pipeline or spec_gpu" > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${tag}_pytest.log
tail -3 gpurun_out/${tag}_pytest.log
cp audio-analysis_b200/libcacophony_b200.so /tmp/lib_default.so
for lib in ${LIBS:-A B}; do cp build_ab/lib$lib.so audio-analysis_b200/libcacophony_b200.so
for v in ${ROW_VARIANTS:-0 1 2 3}; do echo "lib $lib rows variant $v"; CAC_SN_ROWS_VARIANT=$v timeout 300 python tools/prof_signal_noise.py 48 3; done; done 2>&1 | tee gpurun_out/${tag}_variants.log
cp /tmp/lib_default.so audio-analysis_b200/libcacophony_b200.so
for v in ${NCU_VARIANTS:-1}; do
CAC_SN_ROWS_VARIANT=$v timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/${tag}_sn_launches_v$v.csv python tools/prof_signal_noise.py 24 1 > gpurun_out/${tag}_sn_ncu.log 2>&1
done
