#!/bin/bash
# One `ncu --set full` capture of stft_mel_kernel at configs[1] size (1,024 recordings per launch), after the same
# command has exited 0 without ncu; prints the summary metrics.  usage (under gpurun): bash tools/gpu_capture_stft_mel.sh <tag>
tag=${1:-rX}
mkdir -p gpurun_out
# LIB=build_ab/libB.so: capture a variant build (tools/build_variant.sh); CAC_BENCH_RECORDINGS / CAC_BENCH_POOL shrink the launch
if [ -n "$LIB" ]; then cp audio-analysis_b200/libcacophony_b200.so /tmp/lib_default.so; cp $LIB audio-analysis_b200/libcacophony_b200.so; fi
FULL="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-paths"
timeout 600 $FULL > gpurun_out/${tag}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${tag}_plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stft_mel -s 3 -c 1 -f -o gpurun_out/${tag}_stft_mel $FULL > gpurun_out/${tag}_ncu.log 2>&1
python tools/ncu_summary.py gpurun_out/${tag}_stft_mel.ncu-rep | tee gpurun_out/${tag}_stft_mel_summary.txt
python tools/ncu_regions.py gpurun_out/${tag}_stft_mel.ncu-rep > gpurun_out/${tag}_stft_mel_regions.txt 2>&1; head -3 gpurun_out/${tag}_stft_mel_regions.txt
if [ -n "$LIB" ]; then cp /tmp/lib_default.so audio-analysis_b200/libcacophony_b200.so; fi
ls -la gpurun_out/${tag}_*
