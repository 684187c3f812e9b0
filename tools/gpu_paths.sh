#!/bin/bash
# Secondary paths at full BASELINE sizes (configs[2], configs[3], the hop-281 spectra) with the CPU oracle beside them,
# the single-recording latency bench, then ncu --set full captures of ci_bins_kernel and spec4800_kernel<75>
# (each after the same command exited 0 without ncu).  usage (under gpurun): bash tools/gpu_paths.sh <tag>
tag=${1:-rX}
mkdir -p gpurun_out
python tools/bench_paths.py ci morepork hop281 > gpurun_out/${tag}_paths.json 2> gpurun_out/${tag}_paths.err; echo "paths exit $?"
cat gpurun_out/${tag}_paths.json
python tools/bench_single.py > gpurun_out/${tag}_single_recording.json 2> gpurun_out/${tag}_single.err; echo "single exit $?"
cat gpurun_out/${tag}_single_recording.json
for w in ci morepork; do
  k=ci_bins; [ $w = morepork ] && k=spec4800
  python tools/prof_paths.py $w > gpurun_out/${tag}_${w}_plain.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 1 -f -o gpurun_out/${tag}_$w python tools/prof_paths.py $w > gpurun_out/${tag}_${w}_ncu.log 2>&1
done
ls -la gpurun_out | tail -8
