#!/bin/bash
# Full-size capture of the headline step: GPU parity tests, bench line (both arms), launch list and one
# `ncu --set full` capture of stft_mel_kernel at configs[1] size (1,024 recordings per launch).
# usage (under gpurun): bash tools/gpu_full.sh <tag>
tag=${1:-rX}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${tag}_pytest.log
tail -3 gpurun_out/${tag}_pytest.log
python bench.py --steps 5 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"
cat gpurun_out/${tag}_bench.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2> gpurun_out/${tag}_bench_reference.err; echo "reference arm exit $?"
cat gpurun_out/${tag}_bench_reference.json
FULL="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
$FULL > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/${tag}_launches.csv $FULL > gpurun_out/${tag}_ncu1.log 2>&1
$FULL > gpurun_out/${tag}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:stft_mel -s 3 -c 1 -f -o gpurun_out/${tag}_stft_mel $FULL > gpurun_out/${tag}_ncu2.log 2>&1
ls -la gpurun_out | tail -12
