#!/bin/bash
# One GPU-box round: GPU parity tests, full bench line, ncu launch list + one full capture of the top kernel.
# usage (under gpurun): bash tools/gpu_round.sh <tag>
tag=${1:-rX}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${tag}_pytest.log
tail -5 gpurun_out/${tag}_pytest.log
python bench.py --steps 5 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"
cat gpurun_out/${tag}_bench.json
export CAC_BENCH_RECORDINGS=128 CAC_BENCH_POOL=16
SMALL="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
$SMALL > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${tag}_launches.csv $SMALL > gpurun_out/${tag}_ncu1.log 2>&1
$SMALL > gpurun_out/${tag}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:stft_mel -s 3 -c 1 -f -o gpurun_out/${tag}_stft_mel $SMALL > gpurun_out/${tag}_ncu2.log 2>&1
ls -la gpurun_out
