#!/bin/bash
# One GPU-box round (round 2): selected parity tests, the full bench line (with `paths`), the reference arm, ncu launch
# lists of the headline step and of the batched classify prep.  usage (under gpurun): bash tools/gpu_round2.sh <tag> [-k expr]
tag=${1:-rX}
mkdir -p gpurun_out
if [ -n "$2" ]; then
timeout 1200 python -m pytest tests -m gpu -q -s -k "$2" > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${tag}_pytest.log
tail -4 gpurun_out/${tag}_pytest.log
fi
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"
tail -3 gpurun_out/${tag}_bench.err
cat gpurun_out/${tag}_bench.json
timeout 300 python tools/prof_signal_noise.py 48 3 > gpurun_out/${tag}_sn_plain.log 2>&1 && cat gpurun_out/${tag}_sn_plain.log &&
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/${tag}_sn_launches.csv python tools/prof_signal_noise.py 24 1 > gpurun_out/${tag}_sn_ncu.log 2>&1
ls -la gpurun_out | tail -12
