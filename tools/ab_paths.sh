#!/bin/bash
# A/B timing of two library builds on the secondary paths (tools/bench_paths.py): usage ab_paths.sh "morepork ci"
for rep in 1 2; do for v in ${VARIANTS:-A B}; do
cp build_ab/lib$v.so audio-analysis_b200/libcacophony_b200.so
echo "== $v"; CAC_CPU_BASELINE=0 CAC_MOREPORK_RECORDINGS=1024 python tools/bench_paths.py $1 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print({k: (round(v, 4) if isinstance(v, float) else v) for k, v in d.items() if k in ('path', 'ms', 'ms_bins_plus_points', 'ms_spectrogram', 'ms_total', 'recording_hours_per_s', 'hbm_frac_of_measured')})"
done; done
