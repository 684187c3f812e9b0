#!/bin/bash
# quick GPU check: parity tests (optionally a -k filter) + secondary-path timings
tag=${1:-q}; shift
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q "$@" > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${tag}_pytest.log
tail -30 gpurun_out/${tag}_pytest.log
