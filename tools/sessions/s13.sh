#!/usr/bin/env bash
# re-entry baseline: full GPU suite as the round's end runs it, then the bench line (with configs block + e2e tok/s)
set -u
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r2s13_pytest.log 2>&1; echo "suite rc=$?"; tail -5 gpurun_out/r2s13_pytest.log
( time python bench.py ) > gpurun_out/r2s13_bench.json 2> gpurun_out/r2s13_bench.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/r2s13_bench.json
