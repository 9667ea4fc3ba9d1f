#!/usr/bin/env bash
# the no-allocation entry point: alone in a fresh process, then after the whole GPU suite (pytest -x, as the round's end runs it)
set -u
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r2s11_alone.log 2>&1; echo "alone rc=$?"; tail -3 gpurun_out/r2s11_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/r2s11_pytest.log 2>&1; echo "suite rc=$?"; tail -5 gpurun_out/r2s11_pytest.log
