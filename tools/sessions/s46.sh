#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 1500 python tools/fuzz_parity.py --n 260 --seed 11 > gpurun_out/r2s46_fuzz.txt 2>&1; echo "rc=$?"; tail -12 gpurun_out/r2s46_fuzz.txt
