#!/usr/bin/env bash
# the reference's benchmark axes on the final build (estimate / decode / top-k sweeps against its kernels built for sm_100a)
set -u
mkdir -p gpurun_out
timeout 1200 python tools/ref_sweeps.py > gpurun_out/r2s38_ref_sweeps.txt 2> gpurun_out/r2s38_ref_sweeps.err; echo "sweeps rc=$?"; tail -5 gpurun_out/r2s38_ref_sweeps.txt
