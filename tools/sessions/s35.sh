#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python tools/four_ops_profile.py > gpurun_out/r2s35_four_ops.txt 2>&1; echo rc=$?; head -90 gpurun_out/r2s35_four_ops.txt
