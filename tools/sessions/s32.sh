#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fp8_metadata.py tests/test_prefill_fused_gpu.py -x -q 2>&1 | tail -3
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "operator or surface or layer_step" 2>&1 | tail -2
