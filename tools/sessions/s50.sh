#!/bin/bash
# s50: device-side record updater (run-ahead graph) test + the batch-step suite
python -m pytest tests/test_batch_step_gpu.py -x -q -m gpu 2>&1 | tail -15
echo "batch rc=$?"
