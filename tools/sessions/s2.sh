#!/usr/bin/env bash
set -u
export QUEST_B200_LIB=quest_b200/libquest_b200_dbg.so
python tools/step_timeline.py 2>&1 | grep -v Warning | tee gpurun_out/r2s2_timeline_fused.txt
BENCH_UNFUSED=1 python tools/step_timeline.py 2>&1 | grep -v Warning | tee gpurun_out/r2s2_timeline_unfused.txt
