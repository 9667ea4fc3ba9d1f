#!/usr/bin/env bash
# s67: tagged merge in the hand-over decode launch (-DQUEST_DEC_TAG_MERGE=1, libquest_b200_exp.so) against the product build, then the
# bit-equality tests of the layer step on the candidate
set -u
mkdir -p gpurun_out
EXP=$PWD/quest_b200/libquest_b200_exp.so
us() { python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(round(d['us_per_layer'],3), round(d['eager_us_per_layer'],2), round(d['e2e']['ms_per_step'],3))"; }
echo -n "base: "; timeout 60 python bench.py --steps 40 --warmup 5 --no-cpu-baseline --no-extras 2>gpurun_out/r2s67_base.err | tee gpurun_out/r2s67_base.json | us
for rep in 1 2; do
  echo -n "exp rep=$rep: "; QUEST_B200_LIB=$EXP timeout 60 python bench.py --steps 40 --warmup 5 --no-cpu-baseline --no-extras 2>gpurun_out/r2s67_exp_$rep.err | tee gpurun_out/r2s67_exp_$rep.json | us
done
( time QUEST_B200_LIB=$EXP timeout 55 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "handoff_stress or (four_operator_calls and True) or matches_oracle" ) > gpurun_out/r2s67_parity_exp.log 2>&1; echo "parity(exp) rc=$?"; tail -4 gpurun_out/r2s67_parity_exp.log
