#!/usr/bin/env bash
# s65: estimate kernel releases its dependents after its prologue (-DQUEST_TRIG_EST=1) with the top-k kernel releasing the hand-over
# decode launch only from behind its wait (libquest_b200_exp.so) against the product build; parity subset on the candidate if it wins
set -u
mkdir -p gpurun_out
EXP=$PWD/quest_b200/libquest_b200_exp.so
us() { python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(round(d['us_per_layer'],3), {k: round(v['us'],2) for k,v in d['kernels'].items()})"; }
for rep in 1 2; do
  echo -n "base rep=$rep: "; timeout 100 python bench.py --steps 40 --warmup 5 --no-cpu-baseline --no-extras 2>gpurun_out/r2s65_base_$rep.err | tee gpurun_out/r2s65_base_$rep.json | us
  echo -n "exp  rep=$rep: "; QUEST_B200_LIB=$EXP timeout 100 python bench.py --steps 40 --warmup 5 --no-cpu-baseline --no-extras 2>gpurun_out/r2s65_exp_$rep.err | tee gpurun_out/r2s65_exp_$rep.json | us
done
win=$(python - <<'PY'
import json
def f(n): return json.loads(open(n).read().strip().splitlines()[-1])['us_per_layer']
b=min(f('gpurun_out/r2s65_base_1.json'),f('gpurun_out/r2s65_base_2.json')); e=max(f('gpurun_out/r2s65_exp_1.json'),f('gpurun_out/r2s65_exp_2.json'))
print(1 if b-e>=0.2 else 0)
PY
)
echo "candidate wins by >= 0.2 us: $win"
if [ "$win" = "1" ]; then
  ( time QUEST_B200_LIB=$EXP timeout 130 python -m pytest tests/test_gpu_parity.py tests/test_batch_step_gpu.py tests/test_seq_shard_gpu.py tests/test_fp8_metadata.py -x -q -m gpu -k "layer_step or handoff or estimate or batch or shard or fp8 or golden" ) > gpurun_out/r2s65_parity_exp.log 2>&1; echo "parity(exp) rc=$?"; tail -3 gpurun_out/r2s65_parity_exp.log
fi
