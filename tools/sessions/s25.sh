#!/usr/bin/env bash
# hand-over: early against late release of the decode launch; and at the 128K shapes
set -u
mkdir -p gpurun_out
for rep in 1 2; do
for v in "" _late; do
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2s25_bench.json 2> gpurun_out/r2s25_bench.err
  python - "$v" <<'PY'
import json, sys
v = sys.argv[1]
try:
    d = json.load(open("gpurun_out/r2s25_bench.json"))
    print(f"{v or 'early':6s} us/layer {d['us_per_layer']:.2f} frac {d['frac_of_hbm_peak']:.3f}")
except Exception as e:
    print(v, "failed", e)
PY
done
done
for shape in "--hq 32 --hkv 32 --ctx 131072 --budget 8192" "--hq 32 --hkv 8 --ctx 131072 --budget 4096" "--hq 32 --hkv 32 --ctx 65536 --budget 4096"; do
  for v in "" _late; do
    echo "lib$v handoff:    $(QUEST_B200_LIB=quest_b200/libquest_b200$v.so python tools/kernel_breakdown.py $shape 2>&1 | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["shape"]["pages"], d["layer_step"]["us"], d["layer_step_fp8"]["us"])')"
  done
  echo "no handoff:        $(QUEST_NO_HANDOFF=1 python tools/kernel_breakdown.py $shape 2>&1 | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["shape"]["pages"], d["layer_step"]["us"], d["layer_step_fp8"]["us"])')"
done
