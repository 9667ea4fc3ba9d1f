#!/usr/bin/env bash
# bisect of the chained-step regression: HEAD estimate / HEAD decode objects mixed with the new ones
set -u
mkdir -p gpurun_out
for rep in 1 2; do
for v in _base _v3 _v1 _v2 ""; do
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2s17_bench$v.json 2> gpurun_out/r2s17_bench$v.err
  python - "$v" <<'PY'
import json, sys
v = sys.argv[1]
try:
    d = json.load(open(f"gpurun_out/r2s17_bench{v}.json"))
    k = d["kernels"]
    print(f"{v or 'new':8s} us/layer {d['us_per_layer']:.2f} frac {d['frac_of_hbm_peak']:.3f} | est {k['estimate']['us']:.2f} topk {k['topk']['us']:.2f} dec {k['decode']['us']:.2f}")
except Exception as e:
    print(v, "failed", e)
PY
done
done
