#!/usr/bin/env bash
# A/B: programmatic-launch trigger points (EST/TOPK/DEC) with the bin-total top-k, and 8-warp decode CTAs
set -u
mkdir -p gpurun_out
for v in "" _t010 _t110 _t111 _w8 _w8b; do
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2s12_bench$v.json 2> gpurun_out/r2s12_bench$v.err
  python - "$v" <<'PY'
import json, sys
v = sys.argv[1]
try:
    d = json.load(open(f"gpurun_out/r2s12_bench{v}.json"))
    k = d["kernels"]
    print(f"{v or 'base':6s} us/layer {d['us_per_layer']:.2f} frac {d['frac_of_hbm_peak']:.3f} | est {k['estimate']['us']:.2f} topk {k['topk']['us']:.2f} dec {k['decode']['us']:.2f} | ppc {d['config']['decode_plan_pages_per_chunk']}")
except Exception as e:
    print(v, "failed", e)
PY
done
