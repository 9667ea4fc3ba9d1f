#!/usr/bin/env bash
# hand-over A/B over the decode plan's pages per chunk (does the decode grid fit beside the resident top-k CTAs?)
set -u
mkdir -p gpurun_out
for ppc in 4 5 6 8; do
for v in 1 ""; do
  BENCH_PPC=$ppc BENCH_NO_HANDOFF=$v timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2s24_bench.json 2> gpurun_out/r2s24_bench.err
  python - "$v" "$ppc" <<'PY'
import json, sys
v, ppc = sys.argv[1], sys.argv[2]
try:
    d = json.load(open("gpurun_out/r2s24_bench.json"))
    k = d["kernels"]
    print(f"ppc {ppc} {'no handoff' if v else 'handoff   '} us/layer {d['us_per_layer']:.2f} frac {d['frac_of_hbm_peak']:.3f} | est {k['estimate']['us']:.2f} topk {k['topk']['us']:.2f} dec {k['decode']['us']:.2f}")
except Exception as e:
    print(v, "failed", e)
PY
done
done
