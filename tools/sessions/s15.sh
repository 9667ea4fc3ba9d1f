#!/usr/bin/env bash
# FHFMA issue rate, and which kernel's use of it slows the chained layer step (estimate / decode, separately)
set -u
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fhfma_probe tools/fhfma_probe.cu && tools/fhfma_probe | tee gpurun_out/r2s15_fhfma_probe.txt
for rep in 1 2; do
for v in _nof "" _eF_dN _eN_dF; do
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2s15_bench$v.json 2> gpurun_out/r2s15_bench$v.err
  python - "$v" <<'PY'
import json, sys
v = sys.argv[1]
try:
    d = json.load(open(f"gpurun_out/r2s15_bench{v}.json"))
    k = d["kernels"]
    print(f"{v or 'new':8s} us/layer {d['us_per_layer']:.2f} frac {d['frac_of_hbm_peak']:.3f} | est {k['estimate']['us']:.2f} topk {k['topk']['us']:.2f} dec {k['decode']['us']:.2f}")
except Exception as e:
    print(v, "failed", e)
PY
done
done
