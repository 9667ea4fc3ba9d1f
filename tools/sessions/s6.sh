#!/usr/bin/env bash
set -u
python bench.py --steps 20 --warmup 5 > gpurun_out/r2s6_bench.json 2> gpurun_out/r2s6_bench.err; echo "bench rc=$?"; tail -5 gpurun_out/r2s6_bench.err
python -c "
import json;d=json.load(open('gpurun_out/r2s6_bench.json'))
print({k:d[k] for k in ('us_per_layer','frac_of_hbm_peak','e2e_tok_s')}); print(json.dumps(d.get('configs'),indent=1)); print(json.dumps(d.get('e2e_textgen'),indent=1)); print(d['roofline'])"
