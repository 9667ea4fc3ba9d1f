#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
python -m pytest tests/test_prefill_fused_gpu.py -m gpu -x -q > gpurun_out/r2s20_fused.log 2>&1; echo "fused rc=$?"; tail -4 gpurun_out/r2s20_fused.log
python tools/prefill_producer_bench.py 2>&1 | tee gpurun_out/r2s20_producer.txt
python tools/prefill_producer_bench.py --hkv 8 --lens 32768 2>&1 | tee -a gpurun_out/r2s20_producer.txt
( time python bench.py --no-cpu-baseline ) > gpurun_out/r2s20_bench.json 2> gpurun_out/r2s20_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
t=open('gpurun_out/r2s20_bench.json').read()
d=json.loads([l for l in t.splitlines() if l.startswith('{')][0])
print('us/layer', d['us_per_layer'], 'frac', d['frac_of_hbm_peak'], 'e2e', d['e2e']['ms_per_step'], 'tok/s', d.get('e2e_tok_s'))
for k,v in d['configs'].items():
    print(k, {kk: v[kk] for kk in v if kk.endswith('us_per_layer') or kk.startswith('frac') or kk.startswith('exchange_us')})
PY
