#!/usr/bin/env bash
# 8-GPU bench line as the driver launches it (headline weak scaling, configs[3] kv-head sharded, configs[4] sequence sharded)
set -u
mkdir -p gpurun_out
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29537 bench.py --gpus 4 --steps 20 --warmup 5 ) > gpurun_out/r2s29_bench4.json 2> gpurun_out/r2s29_bench4.err; echo "bench4 rc=$?"
python - <<'PY'
import json
t=open('gpurun_out/r2s29_bench4.json').read()
d=json.loads([l for l in t.splitlines() if l.startswith('{')][0])
print('n', d['n_gpus'], 'value', d['value'], 'us/layer', d['us_per_layer'], 'e2e', d['e2e']['value'])
for k,v in d['configs'].items():
    print(k, {kk: v[kk] for kk in v if kk.endswith('us_per_layer') or kk.startswith('frac') or kk.startswith('exchange_us') or kk.startswith('eff')})
PY
