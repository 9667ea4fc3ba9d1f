#!/usr/bin/env bash
# 2-GPU sanity of the bench line (as the driver launches it) and of the sequence-sharded NCCL worker
set -u
mkdir -p gpurun_out
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 5 ) > gpurun_out/r2s28_bench2.json 2> gpurun_out/r2s28_bench2.err; echo "bench2 rc=$?"
python - <<'PY'
import json
t=open('gpurun_out/r2s28_bench2.json').read()
d=json.loads([l for l in t.splitlines() if l.startswith('{')][0])
print('n', d['n_gpus'], 'value', d['value'], 'us/layer', d['us_per_layer'], 'e2e', d['e2e']['value'])
for k,v in d['configs'].items():
    print(k, {kk: v[kk] for kk in v if kk.endswith('us_per_layer') or kk.startswith('frac') or kk.startswith('exchange_us') or kk.startswith('eff')})
PY
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 tests/seqshard_nccl_worker.py ) > gpurun_out/r2s28_seqshard.log 2>&1; echo "seqshard rc=$?"; tail -3 gpurun_out/r2s28_seqshard.log
