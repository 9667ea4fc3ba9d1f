#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_seq_shard_gpu.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29536 tools/bench_configs.py --config 5 --exchange peer --steps 20 --warmup 5 2>&1 | grep -v Warning | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print({k: round(v,2) for k,v in d.items() if k.endswith("us_per_layer")})'
