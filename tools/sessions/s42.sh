#!/usr/bin/env bash
# attend phase with the row squeezed over the head's chunks: parity, phases, headline unchanged?
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_seq_shard_gpu.py tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q -k "not prefill and not estimate and not topk" 2>&1 | tail -2
for r in 0 7; do timeout 600 python tools/seqshard_emulate.py --world 8 --rank $r 2>&1 | grep -v Warning | tail -6; done
timeout 600 python tools/seqshard_emulate.py --world 2 --rank 1 2>&1 | grep -v Warning | tail -6
timeout 600 python tools/seqshard_emulate.py --world 1 --rank 0 2>&1 | grep -v Warning | tail -6
for i in 1 2; do timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2>/dev/null | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print("us/layer", round(d["us_per_layer"],2), "dec", round(d["kernels"]["decode"]["us"],2), "e2e ms", round(d["e2e"]["ms_per_step"],3))'; done
