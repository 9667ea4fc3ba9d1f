#!/bin/bash
# s57: hist top-k over a cluster (rows of 4097..16384): parity, batch/seq-shard suites, timings
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "topk or layer_step" 2>&1 | tail -3
python -m pytest tests/test_batch_step_gpu.py tests/test_seq_shard_gpu.py tests/test_full_size_gpu.py -x -q -m gpu 2>&1 | tail -2
python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 131072 --budget 4096 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('gqa128k', {k:d[k]['us'] for k in d if k!='shape'})"
python tools/kernel_breakdown.py --ctx 131072 --budget 8192 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('mha128k', {k:d[k]['us'] for k in d if k!='shape'})"
python tools/kernel_breakdown.py --ctx 262144 --budget 8192 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('mha256k', {k:d[k]['us'] for k in d if k!='shape'})"
python tools/seqshard_emulate.py --world 8 --rank 7 2>&1 | tail -6
python tools/rank_emulate.py 2>&1 | tail -4
