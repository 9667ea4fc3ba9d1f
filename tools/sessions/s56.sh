#!/bin/bash
# s56: quest_topk on self-counted bin totals: parity, stand-alone timings, the sequence-sharded step's global selection
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "topk" 2>&1 | tail -3
python -m pytest tests/test_seq_shard_gpu.py tests/test_reference_python_on_shim.py -x -q -m gpu 2>&1 | tail -2
python tools/kernel_breakdown.py --ctx 32768 --budget 2048 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('mha32k', {k:d[k]['us'] for k in d if k!='shape'})"
python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 131072 --budget 4096 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('gqa128k', {k:d[k]['us'] for k in d if k!='shape'})"
python tools/kernel_breakdown.py --ctx 65536 --budget 4096 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('mha64k', {k:d[k]['us'] for k in d if k!='shape'})"
python tools/seqshard_emulate.py --world 8 --rank 7 2>&1 | tail -8
