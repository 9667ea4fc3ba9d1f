#!/bin/bash
# s59: A/B 2-CTA cluster vs single CTA at 2049..4096-page rows
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "topk" 2>&1 | tail -2
for dbg in "" "4=1"; do
  echo "QUEST_DBG=$dbg"
  QUEST_DBG=$dbg python tools/kernel_breakdown.py --ctx 65536 --budget 4096 --steps 20 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('mha64k', {k:d[k]['us'] for k in d if k!='shape'})"
  QUEST_DBG=$dbg python tools/kernel_breakdown.py --ctx 49152 --budget 2048 --steps 20 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('mha48k', {k:d[k]['us'] for k in d if k!='shape'})"
done
python tools/seqshard_emulate.py --world 8 --rank 7 2>&1 | tail -6
