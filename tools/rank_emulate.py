#!/usr/bin/env python
"""One rank's share of BASELINE configs[3] (kv-head sharded over `--world` ranks, no collective) on a single GPU: what a rank of the
8-GPU run executes, for per-kernel breakdowns under ncu without an 8-GPU box.
    python tools/rank_emulate.py --world 8 [--steps 20]"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tools import bench_configs as bc  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    bc.apply_debug_env()
    r = bc.measure(4, 0, a.world, dev, layers=2, streams=0, batched=1, steps=a.steps, warmup=a.warmup)
    print(json.dumps({k: r[k] for k in ("us_per_layer", "eager_us_per_layer", "graph_us_per_layer", "parallelism") if k in r}))


if __name__ == "__main__":
    main()
