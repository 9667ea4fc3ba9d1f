#!/usr/bin/env python
"""Times sub-chains of the layer step (CUDA graph over 32 layers) to see how kernels interact under PDL."""
import itertools, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench

torch.cuda.set_device(0)
wl = bench.Workload(torch.device("cuda", 0), 1234)
combos = [("append",), ("estimate",), ("topk",), ("decode",), ("append", "estimate"), ("estimate", "topk"), ("topk", "decode"),
          ("append", "decode"), ("estimate", "topk", "decode"), ("append", "estimate", "topk"), ("estimate", "decode"),
          ("append", "estimate", "topk", "decode")]
for parts in combos:
    st, n = wl.make_raw_step(parts=parts)
    st(); torch.cuda.synchronize()
    g = bench.graph_of(st)
    t = bench.timed(g.replay, 30, 3, False) / 30 / bench.L
    print(f"{'+'.join(parts):40s} {t * 1e6:8.2f} us/layer", flush=True)
