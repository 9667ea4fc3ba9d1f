#!/usr/bin/env python
"""All ranks of BASELINE configs[4] (sequence-sharded step, 8192 pages per rank, budget 8192) played by ONE process on one GPU with
the exchanges stood in for by copies, so that one rank's kernels can be listed under ncu without an 8-GPU box:
    ncu --metrics gpu__time_duration.sum -k regex:'estimate|topk|decode|merge' python tools/seqshard_emulate.py --world 8 --rank 7
Prints the graph-replayed time of the profiled rank's compute phases as well."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import quest_b200.parallel as par  # noqa: E402
import quest_b200.utils as qu  # noqa: E402
from tools.bench_configs import fill, timed, try_graph  # noqa: E402

P, D, H = 16, 128, 32


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--rank", type=int, default=0, help="the rank whose phases are timed / profiled")
    ap.add_argument("--pages", type=int, default=8192)
    ap.add_argument("--budget", type=int, default=8192, help="token budget")
    ap.add_argument("--steps", type=int, default=20)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    R, budget = a.world, a.budget // P
    ics, shards = [], []
    for r in range(R):
        tokens = a.pages * P - (1 if r == R - 1 else 0)
        torch.manual_seed(1234 + 5 + r)
        ic = qu.InferenceController(1, H, D, P, budget, a.pages * P, torch.float16, dev)
        fill(ic, 1, tokens, H, dev)
        if r == R - 1:
            ic.prepare_metadata(1)
        ics.append(ic)
        shards.append(par.SeqShardedDecode(ic, H, R, r, budget))
    n_scored = [a.pages] * (R - 1) + [a.pages - 1]
    for s in shards:
        s.begin_step(n_scored)
    torch.manual_seed(77)
    q = torch.randn(1, H, D, device=dev).half()
    kn, vn = torch.randn(1, H, D, device=dev).half(), torch.randn(1, H, D, device=dev).half()
    gath = torch.stack([s.candidates(q, kn, vn, 0) for s in shards]).clone()
    for s in shards:
        s.select(gath)
    recs = torch.stack([s.attend(q, 0) for s in shards]).clone()
    me = shards[a.rank]
    torch.cuda.synchronize()

    def mine():  # the profiled rank's compute phases, messages of the other ranks already in place
        me.candidates(q, kn, vn, 0)
        me.select(gath)
        me.attend(q, 0)
        me.merge(recs)

    own = int((me.sel[:, :me.k] >= 0).sum().item())
    print(f"rank {a.rank} of {R}: {own / H:.1f} of {me.k} selected pages per head are local")
    for name, fn in (("candidates", lambda: me.candidates(q, kn, vn, 0)), ("select", lambda: me.select(gath)),
                     ("attend", lambda: me.attend(q, 0)), ("merge", lambda: me.merge(recs)), ("all four", mine)):
        fn()
        torch.cuda.synchronize()
        g = try_graph(fn)
        us = timed(g if g is not None else fn, a.steps, 5, False) * 1e6
        print(f"  {name:11s} {us:7.2f} us {'(graph)' if g is not None else '(eager)'}")


if __name__ == "__main__":
    main()
