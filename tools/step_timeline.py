"""Per-CTA phase timeline of one layer step (estimate -> top-k -> decode) inside the bench graph:
  QUEST_EXTRA_FLAGS="-DQUEST_DEC_CLOCKS -DQUEST_EST_CLOCKS -DQUEST_TOPK_HIST_CLOCKS" QUEST_LIB_SUFFIX=_dbg bash quest_b200/csrc/build.sh   (here)
  QUEST_B200_LIB=quest_b200/libquest_b200_dbg.so python tools/step_timeline.py                                     (GPU box)
Replays the bench workload's 32-layer graph and reads the %globaltimer stamps of the LAST layer's launches (the stamps of
earlier layers are overwritten), all on one clock: us relative to the median start of the estimate CTAs."""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import quest_b200._kernels as K  # noqa: E402

EST = ["est: cta start", "est: first ring issued", "est: launch wait done", "est: pages scored"]
DEC = ["dec: cta start", "dec: launch wait done", "dec: ids + q in", "dec: warp 0 rows consumed", "dec: partial out, ticket",
       "dec: head merged"]


def main():
    lib = K.load_library()
    dev = torch.device("cuda", 0)
    wl = bench.Workload(dev, 1234)
    step, _ = wl.make_raw_step(parts=tuple(os.environ.get("PARTS", "layer").split(",")))
    step()
    torch.cuda.synchronize()
    g = bench.graph_of(step)
    for _ in range(5):
        g.replay()
    torch.cuda.synchronize()
    ne, nd = 297, 1024
    be, bd = (ctypes.c_ulonglong * (ne * 4))(), (ctypes.c_ulonglong * (nd * 8))()
    lib.quest_debug_estimate_clocks.restype = lib.quest_debug_decode_clocks.restype = ctypes.c_int
    assert lib.quest_debug_estimate_clocks(be, ne) == 0 and lib.quest_debug_decode_clocks(bd, nd) == 0
    te = np.frombuffer(be, dtype=np.uint64).reshape(ne, 4).astype(np.int64)[:296]
    td = np.frombuffer(bd, dtype=np.uint64).reshape(nd, 8).astype(np.int64)
    tt = None
    if hasattr(lib, "quest_debug_topk_hist_clocks"):
        bt = (ctypes.c_ulonglong * (32 * 4))()
        lib.quest_debug_topk_hist_clocks.restype = ctypes.c_int
        assert lib.quest_debug_topk_hist_clocks(bt, 32) == 0
        tt = np.frombuffer(bt, dtype=np.uint64).reshape(32, 4).astype(np.int64)
    t0 = np.median(te[:, 0])
    print(f"layer step timeline (last layer of the graph), us relative to the median estimate CTA start")
    print(f"{'event':30s} {'n':>5s} {'min':>8s} {'p10':>8s} {'p50':>8s} {'p90':>8s} {'max':>8s}")
    TOPK = ["topk: cta start", "topk: launch wait done", "topk: ids written"]
    for names, t in ((EST, te), (TOPK, tt), (DEC, td)):
        if t is None:
            continue
        order = range(len(names))
        for e in order:
            v = t[:, e]
            v = v[(v > t0 - 100_000) & (v < t0 + 1_000_000)]
            if v.size:
                r = (v - t0) / 1e3
                print(f"{names[e]:30s} {v.size:5d} {r.min():8.2f} {np.percentile(r, 10):8.2f} {np.median(r):8.2f} {np.percentile(r, 90):8.2f} {r.max():8.2f}")


    # estimate CTAs by blockIdx decile: are the late starters the high block indices?
    print("estimate CTAs by blockIdx decile: mean start / mean finish (us)")
    for d in range(10):
        sl = te[d * 296 // 10:(d + 1) * 296 // 10]
        print(f"  blocks {d * 296 // 10:3d}-{(d + 1) * 296 // 10 - 1:3d}: start {(sl[:, 0].mean() - t0) / 1e3:6.2f}  finish {(sl[:, 3].mean() - t0) / 1e3:6.2f}"
              f"  (max finish {(sl[:, 3].max() - t0) / 1e3:6.2f})")


if __name__ == "__main__":
    main()
