"""Per-CTA phase timeline of decode_kernel inside the layer step (csrc/decode.cu built with -DQUEST_DEC_CLOCKS):
  QUEST_EXTRA_FLAGS=-DQUEST_DEC_CLOCKS QUEST_LIB_SUFFIX=_dbg bash quest_b200/csrc/build.sh   (here)
  QUEST_B200_LIB=quest_b200/libquest_b200_dbg.so python tools/decode_timeline.py             (GPU box)
Replays the bench workload's 32-layer graph and reads the %globaltimer stamps of the last layer's decode launch."""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import quest_b200._kernels as K  # noqa: E402

NAMES = ["cta start", "launch wait done", "ids + q in", "warp 0 rows consumed", "partial out, ticket taken", "head merged (32 CTAs)"]


def main():
    lib = K.load_library()
    dev = torch.device("cuda", 0)
    wl = bench.Workload(dev, 1234)
    parts = tuple(os.environ.get("PARTS", "layer").split(","))
    step, _ = wl.make_raw_step(parts=parts)
    step()
    torch.cuda.synchronize()
    g = bench.graph_of(step)
    for _ in range(5):
        g.replay()
    torch.cuda.synchronize()
    n = 1024
    buf = (ctypes.c_ulonglong * (n * 8))()
    lib.quest_debug_decode_clocks.restype = ctypes.c_int
    assert lib.quest_debug_decode_clocks(buf, n) == 0
    t = np.frombuffer(buf, dtype=np.uint64).reshape(n, 8).astype(np.int64)
    w0 = np.median(t[:, 1])
    print(f"decode_kernel, {n} CTAs, parts={parts}; us relative to the median end of the launch wait")
    print(f"{'event':28s} {'n':>5s} {'min':>8s} {'p50':>8s} {'p90':>8s} {'max':>8s}")
    for e, name in enumerate(NAMES):
        v = t[:, e]
        v = v[v > w0 - 1_000_000]
        if v.size:
            r = (v - w0) / 1e3
            print(f"{name:28s} {v.size:5d} {r.min():8.2f} {np.median(r):8.2f} {np.percentile(r, 90):8.2f} {r.max():8.2f}")


if __name__ == "__main__":
    main()
