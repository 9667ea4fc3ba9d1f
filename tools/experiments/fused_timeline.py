"""Per-CTA phase timeline of the fused layer-step kernel (csrc/fused.cu built with -DQUEST_FUSED_CLOCKS):
  QUEST_EXTRA_FLAGS=-DQUEST_FUSED_CLOCKS QUEST_LIB_SUFFIX=_dbg bash quest_b200/csrc/build.sh   (here)
  QUEST_B200_LIB=quest_b200/libquest_b200_dbg.so python tools/fused_timeline.py                 (GPU box)
Runs the bench workload's step in a CUDA graph (32 layers), reads the %globaltimer stamps of the LAST layer's launch and
prints, per event, min / median / max over CTAs relative to the earliest CTA start."""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import quest_b200._kernels as K  # noqa: E402

NAMES = {0: "cta start", 1: "idx+wait done", 10: "topk: scores in", 11: "topk: published", 12: "decode: selection seen",
         13: "decode: pages consumed", 14: "decode+merge done", 15: "exit"}
NAMES.update({2: "estimate g0 signed", 3: "estimate g1 signed", 4: "topk: keys staged", 5: "topk: pass 0", 6: "topk: pass 1",
              7: "topk: pass 2", 8: "topk: pass 3", 9: "topk: ids written"})
ORDER = [0, 1, 2, 3, 10, 4, 5, 6, 7, 8, 9, 11, 12, 13, 14, 15]


def main():
    lib = K.load_library()
    dev = torch.device("cuda", 0)
    wl = bench.Workload(dev, 1234)
    step, _ = wl.make_raw_step(parts=("step",))
    step()
    torch.cuda.synchronize()
    g = bench.graph_of(step)
    for _ in range(5):
        g.replay()
    torch.cuda.synchronize()
    n = 2 * torch.cuda.get_device_properties(dev).multi_processor_count
    buf = (ctypes.c_ulonglong * (n * 16))()
    lib.quest_debug_fused_clocks.restype = ctypes.c_int
    assert lib.quest_debug_fused_clocks(buf, n) == 0
    t = np.frombuffer(buf, dtype=np.uint64).reshape(n, 16).astype(np.int64)
    t = t[t[:, 0] > 0]  # CTAs that ran
    last = t[:, 15].max()
    t = np.where(t > last - 1_000_000, t, 0)  # drop stamps left over from other launch shapes
    t0 = t[:, 0][t[:, 0] > 0].min()
    print(f"{n} CTAs; times in us since the median end of the programmatic-launch wait (cta start: since the first CTA); groups={os.environ.get('QUEST_FUSED_GROUPS', 'default')}")
    print(f"{'event':28s} {'n':>4s} {'min':>8s} {'p50':>8s} {'p90':>8s} {'max':>8s}")
    for e in ORDER:
        v = t[:, e]
        v = v[v >= t0]
        if e == 1:
            w0 = np.median(v)
        if v.size == 0:
            continue
        r = (v - (w0 if e >= 1 else t0)) / 1e3
        print(f"{NAMES[e]:28s} {v.size:4d} {r.min():8.2f} {np.median(r):8.2f} {np.percentile(r, 90):8.2f} {r.max():8.2f}")
    np.save(os.path.join(ROOT, "gpurun_out", "fused_clk.npy"), t - t0)


if __name__ == "__main__":
    main()
