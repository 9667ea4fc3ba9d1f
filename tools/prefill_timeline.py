"""Per-iteration timeline of one CTA of the TMA prefill kernel (instrumented build:
QUEST_EXTRA_FLAGS=-DQUEST_PF_CLOCKS QUEST_LIB_SUFFIX=_dbg bash quest_b200/csrc/build.sh).  SM clock cycles relative to
the iteration's first stamp.  Producer lane: 0 S(j) done seen, 1 PV(j-1) done seen, 2 before P wait, 3 P(j) seen,
4 V(j) seen, 5 PV(j) issued, 6 K(j+2) seen, 7 S(j+2) issued.  Softmax thread 0: 8 before S wait, 9 S(j) seen,
10 math done, 11 PV(j-1) seen, 12 P(j) handed over."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ["QUEST_B200_LIB"] = os.path.join(ROOT, "quest_b200", "libquest_b200_dbg.so")
sys.path.insert(0, ROOT)
import quest_b200._kernels as K  # noqa: E402

H, D, P = 32, 128, 16
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
lib = K.load_library()
pool = torch.randn(n // P, 2, P, H, D, device="cuda", dtype=torch.float16)
q = torch.randn(n, H, D, device="cuda", dtype=torch.float16)
o = torch.zeros_like(q)
table = torch.randperm(n // P, device="cuda").to(torch.int32)
os.environ["QUEST_PF_IMPL"] = "tma"
for _ in range(3):
    assert lib.quest_prefill(q.data_ptr(), o.data_ptr(), n, pool.data_ptr(), table.data_ptr(), n // P, P, 1, H, H, D, P, 0) == 0
torch.cuda.synchronize()
buf = np.zeros((64, 16), dtype=np.uint64)
assert lib.quest_debug_prefill_clocks(buf.ctypes.data_as(C.c_void_p)) == 0
base = buf[:, 8].astype(np.int64)
print("iter | period | producer: Sdone PVdone preP Pseen Vseen PVissued Kseen Sissued | softmax: S-wait-start Sseen math-done PVseen handed")
for j in range(2, 40):
    r = buf[j].astype(np.int64)
    if r[8] == 0:
        break
    rel = [int(x - r[8]) if x else None for x in r[:13]]
    period = int(r[8] - buf[j - 1][8].astype(np.int64))
    print(f"{j:4d} | {period:6d} | " + " ".join(f"{str(x):>7s}" for x in rel[:8]) + " | " + " ".join(f"{str(x):>7s}" for x in rel[8:13]))
