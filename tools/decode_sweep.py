"""BASELINE config 3: the decode kernel alone over page budgets (and two context lengths), ours against the reference's
own kernel (oracle/_ref/libquest_ref.so, sm_100a build of kernels/include/decode), as the reference's
bench_batch_decode.cu:53-86 does: 32 heads x 128, page 16, random distinct page subsets per head, bytes counted as
budget * 262,144.  Both sides are launched back to back from the host (no graph: the reference launches on the legacy
default stream), rotating over NPOOL KV pools so that the 126 MB L2 cannot serve the pages; CUDA events around REPS
launches.  Usage: python tools/decode_sweep.py > profiles/rNN_decode_sweep.txt"""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import quest_b200._kernels as K  # noqa: E402

H, D, P = 32, 128, 16
NPOOL, REPS = 6, 240
DEV = "cuda"


def p(x):
    return C.c_void_p(x.data_ptr())


def timed(fn):
    for i in range(12):
        fn(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(REPS):
        fn(i)
    b.record()
    b.synchronize()
    return a.elapsed_time(b) / REPS * 1e3


def main():
    ref_path = os.path.join(ROOT, "oracle", "_ref", "libquest_ref.so")
    ref = C.CDLL(ref_path) if os.path.exists(ref_path) else None
    torch.manual_seed(0)
    print(f"decode kernel only, H={H} D={D} P={P}, fp16; us per launch (host-launched back to back, {NPOOL} pools rotated), "
          f"GB/s = budget*262144 B / time")
    print(f"{'ctx':>7s} {'budget':>6s} {'MB':>7s} | {'ours us':>8s} {'GB/s':>7s} | {'ref us':>8s} {'GB/s':>7s} | {'speed-up':>8s} {'max|diff|':>9s}")
    for ctx in (32768, 131072):
        n_pages = ctx // P
        pools = [torch.randn(n_pages, 2, P, H, D, device=DEV, dtype=torch.float16) for _ in range(NPOOL)]
        q = torch.randn(1, H, D, device=DEV).half()
        perm = torch.randperm(n_pages, device=DEV).to(torch.int32)
        for budget in (64, 128, 256, 512):
            n_sel = budget - 1
            sel = torch.stack([torch.randperm(n_pages - 1, device=DEV)[:n_sel].sort().values for _ in range(H)])
            phys = perm[sel].contiguous()
            indptr = torch.tensor([0, n_sel], dtype=torch.int32, device=DEV)
            o_our = torch.zeros(1, H, D, dtype=torch.float16, device=DEV)
            o_ref = torch.zeros(1, H, D, dtype=torch.float16, device=DEV)
            last_idx = int(perm[-1])
            # both sides through raw ctypes calls with precomputed arguments (no Python wrapper on either)
            lib = K.load_library()
            ppc, nch = C.c_uint32(), C.c_uint32()
            assert lib.quest_decode_plan(n_sel, H, H, D, P, torch.cuda.get_device_properties(0).multi_processor_count,
                                         C.byref(ppc), C.byref(nch)) == 0
            ws = torch.zeros(lib.quest_decode_workspace_bytes(H, D, nch.value), dtype=torch.uint8, device=DEV)
            st = torch.cuda.current_stream().cuda_stream
            ptrs = [pl.data_ptr() for pl in pools]

            def ours(i):
                lib.quest_decode_run(q.data_ptr(), o_our.data_ptr(), None, ptrs[i % NPOOL], phys.data_ptr(), n_sel, n_sel, last_idx,
                                     P, H, H, D, P, ppc.value, ws.data_ptr(), ws.numel(), st)
            t_our = timed(ours)
            nbytes = budget * 262144
            line = f"{ctx:7d} {budget:6d} {nbytes / 1e6:7.1f} | {t_our:8.2f} {nbytes / t_our / 1e3:7.0f} | "
            if ref is not None:
                ref.ref_decode_begin(p(indptr), H, H, D, P)
                t_ref = timed(lambda i: ref.ref_decode_forward(p(q), p(o_ref), p(pools[i % NPOOL]), p(phys), p(indptr), n_sel, P,
                                                               last_idx, H, H, D, P))
                torch.cuda.synchronize()
                ref.ref_decode_forward(p(q), p(o_ref), p(pools[0]), p(phys), p(indptr), n_sel, P, last_idx, H, H, D, P)
                ours(0)
                torch.cuda.synchronize()
                ref.ref_decode_end()
                err = (o_ref.float() - o_our.float()).abs().max().item()
                line += f"{t_ref:8.2f} {nbytes / t_ref / 1e3:7.0f} | {t_ref / t_our:8.2f} {err:9.1e}"
            else:
                line += "     n/a     n/a |      n/a       n/a"
            print(line, flush=True)
        del pools


if __name__ == "__main__":
    main()
