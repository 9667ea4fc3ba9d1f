"""The reference's three kernel benchmarks on its own axes, ours beside the reference's kernel on the same B200:

  decode    kernels/src/bench/bench_batch_decode.cu:137-142   seqlen 4096..65536 (+131072) x page_budget 64..512
  estimate  kernels/src/bench/bench_max_possible.cu:94-97     seqlen 2048..131072
  top-k     kernels/src/bench/bench_decode_select_k.cu:52-54  len 128..8192 x k 64 / 128 / 256

32 heads x 128, page 16, fp16 (the benchmarks' own shape).  The reference side is its kernels built for sm_100a
(oracle/_ref/libquest_ref.so: decode + merge, estimate); its top-k is raft, not buildable here (rmm / spdlog absent), so
the top-k column beside ours is torch.topk on the same scores (what quest/tests/test_approx_attention.py:80 uses).
Timing: ours both as a CUDA graph of REPS launches (device time, no host in the loop) and as back-to-back stream
launches; the reference launches on the legacy default stream, which a capture cannot record, so it is timed with
back-to-back stream launches only (as its own benchmarks do).  Inputs rotate over NPOOL sets larger than L2 together.
Usage: python tools/ref_sweeps.py [decode] [estimate] [topk] > profiles/rNN_ref_sweeps.txt"""
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import quest_b200._kernels as K  # noqa: E402

H, D, P = 32, 128, 16
DEV = "cuda"
REPS = 200


def p(x):
    return C.c_void_p(x.data_ptr())


def st():
    return torch.cuda.current_stream().cuda_stream


def timed_stream(fn, reps=REPS):
    """Back-to-back launches from the host; us per call."""
    for i in range(10):
        fn(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps):
        fn(i)
    b.record()
    b.synchronize()
    return a.elapsed_time(b) / reps * 1e3


def timed_graph(fn, npool, reps=REPS):
    """fn(i) captured into one graph of `reps` launches; us per call (device time)."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for i in range(npool):
            fn(i)
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(reps):
            fn(i)
    g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    g.replay()
    b.record()
    b.synchronize()
    return a.elapsed_time(b) / reps * 1e3


def load_ref():
    path = os.path.join(ROOT, "oracle", "_ref", "libquest_ref.so")
    return C.CDLL(path) if os.path.exists(path) else None


def sweep_decode(ref, lib, peak):
    print(f"\n== decode (bench_batch_decode.cu axes): bytes = min(budget, pages) * 262144; peak = {peak:.0f} GB/s")
    print(f"{'seqlen':>7s} {'budget':>6s} {'MB':>6s} | {'ours graph us':>13s} {'GB/s':>6s} {'frac':>5s} | {'ours stream':>11s} | "
          f"{'ref stream us':>13s} {'GB/s':>6s} | {'ref/ours':>8s} {'max|diff|':>9s}")
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    for ctx in (4096, 8192, 16384, 32768, 65536, 131072):
        n_pages = ctx // P
        npool = max(2, min(8, (1 << 30) // (n_pages * 262144)))
        pools = [torch.randn(n_pages, 2, P, H, D, device=DEV, dtype=torch.float16) for _ in range(npool)]
        q = torch.randn(1, H, D, device=DEV).half()
        perm = torch.randperm(n_pages, device=DEV).to(torch.int32)
        for budget in (64, 128, 256, 512):
            b = min(budget, n_pages)  # the reference clamps the budget to the page count (bench_batch_decode.cu:47)
            n_sel = b - 1
            sel = torch.stack([torch.randperm(n_pages - 1, device=DEV)[:n_sel].sort().values for _ in range(H)])
            phys = perm[sel].contiguous()
            indptr = torch.tensor([0, n_sel], dtype=torch.int32, device=DEV)
            o_our = torch.zeros(1, H, D, dtype=torch.float16, device=DEV)
            o_ref = torch.zeros(1, H, D, dtype=torch.float16, device=DEV)
            last_idx = int(perm[-1])
            ppc, nch = C.c_uint32(), C.c_uint32()
            assert lib.quest_decode_plan(n_sel, H, H, D, P, sms, C.byref(ppc), C.byref(nch)) == 0
            ws = torch.zeros(lib.quest_decode_workspace_bytes(H, D, nch.value), dtype=torch.uint8, device=DEV)
            ptrs = [pl.data_ptr() for pl in pools]

            def ours(i):
                rc = lib.quest_decode_run(q.data_ptr(), o_our.data_ptr(), None, ptrs[i % npool], phys.data_ptr(), n_sel, n_sel,
                                          last_idx, P, H, H, D, P, ppc.value, ws.data_ptr(), ws.numel(), st())
                assert rc == 0
            t_g, t_s = timed_graph(ours, npool), timed_stream(ours)
            nbytes = b * 262144
            line = (f"{ctx:7d} {budget:6d} {nbytes / 1e6:6.1f} | {t_g:13.2f} {nbytes / t_g / 1e3:6.0f} "
                    f"{nbytes / t_g / 1e3 / peak:5.2f} | {t_s:11.2f} | ")
            if ref is not None:
                ref.ref_decode_begin(p(indptr), H, H, D, P)
                t_r = timed_stream(lambda i: ref.ref_decode_forward(p(q), p(o_ref), p(pools[i % npool]), p(phys), p(indptr),
                                                                    n_sel, P, last_idx, H, H, D, P))
                torch.cuda.synchronize()
                ref.ref_decode_forward(p(q), p(o_ref), p(pools[0]), p(phys), p(indptr), n_sel, P, last_idx, H, H, D, P)
                ours(0)
                torch.cuda.synchronize()
                ref.ref_decode_end()
                err = (o_ref.float() - o_our.float()).abs().max().item()
                line += f"{t_r:13.2f} {nbytes / t_r / 1e3:6.0f} | {t_r / t_s:8.2f} {err:9.1e}"
                assert err <= 2e-3
            print(line, flush=True)
        del pools


def sweep_estimate(ref, lib, peak):
    print(f"\n== estimate (bench_max_possible.cu axes): bytes = (pages - 1) * 32 * 128 * 2 * 2; peak = {peak:.0f} GB/s")
    print(f"{'seqlen':>7s} {'MB':>6s} | {'ours graph us':>13s} {'GB/s':>6s} {'frac':>5s} | {'ours stream':>11s} | "
          f"{'ref stream us':>13s} {'GB/s':>6s} | {'ref/ours':>8s} bit-exact")
    for ctx in (2048, 4096, 8192, 16384, 32768, 65536, 131072):
        n_pages = ctx // P
        n_meta = (n_pages + P - 1) // P
        n = n_pages - 1
        npool = max(2, min(8, (512 << 20) // (n_meta * 262144)))
        metas = [torch.randn(n_meta, 2, P, H, D, device=DEV, dtype=torch.float16) for _ in range(npool)]
        for m in metas:  # max slab >= min slab, as append_kv leaves it
            hi, lo = torch.maximum(m[:, 0], m[:, 1]), torch.minimum(m[:, 0], m[:, 1])
            m[:, 0], m[:, 1] = hi, lo
        meta_idx = torch.randperm(n_meta, device=DEV).to(torch.int32)
        meta_indptr = torch.tensor([0, n_meta], dtype=torch.int32, device=DEV)
        q = torch.randn(1, H, D, device=DEV).half()
        s_our = torch.zeros(H, n, dtype=torch.float16, device=DEV)
        s_ref = torch.zeros(H, n, dtype=torch.float16, device=DEV)
        mlast, mlast_idx = (n_pages - 1) % P + 1, int(meta_idx[-1])
        ptrs = [m.data_ptr() for m in metas]

        def ours(i):
            rc = lib.quest_estimate(q.data_ptr(), s_our.data_ptr(), n, ptrs[i % npool], meta_idx.data_ptr(), H, H, D, P, 2, st())
            assert rc == 0
        t_g, t_s = timed_graph(ours, npool), timed_stream(ours)
        nbytes = n * H * D * 2 * 2
        line = f"{ctx:7d} {nbytes / 1e6:6.1f} | {t_g:13.2f} {nbytes / t_g / 1e3:6.0f} {nbytes / t_g / 1e3 / peak:5.2f} | {t_s:11.2f} | "
        if ref is not None:
            t_r = timed_stream(lambda i: ref.ref_estimate(p(q), p(s_ref), p(metas[i % npool]), p(meta_idx), p(meta_indptr), mlast,
                                                          mlast_idx, H, D, P))
            torch.cuda.synchronize()
            ref.ref_estimate(p(q), p(s_ref), p(metas[0]), p(meta_idx), p(meta_indptr), mlast, mlast_idx, H, D, P)
            ours(0)
            torch.cuda.synchronize()
            same = torch.equal(s_ref.view(torch.int16), s_our.view(torch.int16))
            line += f"{t_r:13.2f} {nbytes / t_r / 1e3:6.0f} | {t_r / t_s:8.2f} {same}"
            assert same
        print(line, flush=True)
        del metas


def sweep_topk(lib):
    print("\n== top-k (bench_decode_select_k.cu axes): 32 heads; the reference's raft kernel cannot be built here -> torch.topk beside ours")
    print(f"{'len':>6s} {'k':>4s} | {'ours graph us':>13s} {'ours stream':>11s} | {'torch.topk graph us':>19s} {'stream':>8s} | "
          f"{'torch/ours (graph)':>18s} same value multiset")
    for n in (128, 256, 512, 1024, 2048, 4096, 8192):
        for k in (64, 128, 256):
            if k > n:
                continue
            npool = 4
            scores = [(torch.randn(H, n, device=DEV) * 30 + 100).half() for _ in range(npool)]
            idx = torch.arange(n, dtype=torch.int32, device=DEV)
            out_v = torch.zeros(H, k, dtype=torch.float16, device=DEV)
            out_i = torch.zeros(H, k, dtype=torch.int32, device=DEV)
            sp = [s.data_ptr() for s in scores]

            def ours(i):
                rc = lib.quest_topk(sp[i % npool], idx.data_ptr(), 0, out_v.data_ptr(), out_i.data_ptr(), H, n, k, st())
                assert rc == 0
            hold = {}

            def tt(i):
                hold["v"] = torch.topk(scores[i % npool], k, dim=1)
            t_g, t_s = timed_graph(ours, npool), timed_stream(ours)
            t_tg, t_ts = timed_graph(tt, npool), timed_stream(tt)
            ours(0), tt(0)
            torch.cuda.synchronize()
            same = torch.equal(out_v.float().sort(dim=1).values, hold["v"].values.float().sort(dim=1).values)
            print(f"{n:6d} {k:4d} | {t_g:13.2f} {t_s:11.2f} | {t_tg:19.2f} {t_ts:8.2f} | {t_tg / t_g:18.2f} {same}", flush=True)
            assert same


def main():
    which = set(sys.argv[1:]) or {"decode", "estimate", "topk"}
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6650.0
    ref, lib = load_ref(), K.load_library()
    torch.manual_seed(0)
    print(f"H={H} D={D} P={P} fp16 on {torch.cuda.get_device_name(0)}; reference kernels: "
          f"{'oracle/_ref/libquest_ref.so (sm_100a build of kernels/include)' if ref is not None else 'not available'}")
    if "decode" in which:
        sweep_decode(ref, lib, peak)
    if "estimate" in which:
        sweep_estimate(ref, lib, peak)
    if "topk" in which:
        sweep_topk(lib)


if __name__ == "__main__":
    main()
