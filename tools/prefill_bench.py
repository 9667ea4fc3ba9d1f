"""causal prefill over the paged KV cache, LongChat-7B shape (32 heads x 128, page 16, fp16),
one layer: our tcgen05 kernels (TMA-fed warp-specialised, and the cp.async-fed v1), our SIMT kernel (short lengths only) and the reference's own kernel
(oracle/_ref/libquest_ref.so: kernels/include/prefill/prefill.cuh built for sm_100a, mma.sync) on the same pool and page
table.  FLOPs counted as the causal attention does them: 4 * D * H * sum_i (kv_len - qo_len + i + 1).  CUDA events around
REPS back-to-back launches after warm-up; the tensor roofline denominator is MEASURED_PEAKS.json's bf16_tflops.
Usage: python tools/prefill_bench.py [len ...] > profiles/rNN_prefill.txt"""
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


def p(x):
    return C.c_void_p(x.data_ptr())


def timed(fn, reps):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    b.synchronize()
    return a.elapsed_time(b) / reps  # ms


def main():
    lens = [int(x) for x in sys.argv[1:]] or [2048, 8192, 32768]
    peak = 2250.0
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]
    except Exception:
        pass
    ref_path = os.path.join(ROOT, "oracle", "_ref", "libquest_ref.so")
    ref = C.CDLL(ref_path) if os.path.exists(ref_path) else None
    if ref is not None and not hasattr(ref, "ref_prefill"):
        ref = None
    only = os.environ.get("PF_ONLY") == "tma"  # profiling runs: just our TMA tensor-core kernel
    if only:
        ref = None
    lib = K.load_library()
    st = torch.cuda.current_stream().cuda_stream
    torch.manual_seed(0)
    print(f"causal prefill, one layer, H={H} D={D} P={P} fp16, qo_len = kv_len = n; ms per launch; "
          f"tensor peak = {peak:.0f} TFLOP/s (MEASURED_PEAKS.json bf16)")
    print(f"{'n':>6s} {'GFLOP':>9s} | {'tma ms':>10s} {'TFLOP/s':>8s} {'frac':>5s} | {'cp.async':>8s} {'TFLOP/s':>8s} | {'simt ms':>8s} {'TFLOP/s':>8s} | "
          f"{'ref ms':>8s} {'TFLOP/s':>8s} | {'speed-up':>8s} {'max|diff|':>9s}")
    for n in lens:
        n_pages = n // P
        pool = torch.randn(n_pages, 2, P, H, D, device=DEV, dtype=torch.float16)
        q = torch.randn(n, H, D, device=DEV, dtype=torch.float16)
        table = torch.randperm(n_pages, device=DEV).to(torch.int32)
        o = {k: torch.zeros(n, H, D, device=DEV, dtype=torch.float16) for k in ("tma", "umma", "simt", "ref")}
        flops = 4.0 * D * H * (n * (n + 1) / 2)
        reps = max(2, min(50, int(2e13 / flops)))

        def ours(impl):
            lib.quest_debug_set_option(K.DBG_PREFILL_IMPL, {"simt": 1, "umma": 2, "tma": 3}[impl])
            rc = lib.quest_prefill(q.data_ptr(), o[impl].data_ptr(), n, pool.data_ptr(), table.data_ptr(), n_pages, P, 1, H, H, D, P, st)
            assert rc == 0, rc

        t_u = timed(lambda: ours("tma"), reps)
        assert lib.quest_debug_prefill_timeouts() == 0
        line = f"{n:6d} {flops / 1e9:9.1f} | {t_u:10.3f} {flops / t_u / 1e9:8.1f} {flops / t_u / 1e9 / peak:5.2f} | "
        if not only:
            t_c = timed(lambda: ours("umma"), reps)
            line += f"{t_c:8.3f} {flops / t_c / 1e9:8.1f} | "
        else:
            line += f"{'-':>8s} {'-':>8s} | "
        if n <= 8192 and not only:
            t_s = timed(lambda: ours("simt"), 2)
            line += f"{t_s:8.2f} {flops / t_s / 1e9:8.1f} | "
        else:
            line += f"{'-':>8s} {'-':>8s} | "
        lib.quest_debug_set_option(K.DBG_PREFILL_IMPL, 0)
        assert lib.quest_debug_prefill_timeouts() == 0 and lib.quest_device_error() == 0
        if ref is not None:
            qo_indptr = torch.tensor([0, n], dtype=torch.int32, device=DEV)
            kv_indptr = torch.tensor([0, n_pages], dtype=torch.int32, device=DEV)
            last_idx = int(table[-1])

            def reff():
                rc = ref.ref_prefill(p(q), p(o["ref"]), n, p(qo_indptr), p(pool), p(table), p(kv_indptr), P, last_idx, H, H, D, P)
                assert rc == 0, rc
            t_r = timed(reff, max(2, reps // 4))
            torch.cuda.synchronize()
            err = (o["ref"].float() - o["tma"].float()).abs().max().item()
            line += f"{t_r:8.3f} {flops / t_r / 1e9:8.1f} | {t_r / t_u:8.2f} {err:9.1e}"
            # a throughput number only counts with parity: every timed size is checked against the reference kernel
            assert err <= 2e-3, f"n={n}: tma kernel differs from the reference prefill kernel by {err}"
            if not only:
                err_c = (o["ref"].float() - o["umma"].float()).abs().max().item()
                assert err_c <= 2e-3, f"n={n}: cp.async tcgen05 kernel differs from the reference prefill kernel by {err_c}"
        else:
            line += f"{'n/a':>8s} {'n/a':>8s} | {'n/a':>8s} {'n/a':>9s}"
        print(line, flush=True)
        del pool, q, o


if __name__ == "__main__":
    main()
