"""Times the REFERENCE's own CUDA kernels (oracle/_ref/libquest_ref.so: estimate, sparse decode) on this GPU for the
headline workload (LongChat-7B layer: H=32, D=128, P=16, ctx 32768, budget 2048) next to ours, stream launches,
CUDA events on the default stream (the reference launches on the legacy default stream = torch's default stream).
The reference's raft top-k is not buildable here, so its top-k leg is absent.  Usage: python tools/ref_time.py"""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import quest_b200._kernels as K  # noqa: E402

H, D, P = 32, 128, 16
CTX = int(os.environ.get("CTX", 32768))
BUDGET_TOK = int(os.environ.get("BUDGET", 2048))
REPS = 200
DEV = "cuda"


def p(x):
    return C.c_void_p(x.data_ptr())


def timeit(fn, flush):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(REPS):
        flush.add_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        tot += a.elapsed_time(b)
    return tot / REPS * 1e3


def main():
    ref = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libquest_ref.so"))
    torch.manual_seed(0)
    n_pages = CTX // P
    n_meta = (n_pages + P - 1) // P
    kv = torch.randn(n_pages, 2, P, H, D, device=DEV).half()
    kmax = kv[:, 0].amax(dim=1)
    kmin = kv[:, 0].amin(dim=1)
    meta = torch.zeros(n_meta, 2, P, H, D, dtype=torch.float16, device=DEV)
    pad = torch.zeros(n_meta * P, H, D, dtype=torch.float16, device=DEV)
    pad[:n_pages] = kmax
    meta[:, 0] = pad.view(n_meta, P, H, D)
    pad[:n_pages] = kmin
    meta[:, 1] = pad.view(n_meta, P, H, D)
    kv_idx = torch.randperm(n_pages, device=DEV).to(torch.int32)
    meta_idx = torch.randperm(n_meta, device=DEV).to(torch.int32)
    meta_indptr = torch.tensor([0, n_meta], dtype=torch.int32, device=DEV)
    q = torch.randn(1, H, D, device=DEV).half()
    n = n_pages - 1
    s_ref = torch.zeros(H, n, dtype=torch.float16, device=DEV)
    s_our = torch.zeros(H, n, dtype=torch.float16, device=DEV)
    mlast = (n_pages - 1) % P + 1
    flush = torch.zeros(256 << 20, dtype=torch.uint8, device=DEV)

    def est_ref():
        ref.ref_estimate(p(q), p(s_ref), p(meta), p(meta_idx), p(meta_indptr), mlast, int_meta_last, H, D, P)

    int_meta_last = int(meta_idx[-1])

    def est_our():
        K.estimate_attn_score(q, s_our, meta, meta_idx, meta_indptr, mlast, int_meta_last, 0, K.EST_META_ORDERED)

    t_er, t_eo = timeit(est_ref, flush), timeit(est_our, flush)
    assert torch.equal(s_ref.view(torch.int16), s_our.view(torch.int16))

    n_sel = BUDGET_TOK // P - 1
    sel = torch.stack([torch.randperm(n, device=DEV)[:n_sel].sort().values for _ in range(H)])
    phys = kv_idx[sel].contiguous()
    indptr = torch.tensor([0, n_sel], dtype=torch.int32, device=DEV)
    o_ref = torch.zeros(1, H, D, dtype=torch.float16, device=DEV)
    o_our = torch.zeros(1, H, D, dtype=torch.float16, device=DEV)
    last_idx = int(kv_idx[-1])
    ref.ref_decode_begin(p(indptr), H, H, D, P)

    def dec_ref():
        ref.ref_decode_forward(p(q), p(o_ref), p(kv), p(phys), p(indptr), n_sel, P, last_idx, H, H, D, P)

    w = K.BatchDecodeWithPagedKVCachePyTorchWrapper(0)
    w.begin_forward(indptr, H, H, D, P, torch.empty(0, dtype=torch.float16), num_pages=n_sel)

    def dec_our():
        w.forward(q, o_our, kv, phys, indptr, P, last_idx, 1.0, 1e4)

    t_dr, t_do = timeit(dec_ref, flush), timeit(dec_our, flush)
    err = (o_ref.float() - o_our.float()).abs().max().item()
    ref.ref_decode_end()
    print(f"ctx={CTX} budget={BUDGET_TOK} H={H} D={D} P={P}; stream launches, L2 flushed, us per call (host launch included)")
    print(f"estimate  reference {t_er:8.2f} us   ours {t_eo:8.2f} us   bit-exact")
    print(f"decode    reference {t_dr:8.2f} us   ours {t_do:8.2f} us   max|diff| {err:.2e}")


if __name__ == "__main__":
    main()
