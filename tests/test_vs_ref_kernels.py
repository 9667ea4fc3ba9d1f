"""GPU parity against the REFERENCE's own CUDA kernels (append + metadata, estimate, sparse decode + merge, prefill) compiled
for sm_100a from /root/reference by oracle/ref_build/build_ref.sh into oracle/_ref/libquest_ref.so (built in the
authoring container by __graft_entry__.build(); the .so travels to the GPU box, the reference sources do not).

Same seeded inputs go through both libraries' C entry points:
  - KV pool + per-page min/max metadata: bit-exact
  - estimate scores (fp16): bit-exact
  - sparse decode attention output: max abs err <= 2e-3 (the reference stores fp16 partials, ours fp32)
  - causal prefill attention output: max abs err <= 2e-3
The reference's raft top-k is not buildable here; top-k parity is covered in test_gpu_parity.py.
"""
import ctypes as C
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libquest_ref.so")
DEV = "cuda"
ATTN_ATOL = 2e-3


@pytest.fixture(scope="module")
def ref():
    if not os.path.exists(REF_SO):
        pytest.skip("oracle/_ref/libquest_ref.so not built (needs /root/reference at build time)")
    return C.CDLL(REF_SO)


def p(x):
    return C.c_void_p(x.data_ptr())


class Pools:
    """Two identical cache states, one filled by the reference kernels and one by ours."""

    def __init__(self, H, D, P, kv_len, seed):
        import quest_b200._kernels as K
        self.K, self.H, self.D, self.P = K, H, D, P
        g = torch.Generator(device="cpu").manual_seed(seed)
        n_pages = (kv_len + P - 1) // P
        n_meta = (n_pages + P - 1) // P
        self.kv_idx = torch.randperm(n_pages + 2, generator=g)[:n_pages].to(torch.int32).to(DEV)
        self.meta_idx = torch.randperm(n_meta + 1, generator=g)[:n_meta].to(torch.int32).to(DEV)
        self.kv_indptr = torch.tensor([0, n_pages], dtype=torch.int32, device=DEV)
        self.meta_indptr = torch.tensor([0, n_meta], dtype=torch.int32, device=DEV)
        self.kv_last_len = (kv_len - 1) % P + 1
        self.meta_last_len = (n_pages - 1) % P + 1
        self.k = torch.randn(kv_len, H, D, generator=g).half().to(DEV)
        self.v = torch.randn(kv_len, H, D, generator=g).half().to(DEV)
        self.q = torch.randn(1, H, D, generator=g).half().to(DEV)
        self.n_pages, self.n_meta, self.kv_len = n_pages, n_meta, kv_len
        shape_kv, shape_m = (n_pages + 2, 2, P, H, D), (n_meta + 1, 2, P, H, D)
        self.kv_ref = torch.zeros(shape_kv, dtype=torch.float16, device=DEV)
        self.kv_our = torch.zeros(shape_kv, dtype=torch.float16, device=DEV)
        self.m_ref = torch.zeros(shape_m, dtype=torch.float16, device=DEV)
        self.m_our = torch.zeros(shape_m, dtype=torch.float16, device=DEV)

    def fill(self, ref):
        """prefill of kv_len-1 tokens, then one decode append (quest/tests/test_estimate.py:117-130)."""
        H, D, P, K = self.H, self.D, self.P, self.K
        n1 = self.kv_len - 1
        pages1 = (n1 + P - 1) // P
        meta1 = (pages1 + P - 1) // P
        last1, mlast1 = (n1 - 1) % P + 1, (pages1 - 1) % P + 1
        kvp1 = torch.tensor([0, pages1], dtype=torch.int32, device=DEV)
        mp1 = torch.tensor([0, meta1], dtype=torch.int32, device=DEV)
        app = torch.tensor([0, n1], dtype=torch.int32, device=DEV)
        k1, v1 = self.k[:n1].contiguous(), self.v[:n1].contiguous()
        torch.cuda.synchronize()
        rc = ref.ref_append_prefill(p(k1), p(v1), p(app), p(self.kv_ref), p(self.kv_idx), p(kvp1), last1,
                                    int(self.kv_idx[pages1 - 1]), p(self.m_ref), p(self.meta_idx), p(mp1), mlast1,
                                    int(self.meta_idx[meta1 - 1]), H, D, P)
        assert rc == 0
        K.append_kv_cache_prefill(k1, v1, self.kv_our, self.kv_idx[:pages1].contiguous(), kvp1, last1, int(self.kv_idx[pages1 - 1]),
                                  self.m_our, self.meta_idx[:meta1].contiguous(), mp1, mlast1, int(self.meta_idx[meta1 - 1]), 0)
        k2, v2 = self.k[n1:].contiguous(), self.v[n1:].contiguous()
        torch.cuda.synchronize()
        rc = ref.ref_append_decode(p(k2), p(v2), p(self.kv_ref), p(self.kv_idx), p(self.kv_indptr), self.kv_last_len,
                                   int(self.kv_idx[-1]), p(self.m_ref), p(self.meta_idx), p(self.meta_indptr),
                                   self.meta_last_len, int(self.meta_idx[-1]), H, D, P)
        assert rc == 0
        K.append_kv_cache_decode(k2, v2, self.kv_our, self.kv_idx, self.kv_indptr, self.kv_last_len, int(self.kv_idx[-1]),
                                 self.m_our, self.meta_idx, self.meta_indptr, self.meta_last_len, int(self.meta_idx[-1]), 0)
        torch.cuda.synchronize()


@pytest.mark.parametrize("kv_len", [33, 257, 1000, 2047, 4099])
def test_append_and_metadata_bitexact_vs_reference_kernels(ref, kv_len):
    s = Pools(32, 128, 16, kv_len, seed=kv_len)
    s.fill(ref)
    assert torch.equal(s.kv_ref.view(torch.int16), s.kv_our.view(torch.int16))
    # metadata entries of existing pages only (both sides leave the rest untouched = zeros here)
    assert torch.equal(s.m_ref.view(torch.int16), s.m_our.view(torch.int16))


@pytest.mark.parametrize("kv_len", [65, 300, 1111, 2047, 8000, 32768])
def test_estimate_bitexact_vs_reference_kernel(ref, kv_len):
    H, D, P = 32, 128, 16
    s = Pools(H, D, P, kv_len, seed=kv_len + 1)
    s.fill(ref)
    n = s.n_pages - 1
    o_ref = torch.zeros((H, n), dtype=torch.float16, device=DEV)
    o_our = torch.zeros((H, n), dtype=torch.float16, device=DEV)
    torch.cuda.synchronize()
    assert ref.ref_estimate(p(s.q), p(o_ref), p(s.m_ref), p(s.meta_idx), p(s.meta_indptr), s.meta_last_len,
                            int(s.meta_idx[-1]), H, D, P) == 0
    for flags in (0, s.K.EST_META_ORDERED):
        s.K.estimate_attn_score(s.q, o_our, s.m_our, s.meta_idx, s.meta_indptr, s.meta_last_len, int(s.meta_idx[-1]), 0, flags)
        torch.cuda.synchronize()
        assert torch.equal(o_ref.view(torch.int16), o_our.view(torch.int16)), f"flags={flags}"


@pytest.mark.parametrize("kv_len,budget", [(300, 8), (2047, 32), (4099, 128), (32768, 128)])
def test_sparse_decode_vs_reference_kernel(ref, kv_len, budget):
    H, D, P = 32, 128, 16
    s = Pools(H, D, P, kv_len, seed=kv_len + budget)
    s.fill(ref)
    n_sel = min(budget, s.n_pages) - 1
    g = torch.Generator(device="cpu").manual_seed(budget)
    # random distinct per-head page subsets of the first n_pages-1 pages (test_batch_decode.cu:61-76), physical ids
    logical = torch.stack([torch.randperm(s.n_pages - 1, generator=g)[:n_sel].sort().values for _ in range(H)])
    phys = s.kv_idx.cpu()[logical].to(torch.int32).to(DEV).contiguous()
    indptr = torch.tensor([0, n_sel], dtype=torch.int32, device=DEV)
    o_ref = torch.zeros((1, H, D), dtype=torch.float16, device=DEV)
    o_our = torch.zeros((1, H, D), dtype=torch.float16, device=DEV)
    torch.cuda.synchronize()
    assert ref.ref_decode_begin(p(indptr), H, H, D, P) == 0
    assert ref.ref_decode_forward(p(s.q), p(o_ref), p(s.kv_ref), p(phys), p(indptr), n_sel, s.kv_last_len, int(s.kv_idx[-1]),
                                  H, H, D, P) == 0
    torch.cuda.synchronize()
    ref.ref_decode_end()
    w = s.K.BatchDecodeWithPagedKVCachePyTorchWrapper(0)
    w.begin_forward(indptr, H, H, D, P, torch.empty(0, dtype=torch.float16))
    w.forward(s.q, o_our, s.kv_our, phys, indptr, s.kv_last_len, int(s.kv_idx[-1]), 1.0, 1e4)
    w.end_forward()
    torch.cuda.synchronize()
    assert torch.isfinite(o_our).all() and torch.isfinite(o_ref).all()
    assert (o_ref.float() - o_our.float()).abs().max().item() <= ATTN_ATOL


@pytest.mark.parametrize("qo_len,kv_len", [(200, 200), (1000, 1000), (333, 2100), (4096, 4096), (8192, 8192), (32768, 32768)])
def test_prefill_vs_reference_kernel(ref, qo_len, kv_len):
    """Causal prefill over the paged cache: our tcgen05 kernels against the reference's BatchPrefillWithPagedKVCache
    kernel (prefill.cuh:674-869, built for sm_100a in its LongChat-7B configuration: 32 MHA heads x 128, page 16)."""
    if not hasattr(ref, "ref_prefill"):
        pytest.skip("libquest_ref.so predates ref_prefill")
    import quest_b200._kernels as K
    H, D, P = 32, 128, 16
    g = torch.Generator(device="cpu").manual_seed(qo_len + kv_len)
    n_pages = (kv_len + P - 1) // P
    pool = torch.randn(n_pages + 3, 2, P, H, D, generator=g).half().to(DEV)
    table = torch.randperm(n_pages + 3, generator=g)[:n_pages].to(torch.int32).to(DEV)
    q = torch.randn(qo_len, H, D, generator=g).half().to(DEV)
    last_len = (kv_len - 1) % P + 1
    o_ref = torch.zeros_like(q)
    qo_indptr = torch.tensor([0, qo_len], dtype=torch.int32, device=DEV)
    kv_indptr = torch.tensor([0, n_pages], dtype=torch.int32, device=DEV)
    torch.cuda.synchronize()
    assert ref.ref_prefill(p(q), p(o_ref), qo_len, p(qo_indptr), p(pool), p(table), p(kv_indptr), last_len, int(table[-1]),
                           H, H, D, P) == 0
    torch.cuda.synchronize()
    from tests.helpers import prefill_impl
    for impl in ("", "tma", "tma_smem", "umma"):
        with prefill_impl(impl):
            for it in range(3 if qo_len >= 4096 else 1):
                o = K.prefill_with_paged_kv_cache(q, pool, table, last_len, True, 0, False, 1.0, 1e4)
                torch.cuda.synchronize()
                assert K.load_library().quest_debug_prefill_timeouts() == 0 and K.load_library().quest_device_error() == 0
                assert (o.float() - o_ref.float()).abs().max().item() <= ATTN_ATOL, (impl, it)
