"""GPU parity tests: the CUDA path (through quest_b200._kernels -> C ABI) against the CPU oracle on the same
seeded inputs.  Bars: metadata, estimate scores and top-k page sets bit-exact; attention max-abs-err <= 2e-3."""
import os

import numpy as np
import pytest
import torch

import quest_b200._kernels as K
import quest_b200.utils as qu
from oracle import quest_oracle as orc
from tests.helpers import OracleCache, golden_inputs, prefill_impl

pytestmark = pytest.mark.gpu
ATTN_ATOL = 2e-3  # north_star: fp16 attention output max abs error
DEV = "cuda:0"


def t(x, dtype=None):
    return torch.from_numpy(np.ascontiguousarray(x)).to(DEV) if dtype is None else torch.from_numpy(np.ascontiguousarray(x)).to(DEV, dtype)


def bits(x):
    return x.detach().cpu().contiguous().view(torch.int16).numpy()


class GpuCache:
    """Device pools driven with the same (shuffled) page tables as an OracleCache."""

    def __init__(self, oc: OracleCache):
        self.oc = oc
        self.kv = torch.full(oc.kv_pool.shape, float("nan"), dtype=torch.float16, device=DEV)
        self.meta = torch.full(oc.meta_pool.shape, float("nan"), dtype=torch.float16, device=DEV)

    def tables(self):
        oc = self.oc
        return (torch.tensor(oc.kv_indices, dtype=torch.int32, device=DEV),
                torch.tensor([0, len(oc.kv_indices)], dtype=torch.int32, device=DEV),
                torch.tensor(oc.meta_indices, dtype=torch.int32, device=DEV),
                torch.tensor([0, len(oc.meta_indices)], dtype=torch.int32, device=DEV))

    def append(self, k, v):
        """Call AFTER oc.append(k, v) so the tables already include the new tokens."""
        oc = self.oc
        kvi, kvp, mi, mp = self.tables()
        f = K.append_kv_cache_prefill if k.shape[0] > 1 else K.append_kv_cache_decode
        f(t(k), t(v), self.kv, kvi, kvp, oc.last_page_len, oc.kv_indices[-1], self.meta, mi, mp,
          oc.meta_last_page_len, oc.meta_indices[-1], 0)

    def assert_equal_oracle(self):
        oc = self.oc
        for pool_g, pool_o, idx in ((self.kv, oc.kv_pool, oc.kv_indices), (self.meta, oc.meta_pool, oc.meta_indices)):
            g = pool_g[torch.tensor(idx, device=DEV)].cpu().numpy()
            o = pool_o[idx]
            written = ~np.isnan(o)
            np.testing.assert_array_equal(g.view(np.int16)[written], o.view(np.int16)[written])
            assert np.isnan(g[~written]).all()  # nothing outside the oracle's footprint was written


def rand_kv(rng, n, H, D):
    return (rng.standard_normal((n, H, D), dtype=np.float32).astype(np.float16),
            rng.standard_normal((n, H, D), dtype=np.float32).astype(np.float16))


def make_pair(kv_len, P, H, D, seed, rounds=None):
    """Oracle + GPU caches filled with a prefill of kv_len-1 tokens and one decode append (reference
    quest/tests/test_estimate.py:117-130), or with an explicit list of append sizes."""
    rng = np.random.default_rng(seed)
    rounds = [kv_len - 1, 1] if rounds is None else rounds
    oc = OracleCache((sum(rounds) + P - 1) // P + 2, P, H, D, seed=seed)
    gc = GpuCache(oc)
    ks, vs = [], []
    for n in rounds:
        k, v = rand_kv(rng, n, H, D)
        oc.append(k, v)
        gc.append(k, v)
        ks.append(k), vs.append(v)
    return oc, gc, np.concatenate(ks), np.concatenate(vs), rng


# ------------------------------------------------------------------ append + metadata (bit-exact)
@pytest.mark.parametrize("P", [1, 3, 7, 16, 31])
@pytest.mark.parametrize("kv_len,H,D", [(17, 32, 128), (97, 4, 64), (409, 8, 128), (1100, 2, 256), (4100, 32, 128)])
def test_append_prefill_then_decode_bitexact(kv_len, H, D, P):
    if kv_len // P > 600:
        pytest.skip("oracle too slow")
    oc, gc, *_ = make_pair(kv_len, P, H, D, seed=kv_len + P)
    gc.assert_equal_oracle()


@pytest.mark.parametrize("P,H,D", [(16, 32, 128), (7, 4, 64), (1, 2, 128)])
def test_append_multi_round_mixed(P, H, D):
    """Partial-page resume, page-boundary decodes and back-to-back prefills (reference test_page.cu:381-618)."""
    rounds = [5, 1, 1, 27, 1, 16, 2, 1, 1, 1, 40, 1]
    oc, gc, *_ = make_pair(0, P, H, D, seed=3, rounds=rounds)
    gc.assert_equal_oracle()


# ------------------------------------------------------------------ estimate (bit-exact)
@pytest.mark.parametrize("kv_len", [27, 61, 113, 482, 577, 1110, 1541, 2047, 3330])
def test_estimate_bitexact_reference_sweep(kv_len):
    H, D, P = 32, 128, 16
    oc, gc, k, v, rng = make_pair(kv_len, P, H, D, seed=kv_len)
    q = rng.standard_normal((1, H, D), dtype=np.float32).astype(np.float16)
    n = len(oc.kv_indices) - 1
    o = torch.empty((H, n), dtype=torch.float16, device=DEV)
    kvi, kvp, mi, mp = gc.tables()
    K.estimate_attn_score(t(q), o, gc.meta, mi, mp, oc.meta_last_page_len, oc.meta_indices[-1], 0)
    want = orc.estimate_scores(q[0], oc.meta_pool, oc.meta_indices, n)
    np.testing.assert_array_equal(bits(o), want.view(np.int16))
    # and within the reference's own tolerance of its torch restatement
    np.testing.assert_allclose(o.float().cpu().numpy(), orc.estimate_scores_torch_ref(q[0], k, P).astype(np.float32),
                               rtol=5e-3, atol=5e-3)


@pytest.mark.parametrize("Hq,Hkv,D,P,kv_len", [(8, 8, 64, 16, 300), (4, 4, 256, 16, 200), (32, 8, 128, 16, 1000),
                                                (8, 2, 128, 7, 150), (6, 3, 64, 1, 40), (32, 32, 128, 32, 2100)])
@pytest.mark.parametrize("flags", [0, K.EST_META_ORDERED, K.EST_META_ORDERED | K.EST_META_STABLE])
def test_estimate_bitexact_shapes_and_gqa(Hq, Hkv, D, P, kv_len, flags):
    """(6,3,64,...) has kv heads that do not divide the row tile -> estimate_rows_kernel; the others take the tile kernel."""
    oc, gc, k, v, rng = make_pair(kv_len, P, Hkv, D, seed=kv_len + D)
    q = rng.standard_normal((1, Hq, D), dtype=np.float32).astype(np.float16)
    n = len(oc.kv_indices) - 1
    o = torch.empty((Hq, n), dtype=torch.float16, device=DEV)
    kvi, kvp, mi, mp = gc.tables()
    torch.cuda.synchronize()  # EST_META_STABLE: nothing that writes metadata may immediately precede the call
    K.estimate_attn_score(t(q), o, gc.meta, mi, mp, oc.meta_last_page_len, oc.meta_indices[-1], 0, flags)
    np.testing.assert_array_equal(bits(o), orc.estimate_scores(q[0], oc.meta_pool, oc.meta_indices, n, Hq).view(np.int16))


@pytest.mark.parametrize("per_q", [0, 1, 2, 3])
@pytest.mark.parametrize("Hq,Hkv,D,P,kv_len", [(32, 8, 128, 16, 5000), (16, 8, 128, 16, 1300), (8, 4, 64, 16, 700), (16, 4, 256, 16, 900),
                                                (16, 2, 128, 16, 777), (64, 8, 128, 16, 2100)])
def test_estimate_bitexact_gqa_layouts(Hq, Hkv, D, P, kv_len, per_q):
    """Groups of 2 and 4 keep the group's q rows in registers and score each kv row once per q head (mixed-precision FMAs);
    DBG_EST_PER_Q = 1 selects the one-q-head-per-lane-row layout for the same shapes (the layout groups of 8 always use), 2 / 3 the
    two layouts of a group of 4 (split over two lane rows / whole group in one)."""
    oc, gc, k, v, rng = make_pair(kv_len, P, Hkv, D, seed=kv_len + Hq)
    q = rng.standard_normal((1, Hq, D), dtype=np.float32).astype(np.float16)
    n = len(oc.kv_indices) - 1
    o = torch.empty((Hq, n), dtype=torch.float16, device=DEV)
    kvi, kvp, mi, mp = gc.tables()
    with K.debug_option(K.DBG_EST_PER_Q, per_q):
        K.estimate_attn_score(t(q), o, gc.meta, mi, mp, oc.meta_last_page_len, oc.meta_indices[-1], 0, K.EST_META_ORDERED)
    np.testing.assert_array_equal(bits(o), orc.estimate_scores(q[0], oc.meta_pool, oc.meta_indices, n, Hq).view(np.int16))


@pytest.mark.parametrize("Hq,Hkv,D,P,n_pages", [(32, 32, 128, 16, 519), (8, 2, 64, 3, 77), (4, 4, 256, 7, 100), (6, 3, 128, 16, 65),
                                                 (8, 4, 128, 16, 333)])
def test_estimate_unordered_random_metadata(Hq, Hkv, D, P, n_pages):
    """Reference kernels/src/test/test_max_possible.cu:29-154: max and min slabs are independent N(0,1) draws (so
    max >= min does NOT hold) behind a shuffled page table; sum_d max(q*max, q*min) must still be bit-exact."""
    rng = np.random.default_rng(n_pages + D)
    n_meta = (n_pages + P - 1) // P
    pool = rng.standard_normal((n_meta + 3, 2, P, Hkv, D), dtype=np.float32).astype(np.float16)
    meta_indices = [int(x) for x in rng.permutation(n_meta + 3)[:n_meta]]
    q = rng.standard_normal((1, Hq, D), dtype=np.float32).astype(np.float16)
    n = n_pages - 1
    o = torch.empty((Hq, n), dtype=torch.float16, device=DEV)
    mi = torch.tensor(meta_indices, dtype=torch.int32, device=DEV)
    mp = torch.tensor([0, n_meta], dtype=torch.int32, device=DEV)
    last_len = n_pages - (n_meta - 1) * P
    K.estimate_attn_score(t(q), o, t(pool), mi, mp, last_len, meta_indices[-1], 0)
    np.testing.assert_array_equal(bits(o), orc.estimate_scores(q[0], pool, meta_indices, n, Hq).view(np.int16))


# ------------------------------------------------------------------ top-k (exact set, deterministic order)
def run_topk(scores, payload, k):
    H, n = scores.shape
    dout = torch.zeros((H, k), dtype=torch.float16, device=DEV)
    iout = torch.full((H, k), -1, dtype=torch.int32, device=DEV)
    K.topk_filtering(t(scores), payload, dout, iout, None, k)
    return dout.cpu().numpy(), iout.cpu().numpy()


@pytest.mark.parametrize("n,k", [(13, 2), (24, 24), (77, 19), (311, 69), (1024, 1024), (1025, 7), (1110, 251), (2047, 127), (2047, 1), (2048, 2048), (3001, 640), (4096, 1), (4097, 333),
                                 (8191, 255), (8192, 4095), (8193, 511), (12000, 300), (16384, 1023), (16385, 64), (65535, 511)])
@pytest.mark.parametrize("dist", ["normal", "narrow", "ties", "const"])
@pytest.mark.parametrize("radix", [0, 1, 2])
def test_topk_exact(n, k, dist, radix):  # radix = 2: no cluster variants (rows beyond 4096: the single-CTA register radix kernels)
    """quest_topk on rows of up to 16384 scores selects in one pass over bin totals the CTA counts itself (topk_hist_kernel<G, true>;
    beyond 4096 one row over a cluster of 4 / 8 CTAs); radix = 1 (DBG_TOPK_NO_HIST) keeps the register radix and radix-cluster
    kernels under test for those lengths; longer rows always use them."""
    if radix and n > 16384:
        pytest.skip("same kernels as radix = 0")
    if radix == 2 and not 4096 < n <= 16384:
        pytest.skip("one CTA per row anyway")
    with K.debug_option(K.DBG_TOPK_NO_HIST, int(radix == 1)), K.debug_option(K.DBG_TOPK_NO_CLUSTER, int(radix == 2)):
        _topk_exact(n, k, dist)


def _topk_exact(n, k, dist):
    H = 32 if n < 10000 else 4
    rng = np.random.default_rng(n * 7 + k)
    if dist == "normal":
        s = (rng.standard_normal((H, n)) * 30 + 100).astype(np.float16)
    elif dist == "narrow":  # every score in one or two bins: the low key bits decide
        s = (rng.random((H, n)) * 3.9 + 128).astype(np.float16)
    elif dist == "ties":  # few distinct values incl. +-0 and negatives: threshold ties everywhere
        s = rng.choice(np.array([-2.0, -0.0, 0.0, 0.5, 1.0, 1.5, 96.0], dtype=np.float16), size=(H, n))
    else:
        s = np.full((H, n), 3.25, dtype=np.float16)
    perm = rng.permutation(n).astype(np.int32)  # payload = shuffled page table
    payload = t(perm).unsqueeze(0).expand(H, -1)
    dout, iout = run_topk(s, payload, k)
    want_pos = orc.topk_select(s, k)
    np.testing.assert_array_equal(iout, perm[want_pos])
    np.testing.assert_array_equal(dout.view(np.int16), np.take_along_axis(s, want_pos.astype(np.int64), 1).view(np.int16))
    assert orc.topk_contract_ok(s, want_pos, k)
    # contiguous [H, n] payload (the reference's layout, controller.py:106) gives the same result
    _, iout2 = run_topk(s, payload.contiguous(), k)
    np.testing.assert_array_equal(iout2, iout)


@pytest.mark.parametrize("n,k", [(13, 2), (24, 24), (77, 19), (311, 69), (1024, 1024), (1025, 7), (2047, 127), (2047, 1), (2048, 2048), (2049, 511),
                                 (3001, 640), (4096, 1), (4097, 333), (6145, 6145), (8191, 255), (8192, 4095), (8192, 8191), (8193, 511),
                                 (12000, 300), (12289, 12288), (16384, 1023)])
@pytest.mark.parametrize("dist", ["normal", "narrow", "ties", "const"])
@pytest.mark.parametrize("cluster", [1, 0])
def test_topk_from_bin_totals_exact(n, k, dist, cluster):
    if not cluster and n <= 4096:
        pytest.skip("no cluster variant at this length")
    with K.debug_option(K.DBG_TOPK_NO_CLUSTER, 1 - cluster):  # 0: rows beyond 4096 pages in one CTA (batched steps with many rows)
        _topk_from_bin_totals_exact(n, k, dist)


def _topk_from_bin_totals_exact(n, k, dist):
    """The layer step's top-k (topk_hist_kernel: threshold from the bin totals the estimate kernel publishes) on crafted
    rows: same positions and order as the oracle, bin totals back at zero afterwards.  "narrow": every score in one or two
    bins, so the low key bits decide; "ties"/"const": the tie quota decides."""
    H = 32
    rng = np.random.default_rng(n * 11 + k)
    if dist == "normal":
        s = (rng.standard_normal((H, n)) * 30 + 100).astype(np.float16)
    elif dist == "narrow":
        s = (rng.random((H, n)) * 3.9 + 128).astype(np.float16)
    elif dist == "ties":
        s = rng.choice(np.array([-2.0, -0.0, 0.0, 0.5, 1.0, 1.5, 96.0], dtype=np.float16), size=(H, n))
    else:
        s = np.full((H, n), 3.25, dtype=np.float16)
    perm = rng.permutation(n).astype(np.int32)
    payload, ts = t(perm), t(s)
    iout = torch.full((H, k), -1, dtype=torch.int32, device=DEV)
    hist = torch.zeros(H * 1024, dtype=torch.int32, device=DEV)
    lib = K.load_library()
    import ctypes as C
    lib.quest_debug_topk_hist.restype = C.c_int
    lib.quest_debug_topk_hist.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p,
                                          C.c_void_p]
    for _ in range(2):  # the second launch finds the totals the first one put back to zero
        assert lib.quest_debug_topk_hist(ts.data_ptr(), payload.data_ptr(), 0, iout.data_ptr(), H, n, k, hist.data_ptr(),
                                         torch.cuda.current_stream().cuda_stream) == 0
        torch.cuda.synchronize()
        want_pos = orc.topk_select(s, k)
        np.testing.assert_array_equal(iout.cpu().numpy(), perm[want_pos])
        assert int(hist.abs().sum()) == 0


def test_topk_reference_test_criterion():
    """quest/tests/test_topk.py:22-64: sum of selected values == torch.topk(...).values.sum()."""
    torch.manual_seed(0)
    for kv_len, kb in [(13, 2), (244, 111), (1110, 251)]:
        w = (torch.randn(32, kv_len, device=DEV) * 3).half()
        dout = torch.randn(32, kb, dtype=torch.float16, device=DEV)
        iout = torch.zeros(32, kb, dtype=torch.int32, device=DEV)
        idx = torch.arange(kv_len, dtype=torch.int32, device=DEV).repeat(32, 1)
        K.topk_filtering(w, idx, dout, iout, torch.zeros(32, 1024, dtype=torch.float16, device=DEV), kb)
        torch.testing.assert_close(dout.float().sum(-1), torch.topk(w.float(), kb, dim=-1).values.sum(-1), rtol=5e-3, atol=5e-3)


# ------------------------------------------------------------------ sparse decode attention
def run_decode(q, gc, sel_phys, oc):
    Hq = q.shape[1]
    n_sel = sel_phys.shape[1]
    w = K.BatchDecodeWithPagedKVCachePyTorchWrapper(0)
    indptr = torch.tensor([0, n_sel], dtype=torch.int32, device=DEV)
    w.begin_forward(indptr, Hq, gc.kv.shape[3], q.shape[2], gc.kv.shape[2], torch.empty(0, dtype=torch.float16))
    o = torch.empty((1, Hq, q.shape[2]), dtype=torch.float16, device=DEV)
    idx = t(sel_phys.astype(np.int32)) if n_sel else torch.zeros((Hq, 0), dtype=torch.int32, device=DEV)
    w.forward(t(q), o, gc.kv, idx, indptr, oc.last_page_len, oc.kv_indices[-1], 1.0, 1e4)
    w.end_forward()
    return o


@pytest.mark.parametrize("P", [1, 3, 7, 16, 32])
@pytest.mark.parametrize("kv_len,budget,Hq,Hkv,D", [(63, 16, 32, 32, 128), (2047, 32, 32, 32, 128), (577, 64, 8, 8, 64),
                                                   (1541, 128, 32, 8, 128), (300, 4, 4, 4, 256), (33, 1, 32, 32, 128)])
def test_sparse_decode_random_page_subsets(kv_len, budget, Hq, Hkv, D, P):
    """Random distinct per-head page subsets (reference test_batch_decode.cu:61-76, sweep :236-257)."""
    if kv_len // P > 700:
        pytest.skip("oracle too slow")
    oc, gc, k, v, rng = make_pair(kv_len, P, Hkv, D, seed=kv_len + P + budget)
    n_pages = len(oc.kv_indices)
    n_sel = min(budget, n_pages) - 1
    logical = np.stack([np.sort(rng.permutation(n_pages - 1)[:n_sel]) for _ in range(Hq)]) if n_sel else np.zeros((Hq, 0), np.int64)
    phys = np.asarray(oc.kv_indices, dtype=np.int32)[logical] if n_sel else np.zeros((Hq, 0), np.int32)
    q = rng.standard_normal((1, Hq, D), dtype=np.float32).astype(np.float16)
    o = run_decode(q, gc, phys, oc)
    want, _ = orc.sparse_decode_attention(q[0], oc.kv_pool, phys, oc.kv_indices[-1], oc.last_page_len)
    assert torch.isfinite(o).all()
    assert np.abs(o[0].float().cpu().numpy() - want).max() <= ATTN_ATOL


@pytest.mark.parametrize("kv_len", [27, 61, 113, 482, 577, 1011])
def test_full_budget_decode_equals_dense(kv_len):
    """quest/tests/test_decode_attention.py:46-111 through the controller surface."""
    H, D, P = 32, 128, 16
    rng = np.random.default_rng(kv_len)
    q, k, v = golden_inputs(kv_len, kv_len, H, D)
    ic = qu.InferenceController(1, H, D, P, 1024, 2048, torch.float16, torch.device(DEV))
    ic.prepare_metadata(kv_len - 1), ic.begin_forward(kv_len - 1)
    qu.append_kv(t(k[:-1]), t(v[:-1]), ic, 0)
    ic.end_forward()
    ic.prepare_metadata(1), ic.begin_forward(1)
    qu.append_kv(t(k[-1:]), t(v[-1:]), ic, 0)
    assert not ic.need_estimate()
    o = qu.decode_sparse_attn(t(q), ic, 0, ic.kv_indices_without_last)
    ic.end_forward()
    want = orc.dense_attention(q, k, v, causal=True)
    assert np.abs(o.float().cpu().numpy() - want).max() <= ATTN_ATOL


# ------------------------------------------------------------------ whole layer step through the operator surface
@pytest.mark.parametrize("kv_len,budget", [(61, 3), (113, 4), (482, 7), (577, 15), (1011, 31), (1541, 55), (3330, 71)])
def test_layer_step_matches_oracle(kv_len, budget):
    """append(prefill) -> append(decode) -> estimate -> topk -> sparse attn, as QuestAttention.forward drives it
    (quest/models/QuestAttention.py:99-158); page set bit-exact, output within 2e-3."""
    H, D, P = 32, 128, 16
    q, k, v = golden_inputs(kv_len * 3 + budget, kv_len, H, D)
    ic = qu.InferenceController(2, H, D, P, budget, 4096, torch.float16, torch.device(DEV))
    ic.prepare_metadata(kv_len - 1), ic.begin_forward(kv_len - 1)
    qu.append_kv(t(k[:-1]), t(v[:-1]), ic, 1)
    ic.end_forward()
    ic.prepare_metadata(1), ic.begin_forward(1)
    qu.append_kv(t(k[-1:]), t(v[-1:]), ic, 1)
    assert ic.need_estimate()
    s = qu.decode_estimate(t(q), ic, 1)
    qu.decode_topk(s, ic)
    o = qu.decode_sparse_attn(t(q), ic, 1, ic.topk_dindices_buffer)
    ic.end_forward()

    oc = OracleCache(len(ic.kv_cache.indicies), P, H, D, shuffle=False)
    oc.append(k[:-1], v[:-1]), oc.append(k[-1:], v[-1:])
    assert oc.kv_indices == ic.kv_cache.indicies
    want_o, want_s, want_sel = orc.quest_decode_step(q[0], oc.kv_pool, oc.kv_indices, oc.last_page_len, oc.meta_pool,
                                                     oc.meta_indices, budget)
    np.testing.assert_array_equal(bits(s), want_s.view(np.int16))
    np.testing.assert_array_equal(ic.topk_dindices_buffer.cpu().numpy(), np.asarray(oc.kv_indices, np.int32)[want_sel])
    assert np.abs(o[0].float().cpu().numpy() - want_o).max() <= ATTN_ATOL


@pytest.mark.parametrize("kv_len,budget", [(61, 3), (482, 7), (1011, 31), (3330, 71)])
def test_decode_layer_step_operator_equals_four_ops(kv_len, budget):
    """quest_b200.utils.decode_layer_step (one native call per layer) against append_kv + decode_estimate + decode_topk +
    decode_sparse_attn on a second controller: output, selected ids and both pools identical, over several tokens."""
    H, D, P = 32, 128, 16
    _, k, v = golden_inputs(kv_len * 5 + budget, kv_len + 3, H, D)
    q = np.random.default_rng(kv_len).standard_normal((4, H, D), dtype=np.float32).astype(np.float16)
    ics = [qu.InferenceController(1, H, D, P, budget, 4096, torch.float16, torch.device(DEV)) for _ in range(2)]
    for ic in ics:
        ic.prepare_metadata(kv_len - 1), ic.begin_forward(kv_len - 1)
        qu.append_kv(t(k[:kv_len - 1]), t(v[:kv_len - 1]), ic, 0)
        ic.end_forward()
    for step in range(4):
        i = kv_len - 1 + step
        tq, tk, tv = t(q[step:step + 1]), t(k[i:i + 1]), t(v[i:i + 1])
        a, b = ics
        a.prepare_metadata(1), a.begin_forward(1)
        qu.append_kv(tk, tv, a, 0)
        assert a.need_estimate()
        sc = qu.decode_estimate(tq, a, 0)
        qu.decode_topk(sc, a)
        o1 = qu.decode_sparse_attn(tq, a, 0, a.topk_dindices_buffer)
        a.end_forward()
        b.prepare_metadata(1), b.begin_forward(1)
        o2 = qu.decode_layer_step(tq, tk, tv, b, 0)
        b.end_forward()
        assert torch.equal(o1, o2), f"step {step}"
        assert torch.equal(a.topk_dindices_buffer, b.topk_dindices_buffer)
        assert a.kv_cache.indicies == b.kv_cache.indicies
        for ca, cb in ((a.kv_cache, b.kv_cache), (a.metadata_cache, b.metadata_cache)):  # written slots only
            idx = torch.tensor(ca.indicies, device=DEV)
            pa, pb = ca.buf_layer(0)[idx].view(torch.int16), cb.buf_layer(0)[idx].view(torch.int16)
            assert torch.equal(pa[:-1], pb[:-1]) and torch.equal(pa[-1, :, :ca.last_page_len], pb[-1, :, :ca.last_page_len])


@pytest.mark.parametrize("Hq,Hkv", [(32, 32), (32, 8)])
def test_four_operators_bound_after_the_first_layer(Hq, Hkv):
    """The four decode operators keep their per-token arguments after the first layer's (fully checked) calls and go to the
    library directly on the other layers (quest_b200.utils._BoundOps): same pools, scores, ids and output as the one-call
    layer step on a twin controller on every layer, over tokens that cross a page boundary; a tensor that does not look like
    the first layer's still gets the reference's error message; nothing stays bound across begin_forward."""
    L, D, P, budget, kv_len = 3, 128, 16, 9, 1006
    g = torch.Generator().manual_seed(Hkv)
    rnd = lambda *s: torch.randn(*s, generator=g).half().to(DEV)
    ics = [qu.InferenceController(L, Hq, D, P, budget, 2048, torch.float16, torch.device(DEV), num_kv_heads=Hkv) for _ in range(2)]
    k0, v0 = rnd(L, kv_len, Hkv, D), rnd(L, kv_len, Hkv, D)
    for ic in ics:
        ic.prepare_metadata(kv_len), ic.begin_forward(kv_len)
        for l in range(L):
            qu.append_kv(k0[l], v0[l], ic, l)
        ic.end_forward()
        assert ic._ops is None   # a prompt binds nothing
    a, b = ics
    for step in range(20):
        a.prepare_metadata(1), a.begin_forward(1)
        b.prepare_metadata(1), b.begin_forward(1)
        for l in range(L):
            tq, tk, tv = rnd(1, Hq, D), rnd(1, Hkv, D), rnd(1, Hkv, D)
            qu.append_kv(tk, tv, a, l)
            sc = qu.decode_estimate(tq, a, l)
            qu.decode_topk(sc, a)
            o1 = qu.decode_sparse_attn(tq, a, l, a.topk_dindices_buffer)
            if l >= 1:
                assert all(x is not None for x in (a._ops.append, a._ops.estimate, a._ops.topk, a._ops.attn))
            o2 = qu.decode_layer_step(tq, tk, tv, b, l)
            assert torch.equal(o1, o2), (step, l)
            assert torch.equal(a.topk_dindices_buffer, b.topk_dindices_buffer), (step, l)
            n = len(b.kv_cache.indicies) - 1
            assert torch.equal(sc.view(torch.int16), b._scores_buffer(n).view(torch.int16)), (step, l)
        if step == 7:  # bound by now: the checked route still answers for tensors that do not fit
            with pytest.raises(RuntimeError, match="contiguous"):
                qu.append_kv(rnd(1, D, Hkv).transpose(1, 2), tv, a, 1)
            with pytest.raises(RuntimeError, match="CHECK_EQ"):
                qu.decode_estimate(rnd(1, Hq, D // 2), a, 1)
            with pytest.raises(RuntimeError):
                qu.decode_topk(sc[:, :-1], a)
            with pytest.raises(RuntimeError, match="dtype"):
                qu.decode_sparse_attn(tq.float(), a, 1, a.topk_dindices_buffer)
        a.end_forward(), b.end_forward()
        assert a._ops is None
    for ca, cb in ((a.kv_cache, b.kv_cache), (a.metadata_cache, b.metadata_cache)):
        idx = torch.tensor(ca.indicies, device=DEV)
        for l in range(L):
            pa, pb = ca.buf_layer(l)[idx].view(torch.int16), cb.buf_layer(l)[idx].view(torch.int16)
            assert torch.equal(pa[:-1], pb[:-1]) and torch.equal(pa[-1, :, :ca.last_page_len], pb[-1, :, :ca.last_page_len])


def assert_workspace_at_rest(ws, Hq):
    """Reserved words, merge tickets and the selection's bin totals are zero after every step (they reset themselves); the one
    word that moves is the step counter of the page-id hand-over at byte 64 (csrc/common.cuh, kWsEpochOff)."""
    w = ws[:4096 + 256 + Hq * 4096].view(torch.int32).clone()
    w[16] = 0
    assert int(w.abs().sum()) == 0


class topk_mode:
    """hist: the layer step's top-k selects from the bin totals the estimate kernel publishes (the default for rows of
    up to 8192 pages); else the stand-alone register radix select (explicit library switch)."""

    def __init__(self, hist):
        # hist == "nohandoff": bin totals, but the decode launch waits for the top-k launch instead of polling for its page ids
        self.opt = K.debug_option(K.DBG_TOPK_NO_HIST, 0 if hist else 1)
        self.opt2 = K.debug_option(K.DBG_NO_HANDOFF, 1 if hist == "nohandoff" else 0)

    def __enter__(self):
        self.opt2.__enter__()
        return self.opt.__enter__()

    def __exit__(self, *exc):
        self.opt2.__exit__(*exc)
        return self.opt.__exit__(*exc)


@pytest.mark.parametrize("H,kv_len,budget,P", [(32, 2500, 32, 16), (32, 32768, 128, 16), (32, 20000, 64, 16), (32, 700, 44, 16),
                                               (16, 9000, 100, 16), (8, 4000, 17, 16), (32, 5000, 8, 8), (32, 40, 2, 16),
                                               (32, 32768 + 16 * 300, 128, 16), (4, 3000, 16, 16), (32, 3000, 16, 32),
                                               (8, 131072, 256, 16), (2, 16 * 2048 + 16, 2047, 16)])
@pytest.mark.parametrize("hist", [True, "nohandoff", False])
def test_layer_step_equals_separate_calls(H, kv_len, budget, P, hist):
    """quest_sparse_decode_step against the three operator calls: scores and selected pages bit-exact, attention within
    tolerance."""
    D = 128
    oc, gc, k, v, rng = make_pair(kv_len, P, H, D, seed=11 + H + budget)
    q = rng.standard_normal((1, H, D), dtype=np.float32).astype(np.float16)
    n = len(oc.kv_indices) - 1
    kk = budget - 1
    kvi, kvp, mi, mp = gc.tables()
    s1 = torch.empty((H, n), dtype=torch.float16, device=DEV)
    K.estimate_attn_score(t(q), s1, gc.meta, mi, mp, oc.meta_last_page_len, oc.meta_indices[-1], 0)
    i1 = torch.zeros((H, kk), dtype=torch.int32, device=DEV)
    K.topk_filtering(s1, kvi[:-1].unsqueeze(0).expand(H, -1), torch.zeros((H, kk), dtype=torch.float16, device=DEV), i1, None, kk)
    o1 = run_decode(q, gc, i1.cpu().numpy(), oc)

    lib = K.load_library()
    import ctypes
    ppc, nch = ctypes.c_uint32(), ctypes.c_uint32()
    assert lib.quest_decode_plan(kk, H, H, D, P, 148, ctypes.byref(ppc), ctypes.byref(nch)) == 0
    ws = torch.zeros(lib.quest_decode_workspace_bytes(H, D, nch.value), dtype=torch.uint8, device=DEV)
    tq = t(q)
    for _ in range(3):  # repeated launches reuse the self-resetting workspace
        s2, i2 = torch.zeros_like(s1), torch.zeros_like(i1)
        o2 = torch.zeros((1, H, D), dtype=torch.float16, device=DEV)
        with topk_mode(hist):
            rc = lib.quest_sparse_decode_step(tq.data_ptr(), o2.data_ptr(), gc.kv.data_ptr(), kvi.data_ptr(), n,
                                              oc.kv_indices[-1], oc.last_page_len, gc.meta.data_ptr(), mi.data_ptr(),
                                              s2.data_ptr(), i2.data_ptr(), kk, H, H, D, P, ppc.value, ws.data_ptr(),
                                              ws.numel(), torch.cuda.current_stream().cuda_stream)
        assert rc == 0
        torch.cuda.synchronize()
        assert_workspace_at_rest(ws, H)
        assert torch.equal(s1, s2) and torch.equal(i1, i2)
        assert (o1.float() - o2.float()).abs().max().item() <= 1e-3
    if n <= 700:  # the numpy oracle is slow beyond this
        want_o, want_s, want_sel = orc.quest_decode_step(q[0], oc.kv_pool, oc.kv_indices, oc.last_page_len, oc.meta_pool,
                                                         oc.meta_indices, budget)
        assert np.abs(o2[0].float().cpu().numpy() - want_o).max() <= ATTN_ATOL
    # the stand-alone decode can share the workspace afterwards
    o3 = run_decode(q, gc, i1.cpu().numpy(), oc)
    assert torch.equal(o1, o3)


@pytest.mark.parametrize("Hq,Hkv,kv_len,budget,P,D", [(32, 32, 2500, 32, 16, 128), (32, 32, 32768, 128, 16, 128), (32, 32, 2049, 16, 16, 128),
                                                      (32, 32, 2065, 16, 16, 128), (32, 8, 3000, 24, 16, 128), (8, 8, 700, 9, 8, 64),
                                                      (32, 32, 17, 2, 16, 128), (32, 32, 9, 1, 16, 128), (6, 3, 1000, 12, 16, 128),
                                                      (32, 8, 131072, 256, 16, 128)])
@pytest.mark.parametrize("hist", [True, "nohandoff", False])
def test_decode_layer_step_equals_four_operator_calls(Hq, Hkv, kv_len, budget, P, D, hist):
    """quest_decode_layer_step (append carried by the estimate kernel) against append_kv + estimate + top-k + decode on a
    second copy of the cache: pools, scores, selected pages and output all bit-identical; repeated over 3 tokens so that
    page and metadata-page boundaries are crossed (kv_len 2049 starts a page, 2065 continues one)."""
    rng = np.random.default_rng(kv_len + budget)
    oc = OracleCache((kv_len + 3 + P - 1) // P + 2, P, Hkv, D, seed=kv_len)
    ga, gb = GpuCache(oc), GpuCache(oc)
    k0, v0 = rand_kv(rng, kv_len - 1, Hkv, D)
    oc.append(k0, v0), ga.append(k0, v0), gb.append(k0, v0)
    lib = K.load_library()
    import ctypes
    for step in range(3):
        k1, v1 = rand_kv(rng, 1, Hkv, D)
        q = rng.standard_normal((1, Hq, D), dtype=np.float32).astype(np.float16)
        oc.append(k1, v1)
        n_pages = len(oc.kv_indices)
        n, kk = n_pages - 1, min(budget, n_pages) - 1
        kvi, kvp, mi, mp = ga.tables()
        # A: the four operator calls
        ga.append(k1, v1)
        s1 = torch.zeros((Hq, n), dtype=torch.float16, device=DEV)
        i1 = torch.zeros((Hq, kk), dtype=torch.int32, device=DEV)
        if kk:
            K.estimate_attn_score(t(q), s1, ga.meta, mi, mp, oc.meta_last_page_len, oc.meta_indices[-1], 0, K.EST_META_ORDERED)
            K.topk_filtering(s1, kvi[:-1].unsqueeze(0).expand(Hq, -1), torch.zeros((Hq, kk), dtype=torch.float16, device=DEV), i1, None, kk)
        o1 = run_decode(q, ga, i1.cpu().numpy(), oc)
        # B: one call
        ppc, nch = ctypes.c_uint32(), ctypes.c_uint32()
        assert lib.quest_decode_plan(kk, Hq, Hkv, D, P, 148, ctypes.byref(ppc), ctypes.byref(nch)) == 0
        ws = torch.zeros(lib.quest_decode_workspace_bytes(Hq, D, nch.value), dtype=torch.uint8, device=DEV)
        s2, i2 = torch.zeros_like(s1), torch.zeros_like(i1)
        o2 = torch.zeros((1, Hq, D), dtype=torch.float16, device=DEV)
        tq, tk, tv = t(q), t(k1), t(v1)
        with topk_mode(hist):
            rc = lib.quest_decode_layer_step(tq.data_ptr(), tk.data_ptr(), tv.data_ptr(), o2.data_ptr(), gb.kv.data_ptr(), kvi.data_ptr(),
                                             n_pages, oc.kv_indices[-1], oc.last_page_len, gb.meta.data_ptr(), mi.data_ptr(),
                                             len(oc.meta_indices), oc.meta_last_page_len, s2.data_ptr(), i2.data_ptr(), kk, Hq, Hkv,
                                             D, P, ppc.value, ws.data_ptr(), ws.numel(), torch.cuda.current_stream().cuda_stream)
        assert rc == 0
        torch.cuda.synchronize()
        assert torch.equal(ga.kv.view(torch.int16), gb.kv.view(torch.int16)), f"kv pool, step {step}"
        assert torch.equal(ga.meta.view(torch.int16), gb.meta.view(torch.int16)), f"metadata pool, step {step}"
        assert torch.equal(s1, s2) and torch.equal(i1, i2) and torch.equal(o1, o2), f"step {step}"
        assert_workspace_at_rest(ws, Hq)
    ga.assert_equal_oracle()
    gb.assert_equal_oracle()


# ------------------------------------------------------------------ golden vectors from the reference's own code
def test_kernel_path_against_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "kernel_path.npz"))
    for i in range(5):
        seed, kv_len, budget, H, D, P = (int(x) for x in g[f"meta_{i}"])
        q, k, v = golden_inputs(seed, kv_len, H, D)
        oc = OracleCache((kv_len + P - 1) // P + 2, P, H, D, seed=seed)
        gc = GpuCache(oc)
        oc.append(k[:-1], v[:-1]), gc.append(k[:-1], v[:-1])
        oc.append(k[-1:], v[-1:]), gc.append(k[-1:], v[-1:])
        n = len(oc.kv_indices) - 1
        s = torch.empty((H, n), dtype=torch.float16, device=DEV)
        kvi, kvp, mi, mp = gc.tables()
        K.estimate_attn_score(t(q), s, gc.meta, mi, mp, oc.meta_last_page_len, oc.meta_indices[-1], 0)
        np.testing.assert_allclose(s.float().cpu().numpy(), g[f"est_{i}"].astype(np.float32), rtol=5e-3, atol=5e-3)
        phys = np.asarray(oc.kv_indices, np.int32)[g[f"apx_topk_{i}"]]
        o = run_decode(q, gc, phys, oc)
        np.testing.assert_allclose(o.float().cpu().numpy(), g[f"apx_o_{i}"], rtol=5e-3, atol=5e-3)


# ------------------------------------------------------------------ rope / rms_norm / prefill / merge
@pytest.mark.parametrize("seq,past,Hq,Hkv,D", [(1, 0, 32, 32, 128), (1, 4097, 32, 32, 128), (37, 100, 8, 2, 64), (130, 0, 4, 4, 256)])
def test_rope(seq, past, Hq, Hkv, D):
    rng = np.random.default_rng(seq + past)
    q = rng.standard_normal((seq, Hq, D), dtype=np.float32).astype(np.float16)
    k = rng.standard_normal((seq, Hkv, D), dtype=np.float32).astype(np.float16)
    tq, tk = t(q), t(k)
    qu.apply_rope_in_place(tq, tk, past)
    np.testing.assert_allclose(tq.float().cpu().numpy(), orc.rope(q, past), rtol=5e-3, atol=5e-3)
    np.testing.assert_allclose(tk.float().cpu().numpy(), orc.rope(k, past), rtol=5e-3, atol=5e-3)


def test_rms_norm():
    rng = np.random.default_rng(0)
    x = rng.standard_normal((1, 5, 4096), dtype=np.float32).astype(np.float16)
    w = rng.standard_normal(4096, dtype=np.float32).astype(np.float16)
    o = qu.rms_norm_forward(t(x), t(w), 1e-5)
    np.testing.assert_allclose(o.float().cpu().numpy(), orc.rms_norm(x, w, 1e-5), rtol=5e-3, atol=5e-3)


PREFILL_CASES = [(13, 33, 32, 32, 128, 16), (502, 1110, 8, 8, 128, 16), (64, 64, 8, 2, 64, 8), (47, 111, 4, 1, 128, 1),
                 (200, 333, 4, 4, 64, 16), (128, 128, 4, 4, 128, 16), (129, 129, 2, 2, 128, 16), (1, 700, 4, 4, 128, 16),
                 (640, 2049, 8, 2, 128, 16), (17, 4100, 4, 4, 128, 16), (257, 321, 6, 2, 64, 32), (300, 500, 4, 2, 128, 64),
                 (70, 135, 2, 1, 64, 128)]


_PREFILL_WANT = {}


def _prefill_case(qo_len, kv_len, Hq, Hkv, D, P, causal, impl, repeats=1):
    if impl.startswith("tma") and (P < 8 or P & (P - 1)):
        pytest.skip("the TMA kernel covers page sizes 2^n >= 8")
    rounds = [kv_len - qo_len, qo_len] if kv_len > qo_len else [qo_len]
    oc, gc, k, v, rng = make_pair(0, P, Hkv, D, seed=qo_len, rounds=[r for r in rounds if r])
    q = rng.standard_normal((qo_len, Hq, D), dtype=np.float32).astype(np.float16)
    kvi, *_ = gc.tables()
    key = (qo_len, kv_len, Hq, Hkv, D, P, causal)  # inputs are a function of the shape (seeded): one oracle run per shape
    if key not in _PREFILL_WANT:
        _PREFILL_WANT[key] = orc.dense_attention(q, k, v, causal=causal)
    want = _PREFILL_WANT[key]
    tq = t(q)
    with prefill_impl(impl):  # "tma_smem": the TMA kernel with P handed over through shared memory instead of tensor memory
        for _ in range(repeats):
            o = K.prefill_with_paged_kv_cache(tq, gc.kv, kvi, oc.last_page_len, causal, 0, False, 1.0, 1e4)
            torch.cuda.synchronize()
            assert K.load_library().quest_debug_prefill_timeouts() == 0 and K.load_library().quest_device_error() == 0
            assert np.abs(o.float().cpu().numpy() - want).max() <= ATTN_ATOL


@pytest.mark.parametrize("impl", ["", "tma", "tma_smem", "umma", "simt"])  # "": the library's own choice by shape
@pytest.mark.parametrize("qo_len,kv_len,Hq,Hkv,D,P", PREFILL_CASES)
def test_prefill_matches_dense_causal(qo_len, kv_len, Hq, Hkv, D, P, impl):
    """quest/tests/test_prefill_attention.py:46-87 + GQA cases of kernels/src/test/test_prefill.cu:407-467; both the
    tcgen05 kernels (TMA-fed and cp.async-fed) and the SIMT kernel on every shape (ragged query tiles, kv tails, page
    sizes 1..128, GQA)."""
    _prefill_case(qo_len, kv_len, Hq, Hkv, D, P, True, impl)


@pytest.mark.parametrize("impl", ["tma", "umma", "simt"])
@pytest.mark.parametrize("qo_len,kv_len,Hq,Hkv,D,P", [(130, 333, 4, 4, 128, 16), (64, 1000, 4, 2, 64, 8)])
def test_prefill_matches_dense_noncausal(qo_len, kv_len, Hq, Hkv, D, P, impl):
    _prefill_case(qo_len, kv_len, Hq, Hkv, D, P, False, impl)


@pytest.mark.parametrize("impl", ["", "tma", "tma_smem", "umma"])
@pytest.mark.parametrize("qo_len,kv_len,Hq,Hkv,D,P", [(2048, 2048, 32, 32, 128, 16), (4096, 4096, 32, 8, 128, 16), (1500, 3100, 48, 12, 64, 32)])
def test_prefill_multi_wave_matches_dense(qo_len, kv_len, Hq, Hkv, D, P, impl):
    """More CTAs than the GPU holds at once (512 / 1024 / 576 against ~296 resident): CTAs of different lengths succeed
    one another on an SM and co-resident CTAs drift apart, so a softmax warp can run tiles ahead of the PV MMAs -- the
    case in which the final wait of the TMA kernel used to alias two phases of its per-tile barrier (round-1 verdict).
    Five launches each."""
    _prefill_case(qo_len, kv_len, Hq, Hkv, D, P, True, impl, repeats=5)


@pytest.mark.parametrize("impl", ["tma", "tma_smem"])
def test_prefill_under_a_competing_stream(impl):
    """The TMA kernel while another stream keeps the SMs and the memory system busy (its CTAs are then scheduled in
    ragged waves and the warps of a CTA see very different latencies): 20 launches, every one against the oracle."""
    qo_len = kv_len = 3072
    Hq, Hkv, D, P = 16, 16, 128, 16
    oc, gc, k, v, rng = make_pair(0, P, Hkv, D, seed=77, rounds=[kv_len])
    q = rng.standard_normal((qo_len, Hq, D), dtype=np.float32).astype(np.float16)
    want = orc.dense_attention(q, k, v, causal=True)
    kvi, *_ = gc.tables()
    tq = t(q)
    side = torch.cuda.Stream()
    junk = torch.randn(64 * 1024 * 1024, device=DEV)
    lib = K.load_library()
    with prefill_impl(impl):
        for it in range(20):
            with torch.cuda.stream(side):
                for _ in range(4):
                    junk.mul_(1.0001).add_(0.5)
            o = K.prefill_with_paged_kv_cache(tq, gc.kv, kvi, oc.last_page_len, True, 0, False, 1.0, 1e4)
            torch.cuda.synchronize()
            assert lib.quest_debug_prefill_timeouts() == 0 and lib.quest_device_error() == 0
            assert np.abs(o.float().cpu().numpy() - want).max() <= ATTN_ATOL, f"launch {it}"


@pytest.mark.parametrize("impl", ["tma", "umma"])
def test_prefill_large_logits_lazy_rescale(impl):
    """Rows whose running maximum keeps growing by large steps (keys scaled up along the sequence): exercises the
    in-TMEM rescale of the tcgen05 kernels' accumulator."""
    qo_len = kv_len = 512
    H, D, P = 2, 128, 16
    rng = np.random.default_rng(5)
    q = rng.standard_normal((qo_len, H, D), dtype=np.float32).astype(np.float16)
    k = (rng.standard_normal((kv_len, H, D), dtype=np.float32) * np.linspace(0.2, 6.0, kv_len)[:, None, None]).astype(np.float16)
    v = rng.standard_normal((kv_len, H, D), dtype=np.float32).astype(np.float16)
    oc = OracleCache(kv_len // P + 2, P, H, D, seed=3)
    gc = GpuCache(oc)
    oc.append(k, v)
    gc.append(k, v)
    kvi, *_ = gc.tables()
    with prefill_impl(impl):
        o = K.prefill_with_paged_kv_cache(t(q), gc.kv, kvi, oc.last_page_len, True, 0, False, 1.0, 1e4)
        torch.cuda.synchronize()
    assert K.load_library().quest_debug_prefill_timeouts() == 0
    want = orc.dense_attention(q, k, v, causal=True)
    assert np.abs(o.float().cpu().numpy() - want).max() <= ATTN_ATOL


def test_merge_states():
    rng = np.random.default_rng(1)
    n, H, D = 5, 32, 128
    o_parts = rng.standard_normal((n, H, D), dtype=np.float32)
    lse = (rng.standard_normal((n, H), dtype=np.float32) * 4)
    o = torch.empty((H, D), dtype=torch.float16, device=DEV)
    l = torch.empty(H, dtype=torch.float32, device=DEV)
    lib = K.load_library()
    to, tl = t(o_parts), t(lse)
    assert lib.quest_merge_states(to.data_ptr(), tl.data_ptr(), n, 0, o.data_ptr(), l.data_ptr(), H, D,
                                  torch.cuda.current_stream().cuda_stream) == 0
    wo, wl = orc.merge_states(o_parts, lse)
    np.testing.assert_allclose(o.float().cpu().numpy(), wo, rtol=2e-3, atol=2e-3)
    np.testing.assert_allclose(l.cpu().numpy(), wl, rtol=1e-4, atol=1e-4)
    # packed records (one all-gather message per rank: o then lse), one part without any rows (lse = -inf)
    lse2 = lse.copy()
    lse2[2] = -np.inf
    rec = t(np.concatenate([o_parts.reshape(n, -1), lse2], axis=1))
    o2 = torch.empty_like(o)
    assert lib.quest_merge_states(rec.data_ptr(), rec.data_ptr() + H * D * 4, n, H * D + H, o2.data_ptr(), None, H, D,
                                  torch.cuda.current_stream().cuda_stream) == 0
    keep = [0, 1, 3, 4]
    wo2, _ = orc.merge_states(o_parts[keep], lse[keep])
    np.testing.assert_allclose(o2.float().cpu().numpy(), wo2, rtol=2e-3, atol=2e-3)


def test_error_behaviour_matches_reference_checks():
    H, D, P = 32, 128, 16
    kv = torch.zeros(4, 2, P, H, D, dtype=torch.float16, device=DEV)
    idx = torch.zeros(2, dtype=torch.int32, device=DEV)
    ptr = torch.tensor([0, 2], dtype=torch.int32, device=DEV)
    k = torch.zeros(1, H, D, dtype=torch.float16, device=DEV)
    with pytest.raises(RuntimeError, match="dispatch with dtype"):
        K.append_kv_cache_decode(k.float(), k.float(), kv, idx, ptr, 1, 0, kv, idx, ptr, 1, 0, 0)
    with pytest.raises(RuntimeError, match="CHECK_EQ"):
        K.append_kv_cache_decode(k.repeat(2, 1, 1), k.repeat(2, 1, 1), kv, idx, ptr, 1, 0, kv, idx, ptr, 1, 0, 0)
    with pytest.raises(RuntimeError, match="kInt32"):
        K.append_kv_cache_decode(k, k, kv, idx.long(), ptr, 1, 0, kv, idx, ptr, 1, 0, 0)
    with pytest.raises(RuntimeError, match="contiguous"):
        K.append_kv_cache_decode(k.transpose(1, 2), k, kv, idx, ptr, 1, 0, kv, idx, ptr, 1, 0, 0)
    with pytest.raises(RuntimeError, match="CHECK_GE"):
        K.topk_filtering(torch.zeros(H, 4, dtype=torch.float16, device=DEV), torch.zeros(H, 4, dtype=torch.int32, device=DEV),
                         torch.zeros(H, 9, dtype=torch.float16, device=DEV), torch.zeros(H, 9, dtype=torch.int32, device=DEV), None, 9)


def test_layer_step_handoff_stress():
    """The page-id hand-over (top-k kernel -> polling decode CTAs, csrc/common.cuh Handoff) over many back-to-back steps that share
    ONE workspace, alternating two budgets (two chunk layouts over the same partial-output area) and replayed from a CUDA graph:
    every step's selected ids and output must equal, bit for bit, the same steps run with the decode launch waiting for the
    top-k launch (DBG_NO_HANDOFF), and no device error may be raised."""
    import ctypes
    H, D, P, kv_len = 32, 128, 16, 9000
    oc, gc, k, v, rng = make_pair(kv_len, P, H, D, seed=77)
    lib = K.load_library()
    n_pages = len(oc.kv_indices)
    n = n_pages - 1
    kvi, kvp, mi, mp = gc.tables()
    budgets = [128, 40]
    plans = []
    for b in budgets:
        ppc, nch = ctypes.c_uint32(), ctypes.c_uint32()
        assert lib.quest_decode_plan(b - 1, H, H, D, P, 148, ctypes.byref(ppc), ctypes.byref(nch)) == 0
        plans.append((b - 1, ppc.value, nch.value))
    ws = torch.zeros(max(lib.quest_decode_workspace_bytes(H, D, p[2]) for p in plans), dtype=torch.uint8, device=DEV)
    steps = 24
    qs = torch.randn(steps, 1, H, D, device=DEV).half()
    scores = torch.empty(H, n, dtype=torch.float16, device=DEV)

    def run(no_handoff):
        outs = torch.zeros(steps, 1, H, D, dtype=torch.float16, device=DEV)
        ids = [torch.full((H, plans[i % 2][0]), -7, dtype=torch.int32, device=DEV) for i in range(steps)]
        ws.zero_()
        with K.debug_option(K.DBG_NO_HANDOFF, 1 if no_handoff else 0):
            def chain():
                st = torch.cuda.current_stream().cuda_stream
                for i in range(steps):
                    kk, ppc, _ = plans[i % 2]
                    assert lib.quest_sparse_decode_step(qs[i].data_ptr(), outs[i].data_ptr(), gc.kv.data_ptr(), kvi.data_ptr(), n,
                                                        oc.kv_indices[-1], oc.last_page_len, gc.meta.data_ptr(), mi.data_ptr(),
                                                        scores.data_ptr(), ids[i].data_ptr(), kk, H, H, D, P, ppc, ws.data_ptr(),
                                                        ws.numel(), st) == 0
            chain()  # eager
            torch.cuda.synchronize()
            eager = outs.clone()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=side):
                    chain()
                for _ in range(5):
                    g.replay()
            side.synchronize()
            torch.cuda.synchronize()
        assert torch.equal(eager, outs)
        return outs.clone(), [x.clone() for x in ids]

    o_ref, i_ref = run(True)
    o_new, i_new = run(False)
    assert lib.quest_device_error() == 0
    assert torch.equal(o_ref, o_new)
    for a, b in zip(i_ref, i_new):
        assert torch.equal(a, b) and int(a.min()) >= 0
