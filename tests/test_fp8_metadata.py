"""fp8 (e4m3) page metadata, SURVEY.md section 8 row f-4: a flag-gated variant of this framework (the reference's metadata is fp16
only).  CPU: the oracle's outward rounding is pinned by its defining properties.  GPU: the pool the library builds equals the
oracle's value for value, the fp8 estimate equals the oracle's fp16-order estimate on the widened bounds bit for bit and is an
upper bound of the fp16 estimate, and the fp8 layer step selects the oracle's top-k of those scores."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import quest_oracle as orc


def test_e4m3_table_matches_the_format():
    v = orc.E4M3_VALUES
    assert v[0x00] == 0 and v[0x01] == 2.0 ** -9 and v[0x08] == 2.0 ** -6 and v[0x38] == 1.0 and v[0x7E] == 448.0
    assert np.isnan(v[0x7F]) and np.isnan(v[0xFF]) and v[0xFE] == -448.0
    fin = v[~np.isnan(v)]
    assert np.array_equal(fin.astype(np.float16).astype(np.float64), fin)  # every e4m3 value is an fp16 value


def test_outward_rounding_properties():
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.standard_normal(20000) * s for s in (1e-3, 0.1, 1.0, 30.0)] + [orc.E4M3_VALUES[:0x7F], [500.0, -500.0, 65504.0]])
    x = x.astype(np.float16)
    up, dn = orc.E4M3_VALUES[orc.e4m3_round_up(x)], orc.E4M3_VALUES[orc.e4m3_round_down(x)]
    xs = np.clip(x.astype(np.float64), -448, 448)
    assert (up >= xs).all() and (dn <= xs).all()
    fin = np.sort(orc.E4M3_VALUES[~np.isnan(orc.E4M3_VALUES)])
    # tight: no representable value lies strictly between the input and its rounding
    assert (np.searchsorted(fin, up, side="left") - np.searchsorted(fin, xs, side="left") <= 0).all()
    assert (np.searchsorted(fin, xs, side="right") - np.searchsorted(fin, dn, side="right") <= 0).all()
    rep = np.isin(xs, fin)
    assert np.array_equal(up[rep], xs[rep]) and np.array_equal(dn[rep], xs[rep])  # representable values are kept


# ------------------------------------------------------------------------------------------------------------------ GPU
gpu = pytest.mark.gpu
DEV = "cuda:0"


def _setup(kv_len, P, Hq, Hkv, D, seed, scale=1.0):
    import quest_b200._kernels as K
    from tests.test_gpu_parity import make_pair
    oc, gc, k, v, rng = make_pair(kv_len, P, Hkv, D, seed=seed)
    if scale != 1.0:  # rescale the stored metadata (exercises subnormal / saturating e4m3 ranges)
        gc.meta.mul_(scale)
        oc.meta_pool[:] = (oc.meta_pool.astype(np.float32) * scale).astype(np.float16)
    meta8 = torch.full(gc.meta.shape, 0x7F, dtype=torch.uint8, device=DEV)  # NaN bytes: unwritten entries show
    kvi, kvp, mi, mp = gc.tables()
    K.metadata_to_fp8(gc.meta, mi, 0, len(oc.kv_indices), meta8)
    return K, oc, gc, rng, meta8, mi


@gpu
@pytest.mark.parametrize("scale", [1.0, 1e-3, 200.0])
@pytest.mark.parametrize("kv_len,P,Hq,Hkv,D", [(1000, 16, 32, 32, 128), (300, 16, 8, 2, 64), (150, 7, 4, 4, 256), (5000, 16, 32, 8, 128)])
def test_fp8_pool_and_estimate(kv_len, P, Hq, Hkv, D, scale):
    K, oc, gc, rng, meta8, mi = _setup(kv_len, P, Hq, Hkv, D, seed=kv_len + D, scale=scale)
    n_pages = len(oc.kv_indices)
    want8 = orc.metadata_to_fp8(oc.meta_pool)
    got8 = meta8.cpu().numpy()
    for j in range(n_pages):  # written entries: equal in value (+0 and -0 are the same bound)
        m, e = oc.meta_indices[j // P], j % P
        np.testing.assert_array_equal(orc.E4M3_VALUES[got8[m, :, e]], orc.E4M3_VALUES[want8[m, :, e]])
    mask = np.ones(got8.shape, bool)
    for j in range(n_pages):
        mask[oc.meta_indices[j // P], :, j % P] = False
    assert (got8[mask] == 0x7F).all()  # nothing else was written
    # scores: the fp16-order estimate on the widened bounds, bit for bit; and an upper bound of the fp16-metadata scores
    q = rng.standard_normal((1, Hq, D), dtype=np.float32).astype(np.float16)
    n = n_pages - 1
    o8 = torch.empty((Hq, n), dtype=torch.float16, device=DEV)
    o16 = torch.empty((Hq, n), dtype=torch.float16, device=DEV)
    tq = torch.from_numpy(q).to(DEV)
    K.estimate_attn_score_fp8(tq, o8, meta8, mi)
    wide = orc.fp8_pool_to_f16(np.where(got8 == 0x7F, 0, got8).astype(np.uint8))
    want = orc.estimate_scores(q[0], wide, oc.meta_indices, n, Hq)
    np.testing.assert_array_equal(o8.cpu().numpy().view(np.int16), want.view(np.int16))
    kvi, kvp, mi2, mp = gc.tables()
    K.estimate_attn_score(tq, o16, gc.meta, mi2, mp, oc.meta_last_page_len, oc.meta_indices[-1], 0, K.EST_META_ORDERED)
    if scale != 200.0:  # 200x pushes bounds past 448, where the pool saturates and the bound is (documentedly) lost
        assert (o8.float() >= o16.float()).all()


@gpu
@pytest.mark.parametrize("kv_len,budget,Hq,Hkv", [(2500, 32, 32, 32), (4100, 64, 32, 8), (20000, 128, 32, 32)])
def test_fp8_layer_step(kv_len, budget, Hq, Hkv):
    """quest_decode_layer_step_fp8 over 3 tokens (crossing a page boundary where the copy of the page that stops being the last
    one must already be current): fp16 pools as the ordinary step leaves them, e4m3 pool = oracle's, scores = oracle's estimate
    on it, selection = oracle's top-k of those scores, output = oracle attention over the selected pages."""
    import quest_b200._kernels as K
    from tests.helpers import OracleCache
    from tests.test_gpu_parity import ATTN_ATOL, GpuCache, rand_kv, t
    P, D = 16, 128
    kv_len = kv_len - (kv_len % P) - 1  # the second appended token opens a new page
    rng = np.random.default_rng(kv_len)
    oc = OracleCache((kv_len + 3 + P - 1) // P + 2, P, Hkv, D, seed=kv_len)
    gc = GpuCache(oc)
    k0, v0 = rand_kv(rng, kv_len - 1, Hkv, D)
    oc.append(k0, v0), gc.append(k0, v0)
    lib = K.load_library()
    meta8 = torch.full(gc.meta.shape, 0x7F, dtype=torch.uint8, device=DEV)
    _, _, mi0, _ = gc.tables()
    K.metadata_to_fp8(gc.meta, mi0, 0, len(oc.kv_indices), meta8)
    kk = budget - 1
    for step in range(3):
        k1, v1 = rand_kv(rng, 1, Hkv, D)
        q = rng.standard_normal((1, Hq, D), dtype=np.float32).astype(np.float16)
        oc.append(k1, v1)
        kvi, kvp, mi, mp = gc.tables()
        n_pages = len(oc.kv_indices)
        n = n_pages - 1
        scores = torch.empty((Hq, n), dtype=torch.float16, device=DEV)
        sel = torch.full((Hq, kk), -1, dtype=torch.int32, device=DEV)
        o = torch.empty((1, Hq, D), dtype=torch.float16, device=DEV)
        ppc, nch = ctypes.c_uint32(), ctypes.c_uint32()
        assert lib.quest_decode_plan(kk, Hq, Hkv, D, P, 148, ctypes.byref(ppc), ctypes.byref(nch)) == 0
        ws = torch.zeros(lib.quest_decode_workspace_bytes(Hq, D, nch.value), dtype=torch.uint8, device=DEV)
        tq, tk, tv = t(q), t(k1), t(v1)
        rc = lib.quest_decode_layer_step_fp8(tq.data_ptr(), tk.data_ptr(), tv.data_ptr(), o.data_ptr(), gc.kv.data_ptr(), kvi.data_ptr(),
                                             n_pages, oc.kv_indices[-1], oc.last_page_len, gc.meta.data_ptr(), meta8.data_ptr(),
                                             mi.data_ptr(), len(oc.meta_indices), oc.meta_last_page_len, scores.data_ptr(),
                                             sel.data_ptr(), kk, Hq, Hkv, D, P, ppc.value, ws.data_ptr(), ws.numel(),
                                             torch.cuda.current_stream().cuda_stream)
        assert rc == 0
        torch.cuda.synchronize()
        gc.assert_equal_oracle()  # fp16 KV + metadata pools exactly as the ordinary append leaves them
        got8 = meta8.cpu().numpy()
        want8 = orc.metadata_to_fp8(oc.meta_pool)
        for j in range(n_pages):
            m, e = oc.meta_indices[j // P], j % P
            np.testing.assert_array_equal(orc.E4M3_VALUES[got8[m, :, e]], orc.E4M3_VALUES[want8[m, :, e]])
        wide = orc.fp8_pool_to_f16(np.where(got8 == 0x7F, 0, got8).astype(np.uint8))
        want_s = orc.estimate_scores(q[0], wide, oc.meta_indices, n, Hq)
        np.testing.assert_array_equal(scores.cpu().numpy().view(np.int16), want_s.view(np.int16))
        want_sel = orc.topk_select(want_s, kk)
        phys = np.asarray(oc.kv_indices, np.int32)
        np.testing.assert_array_equal(sel.cpu().numpy(), phys[want_sel])
        want_o, _ = orc.sparse_decode_attention(q[0], oc.kv_pool, phys[want_sel], oc.kv_indices[-1], oc.last_page_len)
        assert np.abs(o[0].float().cpu().numpy() - want_o).max() <= ATTN_ATOL


@gpu
def test_fp8_argument_checks():
    import quest_b200._kernels as K
    lib = K.load_library()
    z = torch.zeros(64, dtype=torch.float16, device=DEV)
    idx = torch.zeros(4, dtype=torch.int32, device=DEV)
    assert lib.quest_metadata_to_fp8(z.data_ptr(), idx.data_ptr(), 3, 2, z.data_ptr(), 1, 128, 16, None) == -1
    assert lib.quest_metadata_to_fp8(z.data_ptr(), idx.data_ptr(), 0, 1, z.data_ptr(), 1, 96, 16, None) == -2
    assert lib.quest_metadata_to_fp8(None, idx.data_ptr(), 0, 1, z.data_ptr(), 1, 128, 16, None) == -1
    assert lib.quest_estimate_fp8(z.data_ptr(), z.data_ptr(), 4, None, idx.data_ptr(), 1, 1, 128, 16, 0, None) == -1
    assert lib.quest_estimate_fp8(z.data_ptr(), z.data_ptr(), 4, z.data_ptr(), idx.data_ptr(), 3, 2, 128, 16, 0, None) == -1


@gpu
def test_fp8_flag_on_the_operator_surface():
    """InferenceController(fp8_metadata=True): prefill through append_kv, then decode tokens once with the four operators and
    once with decode_layer_step on a twin controller -- same scores (the fp8 estimate), same selected pages, same output; the
    e4m3 pool equals the outward rounding of the fp16 pool for every page."""
    import quest_b200.utils as qu
    L, H, D, P, budget, n0 = 2, 32, 128, 16, 32, 1000
    rng = np.random.default_rng(5)
    a = qu.InferenceController(L, H, D, P, budget, 2048, torch.float16, torch.device(DEV), fp8_metadata=True)
    b = qu.InferenceController(L, H, D, P, budget, 2048, torch.float16, torch.device(DEV), fp8_metadata=True)
    k0 = torch.from_numpy(rng.standard_normal((n0, H, D), dtype=np.float32).astype(np.float16)).to(DEV)
    v0 = torch.from_numpy(rng.standard_normal((n0, H, D), dtype=np.float32).astype(np.float16)).to(DEV)
    for ic in (a, b):
        ic.prepare_metadata(n0), ic.begin_forward(n0)
        for l in range(L):
            qu.append_kv(k0, v0, ic, l)
        ic.end_forward()
    for step in range(20):  # crosses a page boundary
        q = torch.randn(1, H, D, device=DEV).half()
        k1, v1 = torch.randn(1, H, D, device=DEV).half(), torch.randn(1, H, D, device=DEV).half()
        a.prepare_metadata(1), a.begin_forward(1)
        b.prepare_metadata(1), b.begin_forward(1)
        for l in range(L):
            qu.append_kv(k1, v1, a, l)
            sc = qu.decode_estimate(q, a, l)
            qu.decode_topk(sc, a)
            o1 = qu.decode_sparse_attn(q, a, l, a.topk_dindices_buffer)
            o2 = qu.decode_layer_step(q, k1, v1, b, l)
            assert torch.equal(o1, o2), (step, l)
            assert torch.equal(a.topk_dindices_buffer, b.topk_dindices_buffer)
        a.end_forward(), b.end_forward()
    torch.cuda.synchronize()
    n_pages = len(a.kv_cache.indicies)
    for ic in (a, b):
        for l in range(L):
            meta = ic.metadata_cache.buf_layer(l).cpu().numpy()
            got = ic.meta8[l].cpu().numpy()
            want = orc.metadata_to_fp8(meta)
            for j in range(n_pages):
                m, e = ic.metadata_cache.indicies[j // P], j % P
                np.testing.assert_array_equal(orc.E4M3_VALUES[got[m, :, e]], orc.E4M3_VALUES[want[m, :, e]])
