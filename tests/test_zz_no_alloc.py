"""quest_decode_run_static: the sparse decode with its split-KV scratch in a file-scope device array of the library
(no allocation anywhere in the call, include/quest_b200.h).  Called through
the C ABI and compared with quest_decode_run (caller-owned workspace) on the same inputs, two calls in a row -- the
second one finds the first one's contents in the shared array.  Named test_zz_* so that pytest collects it last."""
import ctypes

import pytest
import torch

import quest_b200._kernels as K

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ATTN_ATOL = 2e-3  # the tolerance of the other attention tests (north_star: fp16 output max abs error)


def _case(Hq, Hkv, D, P, n_pages, n_sel, last_len, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    kv = torch.randn(n_pages, 2, P, Hkv, D, generator=g).half().to(DEV)
    q = torch.randn(1, Hq, D, generator=g).half().to(DEV)
    perm = torch.randperm(n_pages, generator=g).to(torch.int32)
    sel = torch.stack([torch.randperm(n_pages - 1, generator=g)[:n_sel].sort().values for _ in range(Hq)])
    phys = perm[sel].contiguous().to(DEV)
    return kv, q, phys, int(perm[-1])


@pytest.mark.parametrize("Hq,Hkv,D,P,n_pages,n_sel,last_len", [(32, 32, 128, 16, 600, 127, 16), (32, 8, 128, 16, 900, 255, 5),
                                                              (8, 8, 64, 8, 300, 40, 8)])
def test_static_scratch_decode_equals_workspace_decode(Hq, Hkv, D, P, n_pages, n_sel, last_len):
    lib = K.load_library()
    st = torch.cuda.current_stream().cuda_stream
    ppc, nch = ctypes.c_uint32(), ctypes.c_uint32()
    assert lib.quest_decode_plan(n_sel, Hq, Hkv, D, P, 148, ctypes.byref(ppc), ctypes.byref(nch)) == 0
    assert nch.value > 1, "the case must use the split-KV scratch"
    assert lib.quest_decode_workspace_bytes(Hq, D, nch.value) <= lib.quest_decode_static_scratch_bytes() == 4 << 20
    ws = torch.zeros(lib.quest_decode_workspace_bytes(Hq, D, nch.value), dtype=torch.uint8, device=DEV)
    for call in range(2):  # two calls in a row, different inputs: the array keeps the previous call's contents
        kv, q, phys, last_idx = _case(Hq, Hkv, D, P, n_pages, n_sel, last_len, seed=100 * call + Hq + n_sel)
        o_ws = torch.zeros(1, Hq, D, dtype=torch.float16, device=DEV)
        o_st = torch.full((1, Hq, D), float("nan"), dtype=torch.float16, device=DEV)
        l_ws = torch.zeros(Hq, dtype=torch.float32, device=DEV)
        l_st = torch.full((Hq,), float("nan"), dtype=torch.float32, device=DEV)
        assert lib.quest_decode_run(q.data_ptr(), o_ws.data_ptr(), l_ws.data_ptr(), kv.data_ptr(), phys.data_ptr(), n_sel, n_sel,
                                    last_idx, last_len, Hq, Hkv, D, P, ppc.value, ws.data_ptr(), ws.numel(), st) == 0
        assert lib.quest_decode_run_static(q.data_ptr(), o_st.data_ptr(), l_st.data_ptr(), kv.data_ptr(), phys.data_ptr(), n_sel,
                                           n_sel, last_idx, last_len, Hq, Hkv, D, P, ppc.value, st) == 0
        torch.cuda.synchronize()
        assert torch.isfinite(o_st).all() and torch.isfinite(l_st).all()
        assert (o_st.float() - o_ws.float()).abs().max().item() <= ATTN_ATOL, f"call {call}"
        assert (l_st - l_ws).abs().max().item() <= 1e-3, f"call {call}"


def test_static_scratch_decode_refuses_plans_beyond_its_array():
    lib = K.load_library()
    Hq, D, P = 64, 256, 16
    kv = torch.zeros(4, 2, P, Hq, D, dtype=torch.float16, device=DEV)
    q = torch.zeros(1, Hq, D, dtype=torch.float16, device=DEV)
    o = torch.zeros_like(q)
    idx = torch.zeros(Hq, 2000, dtype=torch.int32, device=DEV)
    # 2001 pages in chunks of 2 -> 1001 chunks x 64 heads x 256 dims of fp32 partials = 65 MB > 4 MiB
    assert lib.quest_decode_workspace_bytes(Hq, D, 1001) > lib.quest_decode_static_scratch_bytes()
    rc = lib.quest_decode_run_static(q.data_ptr(), o.data_ptr(), None, kv.data_ptr(), idx.data_ptr(), 2000, 2000, 0, 1, Hq, Hq, D, P,
                                     2, torch.cuda.current_stream().cuda_stream)
    assert rc == -5  # QUEST_ERR_SCRATCH_LIMIT
    with pytest.raises(RuntimeError, match="built-in array"):
        K._rc(rc, "quest_decode_run_static")
    torch.cuda.synchronize()
