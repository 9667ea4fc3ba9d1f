"""Python operator surface: mirror of reference quest/utils/__init__.py:11-276 (same names, arguments,
allocation behaviour), on top of quest_b200._kernels."""
from typing import Optional

import torch

import os

import quest_b200._kernels as _kernels
from quest_b200.utils.utils import TensorLayout
from quest_b200.utils.kv_cache import KvCache
from quest_b200.utils.controller import InferenceController
from quest_b200.utils.decode_wrapper import BatchDecodeWithPagedKVCacheWrapper

class _NoNvtx:
    @staticmethod
    def range_push(name):
        pass

    @staticmethod
    def range_pop():
        pass


# NVTX ranges with the reference's names (quest/models/QuestAttention.py:87-150) around the decode operators.  Off by
# default (a push/pop pair costs about a microsecond of host time per range, 8 per layer); QUEST_NVTX=1 or enable_nvtx().
_nvtx = torch.cuda.nvtx if os.environ.get("QUEST_NVTX", "0") == "1" else _NoNvtx


def enable_nvtx(on: bool = True):
    global _nvtx
    _nvtx = torch.cuda.nvtx if on else _NoNvtx


__all__ = [
    'TensorLayout',
    'KvCache',
    'InferenceController',
    "BatchDecodeWithPagedKVCacheWrapper",
    "append_kv",
    "prefill_forward",
    "prefill_layer_step",
    "decode_estimate",
    "decode_topk",
    "decode_sparse_attn",
    "decode_layer_step",
    "enable_nvtx",
    "rms_norm_forward",
    "apply_rope_in_place",
]


def apply_rope_in_place(q: torch.Tensor, k: torch.Tensor, past_kv_len: int, rope_scale: Optional[float] = None,
                        rope_theta: Optional[float] = None):
    """RoPE in place on q, k ``[N, H, D]`` at positions past_kv_len.. (reference __init__.py:25-52)."""
    if rope_scale is None:
        rope_scale = 1.0
    if rope_theta is None:
        rope_theta = 1e4
    _kernels.apply_rope_in_place(q, k, past_kv_len, rope_scale, rope_theta)


def rms_norm_forward(input: torch.Tensor, weight: torch.Tensor, epsilon: float) -> torch.Tensor:
    """reference __init__.py:54-67"""
    o = torch.empty_like(input, dtype=input.dtype, device=input.device)
    _kernels.rms_norm_forward(input, weight, o, epsilon)
    return o


class _BoundOps:
    """What the four decode operators pass to the library that does not change between the layers of one token.  The first
    call of each operator after begin_forward goes through quest_b200._kernels (every check the reference's extension makes)
    and, once that call has succeeded, leaves its constant arguments here; the same operator on the other layers then checks
    only the tensors the caller hands in and makes the library call directly.  A tensor that does not look like the first
    layer's sends the call down the checked route again, which raises the reference's message.  Dropped by begin_forward.
    (The four calls per layer of quest/models/QuestAttention.py:100-150: 1.79 -> 1.11 ms per 32-layer token, tools/four_ops_profile.py.)"""
    __slots__ = ("lib", "kv_ptrs", "meta_ptrs", "append", "estimate", "topk", "attn")

    def __init__(self, ic: InferenceController):
        self.lib = _kernels.load_library()
        self.kv_ptrs, self.meta_ptrs = ic.layer_ptrs()
        self.append = self.estimate = self.topk = self.attn = None


def _ops_of(ic: InferenceController) -> _BoundOps:
    ops = ic._ops
    if ops is None:
        ops = ic._ops = _BoundOps(ic)
    return ops


def _looks_like(t, shape, dtype, device) -> bool:
    return t.shape == shape and t.dtype == dtype and t.device == device and t.is_contiguous()


def append_kv(k: torch.Tensor, v: torch.Tensor, iController: InferenceController, layer_idx: int):
    """Append k/v ``[N, H, D]`` to the KV cache and update the page min/max metadata; dispatches to the
    prefill (N > 1) or decode (N == 1) kernel (reference __init__.py:69-125)."""
    seq_len = k.size(0)
    ops = iController._ops
    if ops is not None and ops.append is not None:
        shape, dtype, device, kv_args, meta_args = ops.append
        if _looks_like(k, shape, dtype, device) and _looks_like(v, shape, dtype, device):
            rc = ops.lib.quest_append_decode(k.data_ptr(), v.data_ptr(), ops.kv_ptrs[layer_idx], *kv_args,
                                             ops.meta_ptrs[layer_idx], *meta_args, _kernels._stream())
            if rc:
                _kernels._rc(rc, "append_kv_cache_decode")
            return
    f = _kernels.append_kv_cache_prefill if seq_len > 1 else _kernels.append_kv_cache_decode
    f(
        k,
        v,
        iController.kv_cache.buf_layer(layer_idx),
        iController.kv_indices_with_last,
        iController.kv_indptr_for_append,
        iController.kv_cache.last_page_len,
        iController.kv_last_page_idx,
        iController.metadata_cache.buf_layer(layer_idx),
        iController.metadata_indices,
        iController.metadata_indptr_for_append,
        iController.metadata_cache.last_page_len,
        iController.metadata_last_page_idx,
        iController.layout,
    )
    if seq_len == 1 and not iController.fp8_metadata:
        ic = iController
        _ops_of(ic).append = (k.shape, k.dtype, k.device,
                              (ic.kv_indices_with_last.data_ptr(), ic.kv_indices_with_last.size(0), int(ic.kv_cache.last_page_len)),
                              (ic.metadata_indices.data_ptr(), ic.metadata_indices.size(0), int(ic.metadata_cache.last_page_len),
                               ic.num_kv_heads, ic.head_dim, ic.page_size))
    if iController.fp8_metadata:  # e4m3 copies of the entries of every page this append touched
        n_pages = iController.kv_indices_with_last.size(0)
        first = (iController.kv_cache.seqlen - seq_len) // iController.page_size
        _kernels.metadata_to_fp8(iController.metadata_cache.buf_layer(layer_idx), iController.metadata_indices, first, n_pages,
                                 iController.meta8[layer_idx])


def prefill_forward(q: torch.Tensor, iController: InferenceController, layer_idx: int,
                    rope_scale: Optional[float] = None, rope_theta: Optional[float] = None) -> torch.Tensor:
    """Causal attention of the new queries over the (already appended) paged cache (reference __init__.py:127-169)."""
    if rope_scale is None:
        rope_scale = 1.0
    if rope_theta is None:
        rope_theta = 1e4
    return _kernels.prefill_with_paged_kv_cache(
        q,
        iController.kv_cache.buf_layer(layer_idx),
        iController.kv_indices_with_last,
        iController.kv_cache.last_page_len,
        True,  # causal
        iController.layout,
        False,  # fp32 QK accumulation
        rope_scale,
        rope_theta,
    )


def prefill_layer_step(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, iController: InferenceController, layer_idx: int,
                       rope_scale: Optional[float] = None, rope_theta: Optional[float] = None,
                       write_back_k: bool = False) -> torch.Tensor:
    """The prompt's attention layer in one native call: what quest/models/QuestAttention.py:85-99 does with
    apply_rope_in_place(q, k, past) -> append_kv(k, v) -> prefill_forward(q).  Call it between begin_forward and end_forward
    like the three operators; the RoPE positions start at the cache length before this append.  q is rotated in place; k is
    left as it was unless write_back_k (nothing on the path reads it again)."""
    _nvtx.range_push("prefill_layer_step")
    o = _kernels.prefill_layer_step(
        q,
        k,
        v,
        iController.kv_cache.buf_layer(layer_idx),
        iController.kv_indices_with_last,
        iController.kv_cache.last_page_len,
        iController.metadata_cache.buf_layer(layer_idx),
        iController.metadata_indices,
        1.0 if rope_scale is None else rope_scale,
        1e4 if rope_theta is None else rope_theta,
        write_back_k,
    )
    if iController.fp8_metadata:
        first = (iController.kv_cache.seqlen - k.size(0)) // iController.page_size
        _kernels.metadata_to_fp8(iController.metadata_cache.buf_layer(layer_idx), iController.metadata_indices, first,
                                 iController.kv_indices_with_last.size(0), iController.meta8[layer_idx])
    _nvtx.range_pop()
    return o


def decode_estimate(q: torch.Tensor, iController: InferenceController, layer_idx: int) -> torch.Tensor:
    """Upper-bound attention score per page and head, last page excluded (reference __init__.py:171-205)."""
    ops = iController._ops
    if ops is not None and ops.estimate is not None:
        shape, dtype, device, oshape, args = ops.estimate
        if _looks_like(q, shape, dtype, device):
            o = torch.empty(oshape, dtype=dtype, device=device)
            rc = ops.lib.quest_estimate(q.data_ptr(), o.data_ptr(), oshape[1], ops.meta_ptrs[layer_idx], *args, _kernels._stream())
            if rc:
                _kernels._rc(rc, "estimate_attn_score")
            return o
    o = torch.empty((iController.num_heads, iController.metadata_cache.seqlen - 1), dtype=q.dtype, device=q.device)
    if iController.fp8_metadata:
        _kernels.estimate_attn_score_fp8(q, o, iController.meta8[layer_idx], iController.metadata_indices)
        return o
    _kernels.estimate_attn_score(
        q,
        o,
        iController.metadata_cache.buf_layer(layer_idx),
        iController.metadata_indices,
        iController.metadata_indptr_for_append,
        iController.metadata_cache.last_page_len,
        iController.metadata_last_page_idx,
        iController.layout,
        # the controller's metadata pool is only ever written by append_kv, which keeps max >= min
        _kernels.EST_META_ORDERED,
    )
    ic, ops = iController, _ops_of(iController)
    # in a decode token (this begin_forward has seen append_kv of one token) the only metadata entry written since the last
    # synchronisation point is the last page's, which is never scored: the other layers' launches may have their first loads in
    # flight while the append kernel in front of them still runs (QUEST_EST_META_STABLE, as in decode_layer_step)
    flags = _kernels.EST_META_ORDERED | (_kernels.EST_META_STABLE if ops.append is not None else 0)
    ops.estimate = (q.shape, q.dtype, q.device, tuple(o.shape),
                    (ic.metadata_indices.data_ptr(), ic.num_heads, ic.num_kv_heads, ic.head_dim, ic.page_size, flags))
    return o


def decode_topk(estimated_attn_score: torch.Tensor, iController: InferenceController):
    """Select the page_budget-1 highest-scoring pages per head into iController.topk_dindices_buffer
    (reference __init__.py:207-237)."""
    ops = iController._ops
    if ops is not None and ops.topk is not None:
        shape, dtype, device, args = ops.topk
        if _looks_like(estimated_attn_score, shape, dtype, device):
            rc = ops.lib.quest_topk(estimated_attn_score.data_ptr(), *args, _kernels._stream())
            if rc:
                _kernels._rc(rc, "topk_filtering")
            return
    page_budet = iController.inference_page_budget - 1  # excluding the last page
    _kernels.topk_filtering(
        estimated_attn_score,
        iController.kv_indices_without_last,
        iController.topk_dout_buffer,
        iController.topk_dindices_buffer,
        iController.topk_buf,
        page_budet,
    )
    ic, e = iController, estimated_attn_score
    _ops_of(ic).topk = (e.shape, e.dtype, e.device,
                        (ic.kv_indices_without_last.data_ptr(), _kernels._check_rows(ic.kv_indices_without_last, "estimated_indices"),
                         ic.topk_dout_buffer.data_ptr(), ic.topk_dindices_buffer.data_ptr(), e.size(0), e.size(1), page_budet))


def decode_sparse_attn(q: torch.Tensor, iController: InferenceController, layer_idx: int, topk_indices: torch.Tensor,
                       rope_scale: Optional[float] = None, rope_theta: Optional[float] = None) -> torch.Tensor:
    """Attention over the selected pages plus the last page (reference __init__.py:239-276)."""
    ops = iController._ops
    if ops is not None and ops.attn is not None:
        shape, dtype, device, ids, pre, post = ops.attn
        if topk_indices is ids and _looks_like(q, shape, dtype, device):
            o = torch.empty_like(q)
            rc = ops.lib.quest_decode_run(q.data_ptr(), o.data_ptr(), None, ops.kv_ptrs[layer_idx], *pre, *post, _kernels._stream())
            if rc:
                _kernels._rc(rc, "BatchDecodeWithPagedKVCache")
            return o
    o = torch.empty_like(q, dtype=q.dtype, device=q.device)
    iController._decode_handler.forward(
        q,
        o,
        iController.kv_cache.buf_layer(layer_idx),
        topk_indices,
        iController.kv_indptr_for_approx_decode,
        iController.kv_cache.last_page_len,
        iController.kv_last_page_idx,
        rope_scale,
        rope_theta,
    )
    ic, w = iController, iController._decode_handler._wrapper
    n_sel, Hq, Hkv, D, P = w._plan
    ppc, _ = w._planned(n_sel, Hq, Hkv, D, P, q.device)
    _ops_of(ic).attn = (q.shape, q.dtype, q.device, topk_indices,
                        (topk_indices.data_ptr(), _kernels._check_rows(topk_indices, "paged_kv_indices"), n_sel,
                         int(ic.kv_last_page_idx), int(ic.kv_cache.last_page_len)),
                        (Hq, Hkv, D, P, ppc, w._ws.data_ptr(), w._ws.numel()))
    return o


class _BoundStep:
    """Everything of a decode_layer_step call that does not change between the layers of one token: validated once
    per begin_forward, so that the per-layer cost is a handful of data_ptr() calls and one ctypes call."""
    __slots__ = ("fn", "mid", "tail", "kv_ptrs", "meta_ptrs", "qshape", "kshape", "dtype", "device")  # meta_ptrs: tuples of pointers

    def __init__(self, ic: InferenceController, q, k, v):
        w = ic._decode_handler._wrapper
        L = ic.kv_cache.buf_layer(0)
        scores = ic._scores_buffer(ic.kv_indices_with_last.size(0) - 1)
        # one fully validated call's worth of checks (shapes, dtypes, devices, plan) without launching anything
        args = w.layer_step_args(q, k, v, torch.empty_like(q), L, ic.kv_indices_with_last, ic.kv_cache.last_page_len,
                                 ic.kv_last_page_idx, ic.metadata_cache.buf_layer(0), ic.metadata_indices,
                                 ic.metadata_cache.last_page_len, scores, ic.topk_dindices_buffer)
        lib = _kernels.load_library()
        self.fn = lib.quest_decode_layer_step_fp8 if ic.fp8_metadata else lib.quest_decode_layer_step
        self.mid = args[5:9]     # kv_indices, num_kv_pages, last idx, last len
        self.tail = args[10:]    # meta_indices ... workspace_bytes (stream appended per call)
        self.kv_ptrs, meta = ic.layer_ptrs()
        if ic._step_meta is None:
            # (metadata pool,) or (metadata pool, its e4m3 copy): the fp8 entry point takes both, in this order
            ic._step_meta = [(m,) + ((ic.meta8[i].data_ptr(),) if ic.fp8_metadata else ()) for i, m in enumerate(meta)]
        self.meta_ptrs = ic._step_meta
        self.qshape, self.kshape, self.dtype, self.device = q.shape, k.shape, q.dtype, q.device


def decode_layer_step(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, iController: InferenceController,
                      layer_idx: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """append_kv + decode_estimate + decode_topk + decode_sparse_attn of one layer in ONE native call (the sequence at
    reference quest/models/QuestAttention.py:100-150 when iController.need_estimate()): three kernel launches, one
    Python->C transition.  Selected ids land in iController.topk_dindices_buffer as with decode_topk.  Without a
    budget to enforce (need_estimate() false) it is append_kv + attention over all pages, as the reference does.
    ``out`` (optional, like q): written instead of a fresh tensor."""
    if not iController.need_estimate():
        _nvtx.range_push("append_kv")
        append_kv(k, v, iController, layer_idx)
        _nvtx.range_pop()
        _nvtx.range_push("full_attn")
        o = decode_sparse_attn(q, iController, layer_idx, iController.kv_indices_without_last)
        _nvtx.range_pop()
        return o if out is None else out.copy_(o)
    st = iController._step
    if st is None:
        st = iController._step = _BoundStep(iController, q, k, v)
    for t_, shape in ((q, st.qshape), (k, st.kshape), (v, st.kshape)):
        if t_.shape != shape or t_.dtype != st.dtype or t_.device != st.device or not t_.is_contiguous():
            raise RuntimeError("decode_layer_step: q must be [1, num_heads, head_dim], k and v [1, num_kv_heads, head_dim]; "
                               "contiguous fp16 tensors on the controller's device")
    o = torch.empty_like(q) if out is None else out
    if out is not None and (out.shape != st.qshape or out.dtype != st.dtype or out.device != st.device or not out.is_contiguous()):
        raise RuntimeError("decode_layer_step: out must look like q")
    # NVTX range names of the reference (quest/models/QuestAttention.py:98-150: append_kv, estimate, topk, approx_attn), nested
    # because the four operators are one native call here: a timeline of this path lines up with one of the reference
    _nvtx.range_push("append_kv")
    _nvtx.range_push("estimate")
    _nvtx.range_push("topk")
    _nvtx.range_push("approx_attn")
    rc = st.fn(q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), st.kv_ptrs[layer_idx], *st.mid,
               *st.meta_ptrs[layer_idx], *st.tail, _kernels._stream())
    _nvtx.range_pop(), _nvtx.range_pop(), _nvtx.range_pop(), _nvtx.range_pop()
    if rc:
        _kernels._rc(rc, "decode_layer_step")
    return o
