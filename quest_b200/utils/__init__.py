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


def append_kv(k: torch.Tensor, v: torch.Tensor, iController: InferenceController, layer_idx: int):
    """Append k/v ``[N, H, D]`` to the KV cache and update the page min/max metadata; dispatches to the
    prefill (N > 1) or decode (N == 1) kernel (reference __init__.py:69-125)."""
    seq_len = k.size(0)
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
    return o


def decode_topk(estimated_attn_score: torch.Tensor, iController: InferenceController):
    """Select the page_budget-1 highest-scoring pages per head into iController.topk_dindices_buffer
    (reference __init__.py:207-237)."""
    page_budet = iController.inference_page_budget - 1  # excluding the last page
    _kernels.topk_filtering(
        estimated_attn_score,
        iController.kv_indices_without_last,
        iController.topk_dout_buffer,
        iController.topk_dindices_buffer,
        iController.topk_buf,
        page_budet,
    )


def decode_sparse_attn(q: torch.Tensor, iController: InferenceController, layer_idx: int, topk_indices: torch.Tensor,
                       rope_scale: Optional[float] = None, rope_theta: Optional[float] = None) -> torch.Tensor:
    """Attention over the selected pages plus the last page (reference __init__.py:239-276)."""
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
        layers = range(ic.kv_cache.pool.num_layers)
        self.kv_ptrs = [ic.kv_cache.buf_layer(i).data_ptr() for i in layers]
        # (metadata pool,) or (metadata pool, its e4m3 copy): the fp8 entry point takes both, in this order
        self.meta_ptrs = [(ic.metadata_cache.buf_layer(i).data_ptr(),) + ((ic.meta8[i].data_ptr(),) if ic.fp8_metadata else ())
                          for i in layers]
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
