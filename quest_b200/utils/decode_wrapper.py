"""Mirror of reference quest/utils/decode_wrapper.py:11-82."""
from typing import Optional

import torch

import quest_b200._kernels as _kernels
from quest_b200.utils.utils import TensorLayout


def _check_kv_layout(kv_layout: str):
    if not hasattr(TensorLayout, kv_layout):
        raise KeyError("Invalide kv_layout {}".format(kv_layout))


class BatchDecodeWithPagedKVCacheWrapper:
    """Owns the split-KV plan and temporary buffer between begin_forward and end_forward."""

    def __init__(self, kv_layout: str = "NHD"):
        _check_kv_layout(kv_layout)
        self.kv_layout = kv_layout
        self._wrapper = _kernels.BatchDecodeWithPagedKVCachePyTorchWrapper(getattr(TensorLayout, kv_layout))

    def begin_forward(self, indptr: torch.Tensor, num_qo_heads: int, num_kv_heads: int, head_dim: int,
                      page_size: int, data_type, num_pages: Optional[int] = None):
        # dtype is passed through an empty placeholder tensor exactly as the reference does (:45-46)
        empty_data = torch.empty(0, dtype=data_type)
        self._wrapper.begin_forward(indptr, num_qo_heads, num_kv_heads, head_dim, page_size, empty_data,
                                    num_pages=num_pages)

    def end_forward(self):
        self._wrapper.end_forward()

    def forward(self, q: torch.Tensor, o: torch.Tensor, paged_kv_data: torch.Tensor, paged_kv_indices: torch.Tensor,
                paged_kv_indptr: torch.Tensor, paged_kv_last_page_len: int, paged_kv_last_page_idx: int,
                rope_scale: Optional[float] = None, rope_theta: Optional[float] = None):
        if rope_scale is None:
            rope_scale = 1.0
        if rope_theta is None:
            rope_theta = 1e4
        self._wrapper.forward(q, o, paged_kv_data, paged_kv_indices, paged_kv_indptr, paged_kv_last_page_len,
                              paged_kv_last_page_idx, rope_scale, rope_theta)
