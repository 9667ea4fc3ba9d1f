"""Descriptor-driven decode step for a batch of sequences (quest_decode_layer_step_batch).

The reference rebuilds its index tensors from Python lists on every token (controller.py:72-129) and passes page counts
and lengths to the kernels as host integers, so a decode step can never be replayed from a CUDA graph, and it asserts
batch 1 (QuestAttention.py:68).  Here everything that changes from token to token sits in one 64-byte record per sequence
in device memory (QuestSeqDesc): the host writes the records into a pinned buffer, one small H2D copy (capturable: the
graph's memcpy node reads the pinned buffer at replay time) refreshes the device copy, and the same captured launches serve
the next token.  The sequences keep their ordinary InferenceControllers (pools, page tables, allocation)."""
import ctypes
from typing import Optional, Sequence

import numpy as np
import torch

from quest_b200 import _kernels
from quest_b200.utils.controller import InferenceController

_BUCKETS = (1024, 2048, 4096, 8192, 10240, 12288, 16384)  # row lengths of the bin-total / register-resident top-k variants


class BatchDecodeStep:
    def __init__(self, controllers: Sequence[InferenceController], max_kv_pages: Optional[int] = None):
        ics = list(controllers)
        if not ics:
            raise ValueError("at least one sequence")
        ic0 = ics[0]
        for ic in ics:
            same = (ic.num_heads, ic.num_kv_heads, ic.head_dim, ic.page_size, ic._page_budget, ic.kv_cache.pool.capacity,
                    ic.metadata_cache.pool.capacity, ic.kv_cache.pool.num_layers, ic.device)
            if same != (ic0.num_heads, ic0.num_kv_heads, ic0.head_dim, ic0.page_size, ic0._page_budget,
                        ic0.kv_cache.pool.capacity, ic0.metadata_cache.pool.capacity, ic0.kv_cache.pool.num_layers, ic0.device):
                raise ValueError("the sequences of a batch must share shapes, budget, pool capacity and device")
        self.ics, self.B, self.dev = ics, len(ics), ic0.device
        self.Hq, self.Hkv, self.D, self.P = ic0.num_heads, ic0.num_kv_heads, ic0.head_dim, ic0.page_size
        self.k = ic0._page_budget - 1
        if self.k < 1:
            raise ValueError("page budget must select at least one page besides the last")
        self.lib = _kernels.load_library()
        cur = max(len(ic.kv_cache.indicies) for ic in ics) + 1
        if max_kv_pages is None:  # the tightest top-k variant that still leaves room to grow
            max_kv_pages = next((b + 1 for b in _BUCKETS if b + 1 >= cur + 2), None)
            if max_kv_pages is None:
                raise ValueError("descriptor mode covers rows of up to 16384 scored pages")
        self.max_kv_pages = int(max_kv_pages)
        self.layer_stride_kv = ic0.kv_cache.pool.buf[0].numel() * 2
        self.layer_stride_meta = ic0.metadata_cache.pool.buf[0].numel() * 2
        sms = torch.cuda.get_device_properties(self.dev).multi_processor_count
        ppc, nch = ctypes.c_uint32(), ctypes.c_uint32()
        _kernels._rc(self.lib.quest_decode_plan(self.k, self.Hq * self.B, self.Hkv * self.B, self.D, self.P, sms,
                                                ctypes.byref(ppc), ctypes.byref(nch)), "quest_decode_plan")
        self.ppc = ppc.value
        self.ws_seq = (self.lib.quest_decode_workspace_bytes(self.Hq, self.D, nch.value) + 15) // 16 * 16
        self.ws = torch.zeros(self.B, self.ws_seq, dtype=torch.uint8, device=self.dev)
        self.score_stride = self.Hq * (self.max_kv_pages - 1)
        self.scores = torch.empty(self.B, self.score_stride, dtype=torch.float16, device=self.dev)
        self.topk = torch.full((self.B, self.Hq, self.k), -1, dtype=torch.int32, device=self.dev)
        self.h_desc = torch.zeros(self.B, 16, dtype=torch.int32).pin_memory()
        self.d_desc = torch.zeros(self.B, 16, dtype=torch.int32, device=self.dev)

    def fits(self) -> bool:
        """False once a sequence outgrew max_kv_pages: build a new BatchDecodeStep (and re-capture)."""
        return all(len(ic.kv_cache.indicies) <= self.max_kv_pages for ic in self.ics)

    def update_host(self):
        """Write the records for the token each controller has just reserved (prepare_metadata(1)); brings the device
        page tables up to date.  Host work only, plus the tables' own small copies."""
        rec64 = self.h_desc.numpy().view(np.uint64)   # [B, 8]
        rec32 = self.h_desc.numpy().view(np.uint32)   # [B, 16]
        for b, ic in enumerate(self.ics):
            ic.sync_tables()
            n = len(ic.kv_cache.indicies)
            if n > self.max_kv_pages:
                raise RuntimeError("a sequence outgrew max_kv_pages; build a new BatchDecodeStep")
            rec64[b, 0] = ic.kv_cache.pool.buf.data_ptr()
            rec64[b, 1] = ic.metadata_cache.pool.buf.data_ptr()
            rec64[b, 2] = ic._kv_table.data_ptr()
            rec64[b, 3] = ic._meta_table.data_ptr()
            rec32[b, 8:13] = (n, ic.kv_cache.last_page_len, len(ic.metadata_cache.indicies),
                              ic.metadata_cache.last_page_len, ic.kv_cache.seqlen - 1)

    def upload(self):
        """Pinned host records -> device records on the current stream (capturable)."""
        self.d_desc.copy_(self.h_desc, non_blocking=True)

    def run_ahead(self, tokens: int):
        """Reserve the next `tokens` tokens of every sequence NOW (pages allocated, device page tables filled) and leave the device
        records at the last token that exists, so that a step whose first call is advance() can be replayed `tokens` times with
        no host work in between: the records move on the device (quest_seq_desc_advance).  The controllers' host state is then
        `tokens` ahead of the device until the run has been replayed in full."""
        self.update_host()   # records of the state before the run (update_host reads the controllers as they are)
        self.upload()
        for ic in self.ics:
            ic.prepare_metadata(int(tokens))
            ic.sync_tables()
        if not self.fits():
            raise RuntimeError("the run outgrows max_kv_pages; build a new BatchDecodeStep")

    def advance(self):
        """One token further on the device (capturable): call it first in the step of a run_ahead() run."""
        _kernels._rc(self.lib.quest_seq_desc_advance(self.d_desc.data_ptr(), self.B, self.P, _kernels._stream()), "seq_desc_advance")

    def rope(self, q: torch.Tensor, k: torch.Tensor, rope_scale: float = 1.0, rope_theta: float = 1e4):
        """apply_rope_in_place for the new token of every sequence at its own position: q [B, Hq, D], k [B, Hkv, D]."""
        _kernels._rc(self.lib.quest_rope_decode_batch(q.data_ptr(), k.data_ptr(), self.d_desc.data_ptr(), self.B, self.Hq,
                                                      self.Hkv, self.D, float(rope_scale), float(rope_theta),
                                                      _kernels._stream()), "rope_decode_batch")

    def layer_step(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, layer: int,
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """append_kv + estimate + top-k + sparse decode of one layer for all sequences: q [B, Hq, D], k, v [B, Hkv, D]
        (contiguous fp16).  Returns o [B, Hq, D]; selected ids are in self.topk[b] (-1 = no page)."""
        for t_, h in ((q, self.Hq), (k, self.Hkv), (v, self.Hkv)):
            if t_.shape != (self.B, h, self.D) or t_.dtype != torch.float16 or not t_.is_contiguous() or t_.device != self.d_desc.device:
                raise RuntimeError("layer_step: q [B, Hq, D], k and v [B, Hkv, D]; contiguous fp16 on the controllers' device")
        o = torch.empty_like(q) if out is None else out
        rc = self.lib.quest_decode_layer_step_batch(
            q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), self.d_desc.data_ptr(), self.B,
            layer * self.layer_stride_kv, layer * self.layer_stride_meta, self.max_kv_pages, self.scores.data_ptr(),
            self.score_stride, self.topk.data_ptr(), self.k, self.Hq, self.Hkv, self.D, self.P, self.ppc, self.ws.data_ptr(),
            self.ws_seq, _kernels._stream())
        if rc:
            _kernels._rc(rc, "decode_layer_step_batch")
        return o
