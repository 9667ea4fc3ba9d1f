"""Sharding of the sparse decode path over the GPUs of one box (one process per GPU, torch.distributed).

Two ways to split a layer step, both from the reference's structure (every kernel there has grid.y = head,
decode_attn.cuh:454-457,1127):

* kv-head / batch sharding: every (sequence, kv head) is independent for append, estimate, top-k, decode and merge,
  so a rank that owns a slice of the heads (and its slice of the pools) runs the ordinary single-GPU path on it.
  No collective belongs to this path (the o_proj all-reduce is the caller's).  `head_shard` is all there is.

* sequence (page) sharding for contexts that do not fit one GPU: a rank owns a contiguous range of pages.  Two small
  exchanges per layer: (1) each rank's best min(k, n_local) (score, global page position) candidates per head are
  all-gathered and every rank computes the same global top-k (the tie rule -- higher score, then lower position --
  is a total order, so no broadcast is needed and the result equals the single-GPU selection bit for bit);
  (2) the partial attention states (o, lse) over each rank's share of the selected pages are all-gathered and merged
  with the log-sum-exp rule (flashinfer state.cuh:53-62; quest_merge_states on the GPU).

The module-level functions take torch tensors on the process group's device (CUDA with NCCL; CPU with gloo in the
tests) and use only torch ops plus the collectives: they state the exchange protocol.  `SeqShardedDecode` is the
CUDA path of the same protocol: every selection, attention and merge step is a libquest_b200 kernel, the two
exchanges are one NCCL all-gather each or one quest_peer_allgather kernel each (stores into the peers' symmetric memory over
NVLink, flag words, bounded wait; csrc/peer.cu).
"""
import ctypes
from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def head_shard(num_kv_heads: int, world: int, rank: int) -> Tuple[int, int]:
    """[start, end) of the kv heads owned by `rank`; the q heads of a GQA group go with their kv head."""
    if num_kv_heads % world:
        raise ValueError(f"{num_kv_heads} kv heads do not split over {world} ranks")
    per = num_kv_heads // world
    return rank * per, (rank + 1) * per


def page_shard(num_pages: int, world: int, rank: int) -> Tuple[int, int]:
    """[start, end) of the logical pages owned by `rank` (balanced to one page; the growing last page ends up on
    the last rank)."""
    return rank * num_pages // world, (rank + 1) * num_pages // world


def _order_key(scores: torch.Tensor, positions: torch.Tensor) -> torch.Tensor:
    """int64 key whose descending order is (score descending, position ascending): the selection rule of
    csrc/topk.cuh.  fp16 scores -> order-preserving 16-bit key (raft select_radix.cuh:82-104 twiddle_in)."""
    bits = scores.contiguous().view(torch.int16).to(torch.int64) & 0xFFFF
    key = torch.where(bits >= 0x8000, 0xFFFF - bits, bits | 0x8000)
    return key * (1 << 32) + ((1 << 32) - 1 - positions.to(torch.int64))


def local_candidates(scores: torch.Tensor, first_position: int, k: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """Best min(k, n_local) pages of each head among this rank's pages.  scores: fp16 [H, n_local] of the pages at
    global positions first_position.. ; returns (scores fp16 [H, c], global positions int64 [H, c])."""
    H, n = scores.shape
    c = min(k, n)
    pos = torch.arange(first_position, first_position + n, device=scores.device).expand(H, n)
    if c == 0:
        return scores[:, :0], pos[:, :0]
    top = torch.topk(_order_key(scores, pos), c, dim=1).indices
    return torch.gather(scores, 1, top), torch.gather(pos, 1, top)


def select_global_topk(scores: torch.Tensor, first_position: int, k: int, pad_to: int,
                       group: Optional[dist.ProcessGroup] = None) -> torch.Tensor:
    """Global top-k page positions per head from page-sharded scores; identical on every rank and equal to the
    single-GPU selection.  pad_to: an upper bound on min(k, n_local) over the ranks (all-gather needs equal shapes).
    Returns int64 [H, k] positions in ascending order."""
    world = dist.get_world_size(group)
    cs, cp = local_candidates(scores, first_position, k)
    H, c = cs.shape
    if c > pad_to:
        raise ValueError("pad_to is smaller than this rank's candidate count")
    ps = torch.full((H, pad_to), -float("inf"), dtype=scores.dtype, device=scores.device)
    pp = torch.full((H, pad_to), (1 << 31) - 1, dtype=torch.int64, device=scores.device)  # padding sorts last
    ps[:, :c], pp[:, :c] = cs, cp
    gs = [torch.empty_like(ps) for _ in range(world)]
    gp = [torch.empty_like(pp) for _ in range(world)]
    dist.all_gather(gs, ps, group=group)
    dist.all_gather(gp, pp, group=group)
    all_s, all_p = torch.cat(gs, dim=1), torch.cat(gp, dim=1)
    top = torch.topk(_order_key(all_s, all_p), k, dim=1).indices
    return torch.gather(all_p, 1, top).sort(dim=1).values


def combine_partials(o: torch.Tensor, lse: torch.Tensor, group: Optional[dist.ProcessGroup] = None) -> torch.Tensor:
    """Merge the ranks' partial attention states.  o: fp32 [H, D] normalised partial output, lse: fp32 [H] base-2
    log-sum-exp (-inf for a rank that attended to nothing).  Returns fp32 [H, D], identical on every rank."""
    world = dist.get_world_size(group)
    go = [torch.empty_like(o) for _ in range(world)]
    gl = [torch.empty_like(lse) for _ in range(world)]
    dist.all_gather(go, o.contiguous(), group=group)
    dist.all_gather(gl, lse.contiguous(), group=group)
    O, L = torch.stack(go), torch.stack(gl)            # [R, H, D], [R, H]
    m = L.max(dim=0).values
    w = torch.exp2(L - m)                              # state.cuh:53-62 with base-2 exponents
    w = torch.where(torch.isfinite(L), w, torch.zeros_like(w))
    return (O * w.unsqueeze(-1)).sum(dim=0) / w.sum(dim=0).unsqueeze(-1)


class SeqShardedDecode:
    """One rank's side of a sequence-sharded decode step on the GPU (BASELINE.json configs[4]).

    The rank holds an ordinary `quest_b200.utils.InferenceController` over ITS pages only (own KV pool, own metadata
    pool, own page tables; filled through append_kv like any sequence).  Ranks own ascending, contiguous ranges of
    the sequence's pages; the last rank also holds the growing last page, which is always attended and never scored
    (reference decode_page.cuh:325-351).  Per layer step:

      1. candidates: [append on the last rank] -> quest_estimate over the local pages -> quest_topk keeps the best
         c = min(k, n_local) of each head: scores into this rank's slot of the exchange buffer, LOCAL physical page
         ids kept at home.
      2. exchange 1 (all-gather of fp16 [Hq, c] scores, ~Hq*c*2 bytes per rank; exchange="nccl" or "peer").  Only scores travel: ranks are
         concatenated in page order and quest_topk breaks ties by input position, so the global selection is the
         single-GPU one bit for bit, on every rank, without positions on the wire.
      3. select: quest_topk over the [Hq, R*c] gathered scores with a payload row that holds this rank's physical
         page ids in its own columns and -1 elsewhere -> [Hq, k] ids where -1 marks "another rank's page".
      4. attend: quest_decode_run over those ids (negative ids contribute no rows; last_page_len = 0 off the last
         rank) -> normalised partial o and base-2 lse per head (-inf if the rank held none of the head's pages).
      5. exchange 2 (all-gather of one packed fp32 record [Hq*D + Hq] per rank) -> quest_merge_states
         (flashinfer state.cuh:53-62) -> fp16 o, identical on every rank.

    The phases are separate methods so that a single process can drive several shards and stand in for the
    collectives (tests/test_gpu_parity.py); `layer_step` runs them with torch.distributed.
    """

    def __init__(self, ic, num_qo_heads: int, world: int, rank: int, page_budget: int,
                 group: Optional[dist.ProcessGroup] = None, exchange: str = "nccl",
                 peer_buffers: Optional[Sequence[torch.Tensor]] = None, fused_candidates: bool = True):
        """exchange: "nccl" (all_gather_into_tensor) or "peer" (quest_peer_allgather: one kernel per exchange that
        stores into every rank's symmetric-memory staging area over NVLink and waits on flag words).
        peer_buffers: for a single process that plays all ranks on one device (tests): the ranks' exchange areas as
        uint8 tensors of peer_area_bytes() each, in rank order; by default the areas are torch symmetric memory.
        fused_candidates: phase 1 as one quest_estimate_select call (False: append + estimate + top-k as separate operators)."""
        from quest_b200 import _kernels
        self._K = _kernels
        self.lib = _kernels.load_library()
        self.ic, self.Hq, self.R, self.r, self.group = ic, int(num_qo_heads), int(world), int(rank), group
        self.k = int(page_budget) - 1  # pages selected among the scored ones; the last page comes on top
        self.Hkv, self.D, self.P = ic.num_kv_heads, ic.head_dim, ic.page_size
        self.has_last = self.r == self.R - 1
        self.dev = ic.device
        self._key = None
        self.fused_candidates = bool(fused_candidates)
        self._ids_in_payload = False
        sms = torch.cuda.get_device_properties(self.dev).multi_processor_count
        ppc, nch = ctypes.c_uint32(), ctypes.c_uint32()
        _kernels._rc(self.lib.quest_decode_plan(self.k, self.Hq, self.Hkv, self.D, self.P, sms, ctypes.byref(ppc),
                                                ctypes.byref(nch)), "quest_decode_plan")
        self.ppc = ppc.value
        self.ws = torch.zeros(self.lib.quest_decode_workspace_bytes(self.Hq, self.D, nch.value), dtype=torch.uint8,
                              device=self.dev)
        self.sel = torch.full((self.Hq, max(self.k, 1)), -1, dtype=torch.int32, device=self.dev)
        self.rec = torch.zeros(self.Hq * self.D + self.Hq, dtype=torch.float32, device=self.dev)
        self.rec_all = torch.empty(self.R, self.Hq * self.D + self.Hq, dtype=torch.float32, device=self.dev)
        self.out = torch.empty(1, self.Hq, self.D, dtype=torch.float16, device=self.dev)
        if exchange not in ("nccl", "peer"):
            raise ValueError("exchange must be 'nccl' or 'peer'")
        self.exchange = exchange
        if exchange == "peer":
            self._setup_peer(peer_buffers)

    # -- exchange areas for the peer-memory all-gather: [candidate staging][record staging][flags cand][flags rec]
    @staticmethod
    def _peer_layout(lib, Hq, D, k, R):
        c_cap = (max(k, 1) + 7) // 8 * 8
        a = lib.quest_peer_staging_bytes(Hq, c_cap * 2, R)
        b = lib.quest_peer_staging_bytes(1, (Hq * D + Hq) * 4, R)
        off_flags = (a + b + 15) // 16 * 16
        return a, off_flags, off_flags + (2 * R * 4 + 15) // 16 * 16

    def peer_area_bytes(self) -> int:
        return self._peer_layout(self.lib, self.Hq, self.D, self.k, self.R)[2]

    def _setup_peer(self, peer_buffers):
        if (self.Hq * self.D + self.Hq) * 4 % 16:
            raise ValueError("peer exchange needs num_qo_heads % 4 == 0")
        off_rec, off_flags, total = self._peer_layout(self.lib, self.Hq, self.D, self.k, self.R)
        if peer_buffers is not None:
            if len(peer_buffers) != self.R or any(b.numel() < total or b.dtype != torch.uint8 for b in peer_buffers):
                raise ValueError("peer_buffers: one uint8 tensor of peer_area_bytes() per rank")
            self._peer_keep = list(peer_buffers)
            bases = [b.data_ptr() for b in peer_buffers]
        else:
            import torch.distributed._symmetric_memory as symm
            area = symm.empty(total, dtype=torch.uint8, device=self.dev)
            area.zero_()
            hdl = symm.rendezvous(area, self.group if self.group is not None else dist.group.WORLD)
            hdl.barrier()  # every rank's flags are zero before anybody's first push
            self._peer_keep = [area, hdl]
            bases = [int(p) for p in hdl.buffer_ptrs]
        arr = ctypes.c_void_p * self.R
        self._peer_ptrs = {
            "cand": (arr(*bases), arr(*[b + off_flags for b in bases])),
            "rec": (arr(*[b + off_rec for b in bases]), arr(*[b + off_flags + self.R * 4 for b in bases])),
        }
        self.peer_state = torch.zeros(2, 1 + self.R, dtype=torch.int32, device=self.dev)  # [cand | rec][timeouts, epochs..]

    def _peer_gather(self, kind: str, src: torch.Tensor, rows: int, row_bytes: int, dst: torch.Tensor):
        staging, flags = self._peer_ptrs[kind]
        state = self.peer_state[0 if kind == "cand" else 1]
        self._K._rc(self.lib.quest_peer_allgather(src.data_ptr(), rows, row_bytes, staging, flags, state.data_ptr(),
                                                  dst.data_ptr(), self.R, self.r, 0, self._st()), "peer all-gather")

    def check_peer(self):
        """Synchronise and raise if any peer wait ran into its time limit (a rank missing from the exchange)."""
        if self.exchange == "peer":
            n = int(self.peer_state[:, 0].sum().item())
            if n:
                raise RuntimeError(f"quest_peer_allgather: {n} waits timed out (a rank did not take part)")

    def _peer_alive(self):
        """Host-side read of the library's sticky error words (no synchronisation): a peer exchange that lost a rank
        poisons its output with NaNs and raises QUEST_DEVERR_PEER, so a step after it fails here instead of returning
        garbage (the exchange itself is also refused with QUEST_ERR_DEVICE from then on)."""
        if self.lib.quest_device_error() & self._K.DEVERR_PEER:
            raise RuntimeError("sequence-sharded decode: an earlier quest_peer_allgather timed out waiting for a rank; "
                               "its outputs were poisoned (quest_device_error_clear() after the group has been repaired)")

    # -- per-layout buffers (the layout changes only when a rank gains a page)
    def begin_step(self, scored_pages_per_rank: Sequence[int]):
        """scored_pages_per_rank[r] = pages of rank r that take part in the selection (all of its pages; on the last
        rank all but the last).  Call after the controller reserved the new token (prepare_metadata on the last rank)."""
        n = [int(x) for x in scored_pages_per_rank]
        if len(n) != self.R:
            raise ValueError("scored_pages_per_rank needs one entry per rank")
        ic = self.ic
        ic.sync_tables()
        local_pages = len(ic.kv_cache.indicies)
        if n[self.r] != local_pages - (1 if self.has_last else 0):
            raise ValueError(f"rank {self.r}: controller holds {local_pages} pages, layout says {n[self.r]} scored")
        if sum(n) < self.k:
            raise ValueError("fewer scored pages than the budget selects: attend to everything instead")
        key = tuple(n)
        if key == self._key:
            return
        self._key = key
        self.n_local = n[self.r]
        self.c_max = min(self.k, max(n))
        if self.exchange == "peer":  # 16-byte rows on the wire: a few extra candidates (or -inf padding) per rank
            self.c_max = (self.c_max + 7) // 8 * 8
        self.c_loc = min(self.c_max, self.n_local)
        Hq, R, c = self.Hq, self.R, self.c_max
        f16 = dict(dtype=torch.float16, device=self.dev)
        self.scores = torch.empty(Hq, max(self.n_local, 1), **f16)
        self.cand_s = torch.full((Hq, c), -float("inf"), **f16)        # my slot of exchange 1 (padding sorts last)
        self.cand_tmp = self.cand_s if self.c_loc == c else torch.empty(Hq, max(self.c_loc, 1), **f16)
        self.cand_p = torch.empty(Hq, max(self.c_loc, 1), dtype=torch.int32, device=self.dev)
        self.gath = torch.empty(R, Hq, c, **f16)
        self.all_s = torch.empty(Hq, R, c, **f16)
        self.pay = torch.full((Hq, R, c), -1, dtype=torch.int32, device=self.dev)

    def _st(self):
        return self._K._stream()

    def candidates(self, q: torch.Tensor, k_new: Optional[torch.Tensor], v_new: Optional[torch.Tensor], layer: int):
        """Phase 1.  q: [1, Hq, D] fp16; k_new, v_new: [1, Hkv, D] (used on the last rank only).  Returns this
        rank's exchange-1 message, fp16 [Hq, c]."""
        ic, lib, rc = self.ic, self.lib, self._K._rc
        kv, meta = ic.kv_cache.buf_layer(layer), ic.metadata_cache.buf_layer(layer)
        self._ids_in_payload = False
        if self.c_loc and self.fused_candidates:
            # one call, two launches: estimate (carrying the append on the last rank) publishing bin totals, then the selection
            # from them with the scores written into this rank's exchange-1 message and the ids straight into its columns of the
            # global payload row (quest_estimate_select) -- instead of append + estimate + top-k + two copies
            app = self.has_last and k_new is not None
            code = lib.quest_estimate_select(
                q.data_ptr(), k_new.data_ptr() if app else None, v_new.data_ptr() if app else None, kv.data_ptr(),
                ic.kv_indices_with_last.data_ptr(), ic.kv_indices_with_last.numel(), ic.kv_cache.last_page_len, meta.data_ptr(),
                ic.metadata_indices.data_ptr(), ic.metadata_indices.numel(), ic.metadata_cache.last_page_len, self.n_local,
                self.scores.data_ptr(), self.cand_s.data_ptr(), self.c_max, self.pay[:, self.r].data_ptr(), self.R * self.c_max,
                self.c_loc, self.Hq, self.Hkv, self.D, self.P, self._K.EST_META_ORDERED | self._K.EST_META_STABLE,
                self.ws.data_ptr(), self.ws.numel(), self._st())
            if code == 0:
                self._ids_in_payload = True
                return self.cand_s
            if code != -2:  # -2: a shape / length the bin-total path does not take (nothing was launched): separate ops below
                rc(code, "estimate + select")
            self.fused_candidates = False
        if self.has_last and k_new is not None:
            rc(lib.quest_append_decode(k_new.data_ptr(), v_new.data_ptr(), kv.data_ptr(), ic.kv_indices_with_last.data_ptr(),
                                       ic.kv_indices_with_last.numel(), ic.kv_cache.last_page_len, meta.data_ptr(),
                                       ic.metadata_indices.data_ptr(), ic.metadata_indices.numel(),
                                       ic.metadata_cache.last_page_len, self.Hkv, self.D, self.P, self._st()), "append")
        if self.c_loc:
            rc(lib.quest_estimate(q.data_ptr(), self.scores.data_ptr(), self.n_local, meta.data_ptr(),
                                  ic.metadata_indices.data_ptr(), self.Hq, self.Hkv, self.D, self.P,
                                  self._K.EST_META_ORDERED, self._st()), "estimate")
            rc(lib.quest_topk(self.scores.data_ptr(), ic.kv_indices_with_last.data_ptr(), 0, self.cand_tmp.data_ptr(),
                              self.cand_p.data_ptr(), self.Hq, self.n_local, self.c_loc, self._st()), "local top-k")
            if self.cand_tmp is not self.cand_s:
                self.cand_s[:, :self.c_loc].copy_(self.cand_tmp)
        return self.cand_s

    def select(self, gathered: Optional[torch.Tensor]) -> torch.Tensor:
        """Phase 3.  gathered: fp16 [R, Hq, c], the ranks' messages in rank order.  Returns int32 [Hq, k]: this rank's
        physical page ids of the globally selected pages, -1 where another rank holds the page."""
        if self.k == 0:
            return self.sel[:, :0]
        if gathered is not None:  # None: all_s already holds [Hq, R, c] (peer exchange unpacks straight into it)
            self.all_s.copy_(gathered.permute(1, 0, 2))
        if self.c_loc and not self._ids_in_payload:
            self.pay[:, self.r, :self.c_loc].copy_(self.cand_p)
        n = self.R * self.c_max
        self._K._rc(self.lib.quest_topk(self.all_s.data_ptr(), self.pay.data_ptr(), n, None, self.sel.data_ptr(), self.Hq,
                                        n, self.k, self._st()), "global top-k")
        return self.sel[:, :self.k]

    def attend(self, q: torch.Tensor, layer: int) -> torch.Tensor:
        """Phase 4.  Returns this rank's exchange-2 message: fp32 [Hq*D + Hq] = normalised partial o, then lse."""
        ic = self.ic
        HD = self.Hq * self.D
        last_idx = ic.kv_last_page_idx if self.has_last else 0
        last_len = ic.kv_cache.last_page_len if self.has_last else 0
        # fp32 partial straight into the record: no fp16 rounding before the cross-rank merge, no copy
        self._K._rc(self.lib.quest_decode_run_partial(q.data_ptr(), self.rec.data_ptr(), self.rec.data_ptr() + HD * 4,
                                                      ic.kv_cache.buf_layer(layer).data_ptr(), self.sel.data_ptr(), self.k, self.k,
                                                      last_idx, last_len, self.Hq, self.Hkv, self.D, self.P, self.ppc,
                                                      self.ws.data_ptr(), self.ws.numel(), self._st()), "decode")
        return self.rec

    def merge(self, records: torch.Tensor) -> torch.Tensor:
        """Phase 5.  records: fp32 [R, Hq*D + Hq].  Returns fp16 [1, Hq, D]."""
        HD = self.Hq * self.D
        self._K._rc(self.lib.quest_merge_states(records.data_ptr(), records.data_ptr() + HD * 4, self.R, HD + self.Hq,
                                                self.out.data_ptr(), None, self.Hq, self.D, self._st()), "merge")
        return self.out

    def layer_step(self, q, k_new, v_new, layer: int) -> torch.Tensor:
        """All five phases with the process group's collectives (NCCL).  Every rank passes the same q; k_new / v_new
        matter on the last rank only.  Returns fp16 [1, Hq, D], identical on every rank."""
        msg = self.candidates(q, k_new, v_new, layer)
        if self.exchange == "peer":
            self._peer_alive()
            if self.k:
                self._peer_gather("cand", msg, self.Hq, self.c_max * 2, self.all_s)
            self.select(None)
            rec = self.attend(q, layer)
            self._peer_gather("rec", rec, 1, rec.numel() * 4, self.rec_all)
            return self.merge(self.rec_all)
        dist.all_gather_into_tensor(self.gath, msg, group=self.group)
        self.select(self.gath)
        rec = self.attend(q, layer)
        dist.all_gather_into_tensor(self.rec_all, rec, group=self.group)
        return self.merge(self.rec_all)
