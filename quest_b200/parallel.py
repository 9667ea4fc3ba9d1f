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

The functions take torch tensors on the process group's device (CUDA with NCCL; CPU with gloo in the tests) and use
only torch ops plus the collectives: they are the host-side glue around the CUDA kernels, not a compute path.
"""
from typing import Optional, Tuple

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
