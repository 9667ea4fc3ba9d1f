"""torchrun worker (one rank per GPU, NCCL) for tests/test_seq_shard_gpu.py::test_two_rank_nccl: every rank builds the
same seeded sequence, keeps its own page range in its own pools, runs SeqShardedDecode.layer_step with the real
collectives and compares with the single-GPU layer step it also runs locally on the whole sequence."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import quest_b200.utils as qu  # noqa: E402
from quest_b200 import parallel as par  # noqa: E402
from tests import seqshard_common as sc  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    try:
        modes = os.environ.get("QUEST_SEQSHARD_EXCHANGES", "nccl,peer").split(",")
        for exchange, kv_len, Hq, Hkv, D, P, budget in [(m, *c) for m in modes for c in ((6000, 32, 32, 128, 16, 48), (9000, 32, 8, 128, 16, 64))]:
            steps = 3
            k, v, dec = sc.seeded_case(kv_len, kv_len, Hq, Hkv, D, steps)
            full = sc.filled_controller(k, v, Hq, Hkv, D, P, budget, kv_len + steps + P, dev)
            ranges = sc.shard_token_ranges(kv_len, P, world)
            ic = sc.shard_controller(k, v, ranges[rank], rank, world, Hq, Hkv, D, P, budget, steps, dev)
            shard = par.SeqShardedDecode(ic, Hq, world, rank, budget, exchange=exchange)
            n_scored = [b - a for a, b in ranges]
            for q, k1, v1 in dec:
                q, k1, v1 = q.to(dev), k1.to(dev), v1.to(dev)
                full.prepare_metadata(1)
                full.begin_forward(1)
                want = qu.decode_layer_step(q, k1, v1, full, 0).clone()
                full.end_forward()
                if rank == world - 1:
                    ic.prepare_metadata(1)
                    n_scored[-1] = len(ic.kv_cache.indicies) - 1
                cnt = torch.tensor(n_scored, device=dev)
                dist.broadcast(cnt, src=world - 1)  # the test's bookkeeping, not the data path
                n_scored = cnt.tolist()
                shard.begin_step(n_scored)
                got = shard.layer_step(q, k1, v1, 0)
                torch.cuda.synchronize()
                shard.check_peer()
                err = (got.float() - want.float()).abs().max().item()
                assert err <= 2e-3, (rank, err)
                # the selection: my share + the others' = the single-GPU set
                lut = {int(p): i for i, p in enumerate(full.kv_cache.indicies)}
                want_sel = [sorted(lut[int(x)] for x in row) for row in full.topk_dindices_buffer.cpu().numpy()]
                first = sum(n_scored[:rank])
                mine = sc.selected_positions(shard.sel[:, :shard.k], ic.kv_cache.indicies[:n_scored[rank]], first)
                pad = torch.full((Hq, shard.k), -1, dtype=torch.int64, device=dev)
                for h, row in enumerate(mine):
                    pad[h, :len(row)] = torch.tensor(row, device=dev)
                allp = [torch.empty_like(pad) for _ in range(world)]
                dist.all_gather(allp, pad)
                got_sel = [sorted(int(x) for x in torch.cat([a[h] for a in allp]).tolist() if x >= 0) for h in range(Hq)]
                assert got_sel == want_sel, rank
        dist.barrier()
        if rank == 0:
            print("seqshard nccl ok (" + ",".join(modes) + ")")
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
