"""world_size-2 gloo test of the multi-GPU host logic (sequence sharding + result gather + max timing)."""
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from stereo_visual_odometry_b200 import dist as D


def _worker(rank, world, port, n_seq, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = D.shard_sequences(n_seq, rank, world)
    local = torch.stack([torch.full((4, 16), float(s), dtype=torch.float64) for s in mine])   # fake trajectories
    g = D.gather_results(local, dst=0)
    ms = D.max_over_ranks(10.0 + rank)
    if rank == 0:
        full = D.interleave_sharded(g, n_seq)
        q.put((full[:, 0, 0].tolist(), ms))
    else:
        assert g is None
        q.put(("rank1", ms))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_gather_two_ranks():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    n_seq = 8
    procs = [ctx.Process(target=_worker, args=(r, 2, 29533, n_seq, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(2)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    full = [r for r in res if r[0] != "rank1"][0]
    assert full[0] == [float(s) for s in range(n_seq)]
    assert all(abs(r[1] - 11.0) < 1e-12 for r in res)


def test_shard_partition_properties():
    for n, w in [(64, 1), (64, 2), (64, 8), (7, 4)]:
        parts = [D.shard_sequences(n, r, w) for r in range(w)]
        assert sorted(sum(parts, [])) == list(range(n))
        assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
