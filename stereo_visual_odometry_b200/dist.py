"""Host-side logic of the multi-GPU path: independent sequences are sharded across ranks with no
data-path collective; the only exchange is one gather of the per-sequence results.  torch.distributed is
plumbing here (NCCL on the GPUs, gloo in the CPU tests)."""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_sequences(n_sequences: int, rank: int, world: int) -> list[int]:
    """sequence s -> rank (s mod world): the ids this rank owns, ascending."""
    return list(range(rank, n_sequences, world))


def gather_results(local: torch.Tensor, dst: int = 0):
    """Gather equally shaped per-rank result tensors on `dst`; returns the list there, None elsewhere."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return [local]
    rank, world = dist.get_rank(), dist.get_world_size()
    out = [torch.empty_like(local) for _ in range(world)] if rank == dst else None
    dist.gather(local, out, dst=dst)
    return out


def max_over_ranks(value: float, device="cpu") -> float:
    """Multi-GPU timings are reported as the max over ranks."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def interleave_sharded(gathered: list[torch.Tensor], n_sequences: int) -> torch.Tensor:
    """Undo shard_sequences: gathered[r][k] is sequence r + k*world."""
    world = len(gathered)
    out = [None] * n_sequences
    for r, g in enumerate(gathered):
        for k, s in enumerate(range(r, n_sequences, world)):
            out[s] = g[k]
    return torch.stack(out)
