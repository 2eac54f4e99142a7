"""Multi-GPU sharding of the batch axis.

Trajectories / scenarios are independent, so the path shards by contiguous blocks of the batch axis, one
process per GPU (torchrun), with NO data-path collective.  The only communication is the optional gather of
results (final states, sensitivity / Jacobian blocks) to every rank -- NCCL all-gather over NVLink on GPUs,
gloo on CPU tensors in the tests.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_bounds(batch: int, rank: int, world: int):
    """Contiguous block [lo, hi) of the batch axis owned by `rank`; sizes differ by at most one."""
    base, rem = divmod(batch, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_last(t: torch.Tensor, rank: int, world: int, batch: int | None = None) -> torch.Tensor:
    """This rank's slice of a batch-last tensor (shared tensors without batch axis are returned as is)."""
    if batch is None:
        batch = t.shape[-1]
    if t.shape[-1] != batch:
        return t
    lo, hi = shard_bounds(batch, rank, world)
    return t[..., lo:hi].contiguous()


def gather_last(t: torch.Tensor, batch: int | None = None, group=None) -> torch.Tensor:
    """All-gathers per-rank batch-last shards into the full batch-last tensor on every rank.

    Shards may have unequal batch sizes (shard_bounds); they are padded to the largest for the collective."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return t
    world = dist.get_world_size(group)
    n_local = t.shape[-1]
    sizes = [torch.zeros(1, dtype=torch.int64, device=t.device) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([n_local], dtype=torch.int64, device=t.device), group=group)
    sizes = [int(s.item()) for s in sizes]
    n_max = max(sizes)
    # batch-last -> batch-first so that each rank's shard is one contiguous block of the gathered buffer
    src = t.movedim(-1, 0).contiguous()
    if n_local < n_max:
        pad = torch.zeros((n_max - n_local,) + tuple(src.shape[1:]), dtype=t.dtype, device=t.device)
        src = torch.cat([src, pad], dim=0)
    out = torch.empty((world * n_max,) + tuple(src.shape[1:]), dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out, src, group=group)
    parts = [out[r * n_max: r * n_max + sizes[r]] for r in range(world)]
    full = torch.cat(parts, dim=0).movedim(0, -1)
    if batch is not None and full.shape[-1] != batch:
        raise RuntimeError(f"gathered batch {full.shape[-1]} != expected {batch}")
    return full
