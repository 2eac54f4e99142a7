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

    ONE collective and ONE assembling copy: the local shard `[..., n_local]` is already one contiguous block, so it is
    gathered as it is into `[world, ..., n_max]`; the batch-last result is then written once (a single strided copy when
    the shards are equal, one copy per rank otherwise).  Shards may have unequal batch sizes (shard_bounds); with `batch`
    given the sizes follow from `shard_bounds` and no size exchange (and no host round trip) is needed."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return t
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    n_local = t.shape[-1]
    lead = tuple(t.shape[:-1])
    if batch is not None:
        sizes = [hi - lo for lo, hi in (shard_bounds(batch, r, world) for r in range(world))]
        if sizes[rank] != n_local:
            raise RuntimeError(f"rank {rank} holds {n_local} of {batch}, shard_bounds says {sizes[rank]}")
    else:
        got = [torch.zeros(1, dtype=torch.int64, device=t.device) for _ in range(world)]
        dist.all_gather(got, torch.tensor([n_local], dtype=torch.int64, device=t.device), group=group)
        sizes = [int(s.item()) for s in got]
    n_max = max(sizes)
    src = t.contiguous()
    if n_local < n_max:  # only the short shards of an uneven split pay a padded copy
        src = torch.zeros(lead + (n_max,), dtype=t.dtype, device=t.device)
        src[..., :n_local] = t
    out = torch.empty((world,) + lead + (n_max,), dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out.view(-1), src.view(-1), group=group)   # flat: rank r's block is out[r]
    if min(sizes) == n_max:
        return out.movedim(0, -2).reshape(lead + (world * n_max,))
    full = torch.empty(lead + (sum(sizes),), dtype=t.dtype, device=t.device)
    off = 0
    for r in range(world):
        full[..., off:off + sizes[r]] = out[r][..., :sizes[r]]
        off += sizes[r]
    return full
