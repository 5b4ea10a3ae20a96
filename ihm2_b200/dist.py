"""Multi-GPU plumbing: problems shard by index, one process per GPU, no collective on the data path
(SURVEY.md section 8e).  The only exchange is the final gather of controls / status / statistics."""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous block [first, first + count) of `total` problems owned by `rank`; the first
    `total % world` ranks take one extra problem."""
    base, extra = divmod(total, world)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def gather_results(local: torch.Tensor, counts: list[int] | None = None) -> torch.Tensor:
    """all-gather along dim 0 (NCCL over NVLink for CUDA tensors, gloo for CPU tensors); ragged shards
    are padded to the largest one and trimmed.  Without an initialised process group returns `local`."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return local
    world = dist.get_world_size()
    if counts is None:
        n = torch.tensor([local.shape[0]], dtype=torch.int64, device=local.device)
        ns = [torch.zeros_like(n) for _ in range(world)]
        dist.all_gather(ns, n)
        counts = [int(v.item()) for v in ns]
    m = max(counts)
    pad = local
    if local.shape[0] < m:
        pad = torch.cat([local, local.new_zeros((m - local.shape[0],) + tuple(local.shape[1:]))], 0)
    out = local.new_empty((world * m,) + tuple(local.shape[1:]))
    dist.all_gather_into_tensor(out, pad.contiguous())
    return torch.cat([out[r * m: r * m + counts[r]] for r in range(world)], 0)
