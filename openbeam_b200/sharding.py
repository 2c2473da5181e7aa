"""Multi-GPU sharding of the path: independent structures / load cases are split into contiguous
index ranges, one process per GPU, no data-path collective; the only exchange is one all-gather of the
32-byte per-system summaries {max|U_f|, sum F_b,x, sum F_b,y, info} written by ``obgpu_summaries_device``.
Works on NCCL (GPU tensors) and gloo (CPU tensors; used by the world_size-2 CPU tests)."""
from __future__ import annotations

import torch
import torch.distributed as dist

SUMMARY_WIDTH = 4


def shard_range(total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous range [lo, hi) of unit indices owned by ``rank`` (sizes differ by at most one)."""
    if not 0 <= rank < world:
        raise ValueError("rank out of range")
    base, rem = divmod(total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def weak_range(per_rank: int, rank: int) -> tuple[int, int]:
    """Weak scaling (bench.py): every rank owns ``per_rank`` units, indices offset by rank."""
    return rank * per_rank, (rank + 1) * per_rank


def gather_summaries(local: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """All-gather of the per-system summaries [n_local, 4] -> [sum n_local, 4] in rank order.  Equal shard sizes
    use one ``all_gather_into_tensor``; ragged shards are padded to the maximum and trimmed."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return local
    world = dist.get_world_size()
    n = torch.tensor([local.shape[0]], dtype=torch.int64, device=local.device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    if out is not None and out.shape[0] == world * local.shape[0]:
        dist.all_gather_into_tensor(out, local.contiguous())   # fast path used by bench.py (equal shards)
        return out
    dist.all_gather(sizes, n)
    sizes = [int(s.item()) for s in sizes]
    m = max(sizes)
    pad = torch.zeros((m, local.shape[1]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    buf = torch.empty((world * m, local.shape[1]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(buf, pad)
    return torch.cat([buf[r * m: r * m + sizes[r]] for r in range(world)], 0)
