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


class PeerGather:
    """The final all-gather done by the solve kernel's own stores (``obgpu_bind_summary_peers``): ``buf`` is the gathered
    array [world*n_local, 4] in symmetric memory (every rank's copy is mapped into every process); ``slots`` are the
    addresses of THIS rank's block inside each rank's copy, which the last step's kernel writes over NVLink while it
    computes.  ``fence()`` is the device-side barrier after which every rank's ``buf`` is complete.  torch is plumbing
    here (allocation, handle exchange, barrier); the data moves by the kernel's bulk stores."""

    def __init__(self, n_local: int, device: torch.device):
        import torch.distributed._symmetric_memory as symm

        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        if self.world > 8:
            raise ValueError("obgpu_bind_summary_peers takes at most 8 destinations")
        self.buf = symm.empty((self.world * n_local, SUMMARY_WIDTH), dtype=torch.float64, device=device)
        self.hdl = symm.rendezvous(self.buf, dist.group.WORLD)
        off = self.rank * n_local * SUMMARY_WIDTH * 8
        self.slots = [int(p) + off for p in self.hdl.buffer_ptrs]

    def fence(self):
        self.hdl.barrier()   # on the current stream


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
