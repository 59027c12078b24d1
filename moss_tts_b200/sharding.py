"""Request sharding across GPUs: utterances are independent, so the batch is split into
contiguous request groups, one per rank, with no collective inside a decode step.  A collective is used
only to gather outputs (lengths, then padded waveforms/codes) at the end."""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(n_items: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous [begin, end) of the items rank `rank` owns; sizes differ by at most one."""
    base, rem = divmod(n_items, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def gather_variable(rows: List[torch.Tensor], device=None, group=None) -> Optional[List[torch.Tensor]]:
    """Gather every rank's list of 1-D tensors on rank 0 (returns None elsewhere), preserving the global
    request order.  Two collectives: all_gather of lengths, then gather of a padded matrix."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    device = device or (rows[0].device if rows else torch.device("cpu"))
    dtype = rows[0].dtype if rows else torch.float32
    counts = torch.tensor([len(rows)], device=device)
    all_counts = [torch.zeros_like(counts) for _ in range(world)]
    dist.all_gather(all_counts, counts, group=group)
    max_rows = int(max(c.item() for c in all_counts))
    lens = torch.zeros(max_rows, dtype=torch.int64, device=device)
    for i, r in enumerate(rows):
        lens[i] = r.numel()
    all_lens = [torch.zeros_like(lens) for _ in range(world)]
    dist.all_gather(all_lens, lens, group=group)
    max_len = int(max(int(l.max().item()) if l.numel() else 0 for l in all_lens))
    pad = torch.zeros(max_rows, max_len, dtype=dtype, device=device)
    for i, r in enumerate(rows):
        pad[i, : r.numel()] = r
    gathered = [torch.zeros_like(pad) for _ in range(world)] if rank == 0 else None
    dist.gather(pad, gathered, dst=0, group=group)
    if rank != 0:
        return None
    out = []
    for rk in range(world):
        for i in range(int(all_counts[rk].item())):
            out.append(gathered[rk][i, : int(all_lens[rk][i].item())].clone())
    return out
