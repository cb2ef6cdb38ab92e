"""Frame-pair sharding across ranks.  The correlation path partitions only along
the pair batch (every pair's volume, pyramid and lookups are independent,
SURVEY.md 8e): no collective touches the data path.  The only communication a
multi-GPU run needs is the max-over-ranks of a device-timed interval."""
from __future__ import annotations

from typing import Tuple

import torch


def shard_pairs(total_pairs: int, world_size: int, rank: int) -> Tuple[int, int]:
    """(first_pair, num_pairs) of `rank`: contiguous, sizes differ by at most one."""
    if not (0 <= rank < world_size) or total_pairs < 0:
        raise ValueError(f"bad shard request: total={total_pairs} world={world_size} rank={rank}")
    base, extra = divmod(total_pairs, world_size)
    count = base + (1 if rank < extra else 0)
    first = rank * base + min(rank, extra)
    return first, count


def max_over_ranks(value_ms: float, device=None) -> float:
    """Slowest rank's interval (what a multi-GPU throughput must be computed from)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value_ms)
    t = torch.tensor([float(value_ms)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
