"""Frame-batch partitioning over the GPUs of one box.

Frames are independent (InstanceNorm statistics are per sample; pipeline.py:1080 processes
them one by one), so ranks take contiguous frame ranges and never communicate on the data
path.  ``torch.distributed`` is used by bench.py only for the start/stop barrier and the
max-over-ranks timing reduction.
"""
from __future__ import annotations

from typing import List, Tuple


def shard_sizes(n_frames: int, world: int) -> List[int]:
    """Contiguous split, remainder spread over the first ranks (sizes differ by at most 1)."""
    if world < 1:
        raise ValueError("world must be >= 1")
    if n_frames < 0:
        raise ValueError("n_frames must be >= 0")
    base, rem = divmod(n_frames, world)
    return [base + (1 if r < rem else 0) for r in range(world)]


def shard_range(n_frames: int, world: int, rank: int) -> Tuple[int, int]:
    """[start, stop) of the frames rank ``rank`` stylises."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world of {world}")
    sizes = shard_sizes(n_frames, world)
    start = sum(sizes[:rank])
    return start, start + sizes[rank]
