"""Frame sharding across the GPUs of one box.

Frames are independent (no temporal state in the CNN or the bicubic kernel), so each rank gets
one contiguous frame range and there is NO data-path collective: outputs are merged on the host
by range order for the encoder.  ``torch.distributed`` is used only for the rendezvous/barrier and
(optionally) to hand small per-rank results to rank 0.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple


def frame_range(rank: int, world: int, n_frames: int) -> Tuple[int, int]:
    """Contiguous [lo, hi) of rank ``rank``: sizes differ by at most one, earlier ranks get the extra."""
    if world <= 0 or not (0 <= rank < world) or n_frames < 0:
        raise ValueError("bad rank/world/n_frames")
    base, rem = divmod(n_frames, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def all_ranges(world: int, n_frames: int) -> List[Tuple[int, int]]:
    return [frame_range(r, world, n_frames) for r in range(world)]


def owner_of(frame: int, world: int, n_frames: int) -> int:
    for r, (lo, hi) in enumerate(all_ranges(world, n_frames)):
        if lo <= frame < hi:
            return r
    raise IndexError(frame)


def merge_ordered(per_rank: Sequence[Sequence], world: int, n_frames: int) -> list:
    """Host reorder: concatenate per-rank outputs by range order and check every range is complete."""
    out = []
    for r, (lo, hi) in enumerate(all_ranges(world, n_frames)):
        if len(per_rank[r]) != hi - lo:
            raise ValueError(f"rank {r} returned {len(per_rank[r])} frames for range [{lo},{hi})")
        out.extend(per_rank[r])
    return out


def gather_to_rank0(local_items: list, n_frames: int):
    """Gather small per-frame results (checksums, tags) onto rank 0 in frame order.
    Uses the default process group (gloo on CPU, nccl's object path on GPU)."""
    import torch.distributed as dist

    world, rank = dist.get_world_size(), dist.get_rank()
    bucket = [None] * world if rank == 0 else None
    dist.gather_object(local_items, bucket, dst=0)
    if rank != 0:
        return None
    return merge_ordered(bucket, world, n_frames)
