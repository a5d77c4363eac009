"""Multi-GPU: frame pairs are independent, so the pair-index range is cut into contiguous blocks,
one per rank (one process per GPU under torchrun).  The only exchange is an all_gather of the
per-pair relative poses [R|t] (12 f64 = 96 B/pair) so every rank can chain the whole trajectory
(SURVEY.md §8e).  No other collective exists on this path.
"""
import torch
import torch.distributed as dist


def shard_range(n_pairs, rank, world):
    """Contiguous block [lo, hi) of pair indices owned by `rank`; equal sizes (the tail ranks
    get one fewer when world does not divide n_pairs)."""
    base, rem = divmod(n_pairs, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def frames_for_block(lo, hi):
    """Consecutive-pair blocks need frames lo..hi inclusive (one-frame halo on the right)."""
    return lo, hi + 1


def gather_relative_poses(R, t, counts=None, group=None):
    """all_gather of per-pair [R|t] rows in rank order -> ([B_total,3,3], [B_total,3]).
    `counts`: pairs per rank when they differ (padded to the max for the collective)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return R, t
    n_local = R.shape[0]
    if counts is None:
        counts = [n_local] * world
    n_max = max(counts)
    packed = torch.zeros(n_max, 12, dtype=torch.float64, device=R.device)
    packed[:n_local, :9] = R.reshape(n_local, 9)
    packed[:n_local, 9:] = t.reshape(n_local, 3)
    out = torch.empty(world * n_max, 12, dtype=torch.float64, device=R.device)
    dist.all_gather_into_tensor(out, packed, group=group)
    rows = torch.cat([out[r * n_max: r * n_max + counts[r]] for r in range(world)], 0)
    return rows[:, :9].reshape(-1, 3, 3).contiguous(), rows[:, 9:].contiguous()
