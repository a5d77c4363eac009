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


class PoseGather:
    """Device path of the one collective for equal pair counts per rank: ONE pack
    kernel writes the rank's [R|t] rows into the all-gather send buffer, NCCL gathers them, ONE
    kernel chains the gathered rows into the trajectory.  No torch glue kernels in between;
    buffers are allocated once.  `collective_ms()` reports the device time of the all-gather."""

    def __init__(self, n_local, device, group=None):
        from . import _lib
        self.lib = _lib.load()
        self._lib = _lib
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.n_local = int(n_local)
        f64 = torch.float64
        self.send = torch.empty(self.n_local, 12, dtype=f64, device=device)
        self.recv = torch.empty(self.world * self.n_local, 12, dtype=f64, device=device)
        self.poses = torch.empty(self.world * self.n_local + 1, 4, 4, dtype=f64, device=device)
        self._evs = []

    def __call__(self, R, t, time_it=False):
        self.pack(R, t)
        return self.collect(time_it)

    def pack(self, R, t):
        """this rank's [R|t] rows -> the all-gather send buffer (current stream)"""
        L = self._lib
        L.check(self.lib.vo_pack_poses(L.ptr(R), L.ptr(t), self.n_local, L.ptr(self.send), L.cur_stream()))

    def collect(self, time_it=False):
        """all-gather of the packed rows + chaining (current stream, which may be a side stream:
        the collective of batch k can then run beside the matching of batch k + 1)"""
        L = self._lib
        st = L.cur_stream()
        if self.world > 1:
            if time_it:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            dist.all_gather_into_tensor(self.recv, self.send, group=self.group)
            if time_it:
                e1.record()
                self._evs.append((e0, e1))
            rows = self.recv
        else:
            rows = self.send
        L.check(self.lib.vo_chain_poses_packed(L.ptr(rows), self.world * self.n_local, None,
                                               L.ptr(self.poses), st))
        return self.poses

    def collective_ms(self):
        """mean device time of the timed all-gathers (call after a synchronize)"""
        if not self._evs:
            return 0.0
        ms = sum(a.elapsed_time(b) for a, b in self._evs) / len(self._evs)
        self._evs = []
        return ms
