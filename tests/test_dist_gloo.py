"""CPU, world_size 2 over gloo: pair sharding + pose all_gather order / bit-identity vs the
unsharded run (the NCCL path differs only in the backend)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import reference_path as RP
from visual_odometry_b200.dist import shard_range, frames_for_block, gather_relative_poses


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _fake_poses(n):
    rng = np.random.default_rng(7)
    R = np.stack([RP.cv2.Rodrigues(rng.normal(0, 0.1, 3))[0] for _ in range(n)])
    return R, rng.normal(0, 1, (n, 3))


def _worker(rank, world, port, n_pairs, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    R, t = _fake_poses(n_pairs)
    lo, hi = shard_range(n_pairs, rank, world)
    counts = [shard_range(n_pairs, r, world)[1] - shard_range(n_pairs, r, world)[0] for r in range(world)]
    Rg, tg = gather_relative_poses(torch.from_numpy(R[lo:hi]), torch.from_numpy(t[lo:hi]), counts)
    q.put((rank, Rg.numpy().copy(), tg.numpy().copy()))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_ranges_cover_and_halo():
    for n, w in [(999, 2), (4096, 8), (7, 4), (3, 8)]:
        spans = [shard_range(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        assert max(h - l for l, h in spans) - min(h - l for l, h in spans) <= 1
    assert frames_for_block(10, 20) == (10, 21)   # one-frame halo on the right


def test_gather_is_bit_identical_to_unsharded_world2():
    n_pairs, world = 37, 2     # uneven split: 19 + 18
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_pairs, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    R, t = _fake_poses(n_pairs)
    for rank, Rg, tg in got:
        assert np.array_equal(Rg, R) and np.array_equal(tg, t)     # order + bits
    # chaining the gathered poses = chaining the unsharded ones
    assert np.array_equal(RP.chain_poses(got[0][1], got[0][2]), RP.chain_poses(R, t))
