"""N>1 host logic on CPU: world_size-2 gloo.  The step path has no collective; what is exercised is
env sharding and the end-of-rollout reduction bench.py performs (max time, summed work and stats)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nclone_b200.sharding import shard_env_range, shard_env_to_map, reduce_rollout


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard_env_range(1001, rank, world)
    stats = torch.tensor([10 + rank, 1 + rank, 9, 0], dtype=torch.int64)
    t, w, s = reduce_rollout([100.0 + 50 * rank, 7.0], [hi - lo, 3.0], stats, dist)
    out[rank] = (lo, hi, t, w, s.tolist())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_reduction():
    world = 2
    mgr = mp.Manager(); out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    (lo0, hi0, t0, w0, s0), (lo1, hi1, t1, w1, s1) = out[0], out[1]
    assert (lo0, hi0, lo1, hi1) == (0, 501, 501, 1001)
    assert t0 == t1 == [150.0, 7.0]                 # max over ranks
    assert w0 == w1 == [1001.0, 6.0]                # summed work: every env counted once
    assert s0 == s1 == [21, 3, 18, 0]


def test_shards_partition_and_rotate():
    for total, world in ((65536, 8), (1001, 3), (5, 8)):
        cover = []
        for r in range(world):
            lo, hi = shard_env_range(total, r, world)
            cover += list(range(lo, hi))
        assert cover == list(range(total))
    a, b = shard_env_to_map(4096, 0, 100), shard_env_to_map(4096, 1, 100)
    assert a.min() == 0 and a.max() == 99 and not np.array_equal(a, b)
    assert np.bincount(a, minlength=100).min() >= 40


def test_single_process_passthrough():
    t, w, s = reduce_rollout([1.5], [2.0], torch.tensor([1, 2, 3, 4]))
    assert t == [1.5] and w == [2.0] and s.tolist() == [1, 2, 3, 4]
