"""World-size-2 gloo tests (CPU) of the multi-GPU host logic: direction shards + the 8-byte MAX
all-reduce that replaces Hough::getLine, and frame ownership of the batched replay.  The shard
winners come from the ORACLE here (test infrastructure); on the GPU box the same ShardedHough
class sits on h3d_create_sharded (tests/test_sharding_gpu.py)."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


class _OracleShard:
    """Oracle Hough restricted to directions [j0, j1): key of the best cell of the shard."""

    def __init__(self, orc, g, mn, mx, dx, j0, j1):
        self.h = orc.hough(g, mn, mx, dx)
        self.j0, self.j1 = j0, j1

    def add(self, xyz):
        self.h.add(xyz)

    def subtract(self, xyz):
        self.h.subtract(xyz)

    def get_line_key(self):
        v = self.h.votes().reshape(-1, self.h.num_b)[:, self.j0:self.j1]
        if v.size == 0 or v.max() == 0:
            return 0
        m = int(v.max())
        cell, j = np.nonzero(v == m)
        flat = int((cell.astype(np.int64) * self.h.num_b + j + self.j0).min())
        return (m << 32) | (0xFFFFFFFF - flat)


def _worker(rank, world, port, out):
    import torch.distributed as dist
    from oracle import oracle as O
    from pointcloud_segmentation_b200 import sharding, synth
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    orc = O.Oracle()
    pts = synth.steel_tower(600, seed=5)
    x, _ = orc.shift_to_origin(pts)
    mn, mx = orc.min_max(x)
    g = 3
    num_b = len(orc.sphere(g))
    j0, j1 = sharding.direction_shards(num_b, world)[rank]
    sh = sharding.ShardedHough(mn, mx, 0.0, rank, world, num_b, backend=_OracleShard(orc, g, mn, mx, 0.0, j0, j1))
    sh.add(x)
    k1 = sh.get_line_key()
    sh.subtract(x[:150])
    k2 = sh.get_line_key()
    out.put((rank, k1, k2))
    dist.barrier()
    dist.destroy_process_group()


def test_direction_sharded_getline_over_gloo():
    from oracle import oracle as O
    from pointcloud_segmentation_b200 import sharding, synth
    O.build(ref=False)
    world, port = 2, 29611 + os.getpid() % 200
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    got = [out.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    orc = O.Oracle()
    pts = synth.steel_tower(600, seed=5)
    x, _ = orc.shift_to_origin(pts)
    mn, mx = orc.min_max(x)
    h = orc.hough(3, mn, mx, 0.0)
    h.add(x)
    v1, i1, _, _ = h.get_line()
    h.subtract(x[:150])
    v2, i2, _, _ = h.get_line()
    for _, k1, k2 in got:                       # every rank sees the global winner, ties included
        assert sharding.unpack_key(k1) == (v1, i1)
        assert sharding.unpack_key(k2) == (v2, i2)


def test_shard_bookkeeping():
    from pointcloud_segmentation_b200 import sharding
    for nb, w in ((20481, 8), (1281, 3), (6, 8), (21, 4)):
        sh = sharding.direction_shards(nb, w)
        assert sh[0][0] == 0 and sh[-1][1] == nb
        assert all(a[1] == b[0] for a, b in zip(sh, sh[1:]))
        assert max(j1 - j0 for j0, j1 in sh) - min(j1 - j0 for j0, j1 in sh) <= 1
    owned = sorted(i for r in range(8) for i in sharding.frames_for_rank(1000, r, 8))
    assert owned == list(range(1000))
    assert sharding.unpack_key(sharding.pack_key(37, 123456)) == (37, 123456)
    assert sharding.unpack_key(0) == (0, 0)
    assert sharding.pack_key(5, 10) > sharding.pack_key(5, 11) > sharding.pack_key(4, 0)   # lowest index wins ties
