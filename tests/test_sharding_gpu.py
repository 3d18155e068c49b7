"""Direction-sharded accumulator on the device: ShardedHough over h3d_create_sharded,
winner combined with an 8-byte MAX all-reduce (NCCL when two GPUs are present)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pointcloud_segmentation_b200 import sharding, synth  # noqa: E402

pytestmark = pytest.mark.gpu
G, MINVOTES, NLINES = 4, 10, 6


def _cloud():
    return synth.steel_tower(3000, seed=12)


def _reference_lines(h3d, pts):
    tr = h3d.hough3dlines(pts, G, 0.0, MINVOTES, NLINES, mode=1)
    return [(it["votes"], it["argmax"], len(it["inliers2"])) for it in tr.iters if it["status"] == 0], tr


def test_sharded_loop_single_rank_matches_device_loop(h3d):
    pts = _cloud()
    want, tr = _reference_lines(h3d, pts)
    got = sharding.hough3dlines_sharded(pts, G, tr.dx, MINVOTES, NLINES, rank=0, world=1)
    assert [(v, f, n) for v, f, _, _, n in got] == want
    for (_, _, a, b, _), it in zip(got, [i for i in tr.iters if i["status"] == 0]):
        assert np.array_equal(a, it["a2"]) and np.array_equal(b, it["b2"])


def _worker(rank, world, port, dx, out):
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    lines = sharding.hough3dlines_sharded(_cloud(), G, dx, MINVOTES, NLINES, rank, world,
                                          device=torch.device("cuda", rank))
    out.put((rank, [(v, f, n) for v, f, _, _, n in lines]))
    dist.barrier()
    dist.destroy_process_group()


def test_direction_sharded_loop_over_nccl(h3d):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    want, tr = _reference_lines(h3d, _cloud())
    world, port = 2, 29811 + os.getpid() % 100
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, tr.dx, out)) for r in range(world)]
    for p in procs:
        p.start()
    got = [out.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for _, lines in got:
        assert lines == want
