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


def _same_trace(got, want):
    assert (got.rc, got.n_lines, got.n_iters, got.num_x, got.num_b) == (want.rc, want.n_lines, want.n_iters, want.num_x, want.num_b)
    for g, w in zip(got.iters, want.iters):
        assert (g["votes"], g["argmax"], g["status"]) == (w["votes"], w["argmax"], w["status"])
        for key in ("a_raw", "b_raw", "a1", "b1", "a2", "b2", "inliers1", "inliers2"):
            assert np.array_equal(g[key], w[key]), key


def test_device_sharded_loop_single_rank_comm(h3d):
    """h3d_hough3dlines_sharded with a one-rank NCCL communicator: the in-stream all-reduce path on one GPU."""
    pts = _cloud()
    want = h3d.hough3dlines(pts, G, 0.0, MINVOTES, NLINES, mode=1)
    comm = h3d.Comm(h3d.Comm.unique_id(), 0, 1)
    got = h3d.hough3dlines(pts, G, 0.0, MINVOTES, NLINES, mode=1, comm=comm)
    comm.close()
    _same_trace(got, want)
    big = synth.steel_tower(20000, seed=3)          # beyond the register-resident form on both sides
    want = h3d.hough3dlines(big, 5, 0.0, MINVOTES, 0, mode=1)            # unbounded line count
    comm = h3d.Comm(h3d.Comm.unique_id(), 0, 1)
    got = h3d.hough3dlines(big, 5, 0.0, MINVOTES, 0, mode=1, comm=comm)
    comm.close()
    _same_trace(got, want)


def _dev_worker(rank, world, uid_q, out):
    from pointcloud_segmentation_b200 import capi
    torch.cuda.set_device(rank)
    uid = uid_q.get(timeout=120)
    comm = capi.Comm(uid, rank, world)
    res = []
    for pts, g, nlines in ((_cloud(), G, NLINES), (synth.steel_tower(20000, seed=3), 5, 0)):
        tr = capi.hough3dlines(pts, g, 0.0, MINVOTES, nlines, mode=1, comm=comm)
        res.append([(it["votes"], it["argmax"], it["status"], it["a2"].tobytes(), it["b2"].tobytes(),
                     it["inliers2"].tobytes()) for it in tr.iters])
    comm.close()
    out.put((rank, res))


def test_device_sharded_loop_over_nccl(h3d):
    """Two ranks, two GPUs: every rank's records equal the single-GPU loop's, bit for bit."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    want = []
    for pts, g, nlines in ((_cloud(), G, NLINES), (synth.steel_tower(20000, seed=3), 5, 0)):
        tr = h3d.hough3dlines(pts, g, 0.0, MINVOTES, nlines, mode=1)
        want.append([(it["votes"], it["argmax"], it["status"], it["a2"].tobytes(), it["b2"].tobytes(),
                      it["inliers2"].tobytes()) for it in tr.iters])
    world = 2
    ctx = mp.get_context("spawn")
    out, uid_qs = ctx.Queue(), [ctx.Queue() for _ in range(world)]
    procs = [ctx.Process(target=_dev_worker, args=(r, world, uid_qs[r], out)) for r in range(world)]
    for p in procs:
        p.start()
    uid = h3d.Comm.unique_id()
    for q in uid_qs:
        q.put(uid)
    got = [out.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for _, res in got:
        assert res == want
