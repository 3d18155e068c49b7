#!/usr/bin/env python
"""BASELINE.json configs[3]: 1 M-point infrastructure cloud, g6, accumulator sharded by DIRECTION over the
ranks.  Run with torchrun (one rank per GPU) or plainly for one GPU:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
      tools/c4_sharded.py [--nlines 50] [--points 1000000]

Every rank holds all points and directions [j0, j1); Hough::getLine = 8-byte MAX all-reduce (NCCL).  Prints
one JSON line from rank 0: per-shard add time (max over ranks), the all-reduce latency, the loop time and a
digest of the detected lines (identical on every rank and for every N)."""
import argparse
import hashlib
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pointcloud_segmentation_b200 import capi as h3d, sharding, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--nlines", type=int, default=50)
ap.add_argument("--points", type=int, default=1_000_000)
ap.add_argument("--skip-host-loop", action="store_true", help="do not time the host-driven ShardedHough loop (6 s)")
args = ap.parse_args()
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


def max_over_ranks(v):
    if world == 1:
        return v
    t = torch.tensor([v], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


pts = synth.infrastructure(args.points)
h3d.init_space(6)
x, shift = h3d.shift_to_origin(pts)
mn, mx = h3d.min_max(x)
d = float(np.sqrt(((mx - mn) ** 2).sum()))
dx = d / 64.0
j0, j1 = sharding.direction_shards(20481, world)[rank]
# vote kernel of this rank's shard, CUDA events inside the library
hg = h3d.Hough(mn, mx, dx, shard=(j0, j1))
hg.add(x)
us = hg.bench_vote(reps=5)
hg.close()
vote_us = max_over_ranks(us)
# the 8-byte MAX all-reduce that replaces getLine's global scan
ar_us = 0.0
if world > 1:
    t = torch.zeros(1, dtype=torch.int64, device=dev)
    for _ in range(20):
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(200):
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e1.record()
    torch.cuda.synchronize()
    ar_us = max_over_ranks(e0.elapsed_time(e1) * 1e3 / 200)
# the whole loop, direction-sharded
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
t0 = time.perf_counter()
lines = None if args.skip_host_loop else sharding.hough3dlines_sharded(pts, 6, dx, 10, args.nlines, rank, world, device=dev)
torch.cuda.synchronize()
loop_ms = None if args.skip_host_loop else max_over_ranks((time.perf_counter() - t0) * 1e3)
# the same loop DEVICE-RESIDENT: h3d_hough3dlines_sharded, the all-reduce enqueued on the loop's stream
comm = None
if world > 1:
    uid = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        uid = torch.tensor(list(h3d.Comm.unique_id()), dtype=torch.uint8, device=dev)
    dist.broadcast(uid, 0)
    comm = h3d.Comm(bytes(uid.cpu().tolist()), rank, world)
dev_ms = []
for rep in range(3):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    tr = h3d.hough3dlines(pts, 6, 0.0, 10, args.nlines, mode=1, comm=comm)
    dev_ms.append(max_over_ranks((time.perf_counter() - t0) * 1e3))
dev_lines = [(it["votes"], it["argmax"], it["a2"], it["b2"], len(it["inliers2"])) for it in tr.iters if it["status"] == 0]
dev_same_as_host_loop = None if lines is None else len(dev_lines) == len(lines) and all(
    a[0] == b[0] and a[1] == b[1] and np.array_equal(a[2], b[2]) and np.array_equal(a[3], b[3]) and a[4] == b[4]
    for a, b in zip(dev_lines, lines))
lines = dev_lines
if comm is not None:
    comm.close()
dig = hashlib.sha256()
for v, flat, a, b, n in lines:
    dig.update(np.array([v, flat, n], dtype=np.uint64).tobytes() + np.asarray(a).tobytes() + np.asarray(b).tobytes())
digest = dig.hexdigest()[:16]
same = True
if world > 1:
    mine = torch.tensor(list(bytes.fromhex(digest)), dtype=torch.int64, device=dev)
    lo, hi = mine.clone(), mine.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    same = bool((lo == hi).all().item())
if rank == 0:
    votes = args.points * 20481
    print(json.dumps({"config": "configs[3]: 1M-point infrastructure cloud, g6, dx=d/64, direction-sharded",
                      "n_gpus": world, "points": args.points, "nlines": args.nlines, "directions_per_rank": j1 - j0,
                      "add_us_max_over_ranks": vote_us, "add_gvotes_per_s_whole_job": votes / vote_us / 1e3,
                      "argmax_allreduce_us": ar_us, "device_loop_ms_max_over_ranks": dev_ms,
                      "device_loop_same_lines_as_host_driven_loop": dev_same_as_host_loop,
                      "device_loop_votes_cast_rank0": tr.votes_cast,
                      "host_driven_loop_ms_max_over_ranks": loop_ms, "lines": len(lines),
                      "lines_digest": digest, "digest_identical_on_all_ranks": same,
                      "note": "device loop = h3d_hough3dlines_sharded (host points in, records out; one in-stream 8-byte ncclAllReduce per iteration); host-driven loop = Hough-class calls (h3d_add/subtract/get_line_key + host-array "
                              "K3 entry points), wall clock bracketed by device syncs and a barrier"}))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
