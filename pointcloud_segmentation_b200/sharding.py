"""Multi-GPU host logic: one process per GPU, torch.distributed for the plumbing.

Two ways the path shards, neither needs a bulk collective:
  * FRAMES are independent (one Hough per frame, wrap.h:220/347): frame i goes to rank
    i mod world, nothing is exchanged (`frames_for_rank`).
  * DIRECTIONS: a vote for direction j touches only plane j (lib/hough.cpp:90), so rank r owns
    directions [j0, j1) of the sphere, every rank holds all points, and Hough::getLine becomes
    the maximum over ranks of the packed key (votes << 32 | 0xFFFFFFFF - flat_index), one
    8-byte MAX all-reduce per iteration (`ShardedHough`).  The line fit (inliers, LSQ, remove) is
    computed redundantly and deterministically by every rank, so it needs no collective.
All compute goes through the C-ABI (capi); there is no CPU fallback here either.
"""
from __future__ import annotations

import numpy as np


def direction_shards(num_b: int, world: int):
    """[j0, j1) per rank; ragged when world does not divide num_b, empty ranges when world > num_b."""
    return [(num_b * r // world, num_b * (r + 1) // world) for r in range(world)]


def frames_for_rank(n_frames: int, rank: int, world: int):
    """Round-robin frame ownership of the batched replay (BASELINE configs[2])."""
    return list(range(rank, n_frames, world))


def pack_key(votes: int, flat: int) -> int:
    return (int(votes) << 32) | (0xFFFFFFFF - int(flat))


def unpack_key(key: int):
    votes = int(key) >> 32
    return votes, (0xFFFFFFFF - (int(key) & 0xFFFFFFFF)) if votes else 0


def allreduce_key(key: int, group=None, device=None) -> int:
    """Global Hough::getLine winner: MAX over ranks of the packed key (ties -> lowest flat index,
    lib/hough.cpp:107-112).  votes < 2^31, so the key is a non-negative int64."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return int(key)
    t = torch.tensor([int(key)], dtype=torch.int64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return int(t.item())


class ShardedHough:
    """`Hough` whose direction sphere is split over the ranks of a process group.

    get_key: callable returning this rank's packed shard winner; the default builds a
    capi.Hough over [j0, j1) (h3d_create_sharded).  Tests inject an oracle-backed callable to
    exercise the collective logic on CPU (gloo)."""

    def __init__(self, minP, maxP, dx, rank, world, num_b, group=None, device=None, backend=None):
        self.rank, self.world, self.group, self.device = rank, world, group, device
        self.j0, self.j1 = direction_shards(num_b, world)[rank]
        self.backend = backend
        if backend is None and self.j1 > self.j0:
            from . import capi
            self.backend = capi.Hough(minP, maxP, dx, shard=(self.j0, self.j1))
        self.ref = self.backend            # any shard can do the inverse map (it depends on j only)

    def add(self, xyz):
        if self.backend is not None:
            self.backend.add(xyz)

    def subtract(self, xyz):
        if self.backend is not None and len(xyz):
            self.backend.subtract(xyz)

    def get_line_key(self) -> int:
        local = self.backend.get_line_key() if self.backend is not None else 0   # empty shard: key 0
        return allreduce_key(local, self.group, self.device)

    def get_line(self):
        """(votes, a, b, flat) of the global winner, identical on every rank."""
        votes, flat = unpack_key(self.get_line_key())
        a, b = self.ref.line_from_index(flat)
        return votes, a, b, flat

    def close(self):
        if self.backend is not None and hasattr(self.backend, "close"):
            self.backend.close()


def hough3dlines_sharded(points, granularity, opt_dx, minvotes, nlines, rank, world, group=None, device=None):
    """The iterative loop of wrap.h:172-259,339-347 with a direction-sharded accumulator.
    Returns [(votes, flat, a2 + shift, b2, n_inliers)] -- identical on every rank."""
    from . import capi
    capi.init_space(granularity)
    x = np.ascontiguousarray(points, dtype=np.float64).reshape(-1, 3)
    x = x[np.isfinite(x).all(axis=1)]
    mn, mx = capi.min_max(x)
    d = float(np.sqrt(((mx - mn) ** 2).sum()))
    if d == 0.0 or opt_dx >= d:
        return []
    x, shift = capi.shift_to_origin(x)
    mn, mx = capi.min_max(x)
    num_b = len(capi.sphere(granularity)[0])
    h = ShardedHough(mn, mx, opt_dx, rank, world, num_b, group, device)
    h.add(x)
    lines, y = [], np.zeros((0, 3))
    while True:
        h.subtract(y)                                                  # wrap.h:241
        votes, a, b, flat = h.get_line()                               # wrap.h:243
        y = x[capi.points_close_to_line(x, a, b, opt_dx)]              # wrap.h:245
        rc, a, b = capi.orthogonal_lsq(y)                              # wrap.h:247
        if rc == 0.0:
            break
        y = x[capi.points_close_to_line(x, a, b, opt_dx)]              # wrap.h:250
        if len(y) < minvotes:
            break
        rc, a, b = capi.orthogonal_lsq(y)                              # wrap.h:254
        if rc == 0.0:
            break
        lines.append((votes, flat, a + shift, b, len(y)))              # wrap.h:257
        x = capi.remove_points(x, y)                                   # wrap.h:339
        if not (len(x) > 1 and (nlines == 0 or nlines > len(lines))):  # wrap.h:341-342
            break
    h.close()
    return lines
