"""ROS-free replay of node.cpp:286-306 on the GPU: frames in flight -> device post-tests with one pose per
frame -> world model (segFiltering) in frame order, against the oracle chain (C restatement of the frame +
oracle/world.py) -- SURVEY 8(f) rank 4 on top of ranks 1-3.  World segments, intersection matrix and both CSV
files are compared bit for bit."""
import math

import numpy as np
import pytest

from oracle import world as OW
from pointcloud_segmentation_b200 import synth
from test_world_cpu import DIAG, MIN_W, _same_world

pytestmark = pytest.mark.gpu

G, MINVOTES, NLINES = 4, 12, 10
FLOOR = 0.3


def _cfg(mod):
    return mod.SegmentConfig.make(np.sqrt(3) * float(synth.LEAF), 0.95, 1.5, [0.03, 0.05])


def _pose(mod, p):
    x, y, z, yaw = (float(v) for v in p)
    return mod.Pose.make([math.cos(yaw / 2), 0.0, 0.0, math.sin(yaw / 2)], [x, y, z], FLOOR)


def _frames(n):
    poses = synth.tower_trajectory(1000)[10:10 + 6 * n:6]
    frames = [synth.tof_frame(pose=tuple(p), seed=500 + k).astype(np.float32) for k, p in enumerate(poses)]
    return poses, frames


def _oracle_world(oracle, oracle_mod, poses, frames):
    want = OW.World(DIAG, MIN_W)
    kept = []
    for p, f in zip(poses, frames):
        filt = oracle.prefilter(f, synth.WINDOW, float(synth.LEAF)).astype(np.float64)
        trace = oracle.hough3dlines(filt, G, synth.OPT_DX, MINVOTES, NLINES, mode=0)
        _, segs = oracle.segments(filt, trace, _cfg(oracle_mod), _pose(oracle_mod, p))
        drone = [OW.Seg(tuple(s.a_world), tuple(s.b_world), s.t_min, s.t_max, s.radius, s.n_points, s.pca_coeff,
                        tuple(s.pca_eigenvalues)) for s in segs if s.status == 0 and s.world_status == 1]
        kept.append(len(drone))
        want.update(drone)
    return want, kept


def test_replay_world_matches_oracle_chain(h3d, oracle, oracle_mod, tmp_path):
    poses, frames = _frames(10)
    want, kept = _oracle_world(oracle, oracle_mod, poses, frames)
    assert sum(kept) > len(want.segments) > 2                     # beams were seen again and fused
    # frames in flight, one pose per frame
    rep = h3d.Replay(3, 40, G, synth.WINDOW, synth.LEAF, synth.OPT_DX, MINVOTES, NLINES,
                     max_raw_points=frames[0].shape[0], max_frames=len(frames))
    rep.enable_segments(_cfg(h3d))
    world = h3d.World(DIAG, MIN_W)
    rep.run_world([f.ctypes.data for f in frames], [f.shape[0] for f in frames], frames[0].shape[1], False,
                  [_pose(h3d, p) for p in poses], world)
    _same_world(world, want)
    # the same replay with two host threads for the three pipelines (split-phase workers): poses and segment
    # records must follow their frames
    old_threads = h3d.set_replay_threads(2)
    try:
        world_mux = h3d.World(DIAG, MIN_W)
        rep.run_world([f.ctypes.data for f in frames], [f.shape[0] for f in frames], frames[0].shape[1], False,
                      [_pose(h3d, p) for p in poses], world_mux)
        _same_world(world_mux, want)
        world_mux.close()
    finally:
        h3d.set_replay_threads(old_threads)
    seg_csv, int_csv = tmp_path / "segments.csv", tmp_path / "intersections.csv"
    world.save_csv(str(seg_csv), str(int_csv))
    assert (seg_csv.read_text(), int_csv.read_text()) == want.csv()
    rep.close()
    # one frame at a time through a single pipeline, as the node's worker thread does
    pipe = h3d.Pipeline(G, synth.WINDOW, synth.LEAF, synth.OPT_DX, MINVOTES, NLINES, frames[0].shape[0])
    pipe.enable_segments(_cfg(h3d))
    world2 = h3d.World(DIAG, MIN_W)
    for k, (p, f) in enumerate(zip(poses, frames)):
        pipe.set_pose(_pose(h3d, p))
        pipe.run(f)
        _, segs = pipe.segments()
        assert world2.update_frame(segs) == kept[k]
    _same_world(world2, want)
    pipe.close()
    world.close()
    world2.close()
