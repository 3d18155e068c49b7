"""sensor_msgs/PointCloud2 adapter (node.cpp:286-289, 415): the message bytes
go to the device as they are; results must equal the plain (n, 4) float frame, bit for bit."""
import numpy as np
import pytest

from pointcloud_segmentation_b200 import synth

pytestmark = pytest.mark.gpu


def _message(frame, point_step, offs, width, height, row_pad=0, fill=0xA5):
    """Byte buffer of a PointCloud2 with FLOAT32 x,y,z at byte offsets `offs`; other bytes are junk."""
    n = frame.shape[0]
    assert n == width * height
    row_step = width * point_step + row_pad
    buf = np.full(height * row_step, fill, dtype=np.uint8)
    xyz = np.ascontiguousarray(frame[:, :3], dtype="<f4")
    for i in range(3):
        col = xyz[:, i].copy().view(np.uint8).reshape(n, 4)
        idx = (np.arange(n) // width) * row_step + (np.arange(n) % width) * point_step + offs[i]
        for b in range(4):
            buf[idx + b] = col[:, b]
    return buf, row_step


def _records(pipe):
    r = pipe.res
    import ctypes as C
    it = bytes(pipe.iters)[: r.n_iters * C.sizeof(type(pipe.iters[0]))]
    return (r.rc, r.n_iters, r.n_lines, r.n_input, tuple(r.shift), r.pool_used, it, pipe.pool[: r.pool_used].tobytes())


@pytest.mark.parametrize("prefilter", [True, False], ids=["k0", "nofilter"])
@pytest.mark.parametrize("name,point_step,offs,row_pad", [
    ("pcl_xyz", 16, (0, 4, 8), 0),            # pcl::PointXYZ: read in place, aligned quadruples
    ("xyz_after_intensity", 32, (4, 8, 12), 0),   # read in place, unaligned start
    ("row_padding", 16, (0, 4, 8), 32),       # organised cloud with padded rows: gather
    ("scattered", 24, (0, 8, 20), 0),         # x,y,z not adjacent: gather
])
def test_cloud2_matches_plain_frame(h3d, name, point_step, offs, row_pad, prefilter):
    frame = synth.tof_frame(pose=(-1.1, 0.0, 0.8, 0.0), seed=4).astype(np.float32)
    if not prefilter:
        frame = frame[::9][: 4000].copy()      # the loop on the raw points (wrap.h:175-189 widening only)
        width, height = frame.shape[0] // 8, 8
        frame = frame[: width * height]
    else:
        width, height = synth.TOF_W, synth.TOF_H
    pipe = h3d.Pipeline(4, synth.WINDOW, synth.LEAF, synth.OPT_DX if prefilter else 0.06, 12, 8,
                        max_raw_points=frame.shape[0], prefilter=prefilter)
    pipe.run(frame)
    want = _records(pipe)
    buf, row_step = _message(frame, point_step, offs, width, height, row_pad)
    for _ in range(2):
        pipe.run_cloud2(buf, width, height, point_step, row_step, *offs)
        assert _records(pipe) == want
    assert want[2] > 0
    pipe.close()


def test_cloud2_rejects_what_it_cannot_read(h3d):
    frame = synth.tof_frame().astype(np.float32)
    pipe = h3d.Pipeline(4, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 4, max_raw_points=frame.shape[0])
    buf, row_step = _message(frame, 16, (0, 4, 8), synth.TOF_W, synth.TOF_H)
    for kwargs in (dict(is_bigendian=1),):
        with pytest.raises(h3d.H3dError):
            pipe.run_cloud2(buf, synth.TOF_W, synth.TOF_H, 16, row_step, 0, 4, 8, **kwargs)
    with pytest.raises(h3d.H3dError):
        pipe.run_cloud2(buf, synth.TOF_W, synth.TOF_H, 16, row_step, 0, 4, 14)        # misaligned / outside point_step
    with pytest.raises(h3d.H3dError):
        pipe.run_cloud2(buf[:-4], synth.TOF_W, synth.TOF_H, 16, row_step, 0, 4, 8)    # short buffer
    with pytest.raises(h3d.H3dError):
        pipe.run_cloud2(buf, synth.TOF_W, synth.TOF_H, 16, row_step - 4, 0, 4, 8)     # row_step < width * point_step
    with pytest.raises(h3d.H3dError):
        pipe.run_cloud2(buf, synth.TOF_W * 2, synth.TOF_H, 16, row_step * 2, 0, 4, 8)  # more points than max_raw_points
    empty = pipe.run_cloud2(np.zeros(0, np.uint8), 0, 0, 16, 0, 0, 4, 8)
    assert empty.rc == 1 and empty.n_lines == 0
    pipe.close()
