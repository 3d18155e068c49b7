"""ctypes bindings of libh3d.so (include/h3d.h) for the test and bench harness.

The product is the C-ABI library plus the C++ drop-in headers; this module only
lets pytest / bench.py call the same entry points a C++ consumer would.  It
raises at import of the library if libh3d.so is missing -- there is no fallback.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libh3d.so")

_lib = None


class H3dError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"h3d error {code}: {msg}")
        self.code = code


class Iter(C.Structure):
    _fields_ = [
        ("votes", C.c_uint32), ("status", C.c_uint32), ("argmax", C.c_uint64),
        ("a_raw", C.c_double * 3), ("b_raw", C.c_double * 3),
        ("a1", C.c_double * 3), ("b1", C.c_double * 3), ("rc1", C.c_double),
        ("a2", C.c_double * 3), ("b2", C.c_double * 3), ("rc2", C.c_double),
        ("n1", C.c_uint32), ("n2", C.c_uint32), ("off1", C.c_uint32), ("off2", C.c_uint32),
    ]


class Result(C.Structure):
    _fields_ = [
        ("rc", C.c_int), ("n_iters", C.c_uint32), ("n_lines", C.c_uint32), ("n_input", C.c_uint32),
        ("shift", C.c_double * 3), ("d", C.c_double), ("dx", C.c_double), ("max_x", C.c_double),
        ("num_x", C.c_uint64), ("num_b", C.c_uint64), ("pool_used", C.c_uint64),
        ("votes_cast", C.c_uint64),
    ]


class PipelineConfig(C.Structure):
    _fields_ = [
        ("granularity", C.c_int), ("window", C.c_float * 6), ("leaf", C.c_float),
        ("opt_dx", C.c_double), ("opt_minvotes", C.c_int), ("opt_nlines", C.c_int),
        ("max_raw_points", C.c_size_t), ("max_iters", C.c_size_t), ("prefilter", C.c_int),
    ]


class Cloud2Layout(C.Structure):
    """h3d_cloud2_layout: sensor_msgs/PointCloud2 geometry (width, height, steps, x/y/z byte offsets)."""
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("point_step", C.c_uint32), ("row_step", C.c_uint32),
                ("off_x", C.c_uint32), ("off_y", C.c_uint32), ("off_z", C.c_uint32), ("is_bigendian", C.c_uint32)]


class SegmentConfig(C.Structure):
    """h3d_segment_config: arguments of the post-tests (wrap.h:167-170)."""
    _fields_ = [("diag_voxel", C.c_double), ("min_pca_coeff", C.c_double), ("rad_2_leaf_ratio", C.c_double),
                ("n_radius", C.c_uint32), ("pad_", C.c_uint32), ("radius_sizes", C.c_double * 8)]

    @classmethod
    def make(cls, diag_voxel, min_pca_coeff, rad_2_leaf_ratio, radius_sizes):
        c = cls()
        c.diag_voxel, c.min_pca_coeff, c.rad_2_leaf_ratio = diag_voxel, min_pca_coeff, rad_2_leaf_ratio
        c.n_radius = len(radius_sizes)
        for k, r in enumerate(radius_sizes[:8]):
            c.radius_sizes[k] = r
        return c


class Segment(C.Structure):
    """h3d_segment: one tested line; status 0 = pushed to computed_lines (wrap.h:324-335)."""
    _fields_ = [("iter", C.c_uint32), ("status", C.c_uint32), ("n_points", C.c_uint32), ("min_nb_points", C.c_int32),
                ("a", C.c_double * 3), ("b", C.c_double * 3), ("t_min", C.c_double), ("t_max", C.c_double),
                ("radius_raw", C.c_double), ("radius", C.c_double), ("max_dist", C.c_double),
                ("pca_coeff", C.c_double), ("pca_eigenvalues", C.c_double * 3), ("length", C.c_double),
                ("a_world", C.c_double * 3), ("b_world", C.c_double * 3), ("world_status", C.c_uint32),
                ("pad_", C.c_uint32)]

    def as_tuple(self):
        return (self.iter, self.status, self.n_points, self.min_nb_points, tuple(self.a), tuple(self.b), self.t_min,
                self.t_max, self.radius_raw, self.radius, self.max_dist, self.pca_coeff, tuple(self.pca_eigenvalues),
                self.length, tuple(self.a_world), tuple(self.b_world), self.world_status)


class Pose(C.Structure):
    """Sensor pose for drone2WorldSeg / heighSegmentCutoff (node.cpp:429-470): q = (w, x, y, z)."""
    _fields_ = [("q", C.c_double * 4), ("position", C.c_double * 3), ("floor_trim_height", C.c_double)]

    @classmethod
    def make(cls, q_wxyz, position, floor_trim_height):
        p = cls()
        for k in range(4):
            p.q[k] = float(q_wxyz[k])
        for k in range(3):
            p.position[k] = float(position[k])
        p.floor_trim_height = float(floor_trim_height)
        return p


def lib():
    """Load libh3d.so (built in-tree by `make -C pointcloud_segmentation_b200`)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `make -C {HERE}` "
                "(or __graft_entry__.build()); there is no CPU fallback")
        _lib = C.CDLL(LIB_PATH)
        _lib.h3d_last_error.restype = C.c_char_p
    return _lib


def check(rc):
    if rc != 0:
        raise H3dError(rc, lib().h3d_last_error().decode(errors="replace"))


def _d3(v):
    return (C.c_double * 3)(*[float(x) for x in v])


def _pts(xyz):
    a = np.ascontiguousarray(xyz, dtype=np.float64).reshape(-1, 3)
    return a, a.ctypes.data_as(C.POINTER(C.c_double)), C.c_size_t(a.shape[0])


def device_info():
    sm, cc, smem = C.c_int(), C.c_int(), C.c_size_t()
    check(lib().h3d_device_info(C.byref(sm), C.byref(cc), C.byref(smem)))
    return {"sm_count": sm.value, "cc": cc.value, "smem_optin": smem.value}


def init_space(g):
    check(lib().h3d_init_space(C.c_uint(g)))


def sphere(g):
    nv, nt = C.c_size_t(), C.c_size_t()
    check(lib().h3d_sphere(C.c_int(g), None, C.c_size_t(0), C.byref(nv), None, C.c_size_t(0), C.byref(nt)))
    v = np.empty((nv.value, 3), dtype=np.float64)
    t = np.empty(nt.value, dtype=np.uint32)
    check(lib().h3d_sphere(C.c_int(g), v.ctypes.data_as(C.POINTER(C.c_double)), nv, C.byref(nv),
                           t.ctypes.data_as(C.POINTER(C.c_uint)), nt, C.byref(nt)))
    return v, t


class Hough:
    """Mirror of the reference's `Hough` class (lib/hough.h:21-46) over the C-ABI."""

    def __init__(self, minP, maxP, dx, shard=None):
        self._h = C.c_void_p()
        if shard is None:
            check(lib().h3d_create(_d3(minP), _d3(maxP), C.c_double(dx), C.byref(self._h)))
        else:
            check(lib().h3d_create_sharded(_d3(minP), _d3(maxP), C.c_double(dx), C.c_size_t(shard[0]),
                                           C.c_size_t(shard[1]), C.byref(self._h)))
        nx, nb, dxx, mx = C.c_size_t(), C.c_size_t(), C.c_double(), C.c_double()
        check(lib().h3d_info(self._h, C.byref(nx), C.byref(nb), C.byref(dxx), C.byref(mx)))
        self.num_x, self.num_b, self.dx, self.max_x = nx.value, nb.value, dxx.value, mx.value

    def add(self, xyz):
        a, p, n = _pts(xyz)
        check(lib().h3d_add(self._h, p, n))

    def subtract(self, xyz):
        a, p, n = _pts(xyz)
        check(lib().h3d_subtract(self._h, p, n))

    def get_line(self):
        a, b = (C.c_double * 3)(), (C.c_double * 3)()
        votes, flat = C.c_uint(), C.c_uint64()
        check(lib().h3d_get_line(self._h, a, b, C.byref(votes), C.byref(flat)))
        return votes.value, np.array(a[:]), np.array(b[:]), flat.value

    def get_line_key(self):
        k = C.c_uint64()
        check(lib().h3d_get_line_key(self._h, C.byref(k)))
        return k.value

    def line_from_index(self, flat):
        """Inverse map of Hough::getLine (lib/hough.cpp:114-132) for a flat accumulator index."""
        a, b = (C.c_double * 3)(), (C.c_double * 3)()
        check(lib().h3d_line_from_index(self._h, C.c_uint64(flat), a, b))
        return np.array(a[:]), np.array(b[:])

    def votes(self):
        cells = self.num_x * self.num_x * self.num_b
        out = np.empty(cells, dtype=np.uint32)
        check(lib().h3d_download_votes(self._h, out.ctypes.data_as(C.POINTER(C.c_uint)), C.c_size_t(cells)))
        return out

    def bench_vote(self, reps=10, ablate=0, D=0):
        us = C.c_float()
        check(lib().h3d_bench_vote(self._h, C.c_int(reps), C.c_int(ablate | (D << 8)), C.byref(us)))
        return us.value

    def close(self):
        if self._h:
            lib().h3d_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Trace:
    """Result of one hough3dlines() call: header fields + per-iteration records with inlier lists."""

    def __init__(self, res, iters, pool):
        self.rc, self.n_lines, self.n_input = res.rc, res.n_lines, res.n_input
        self.n_iters = res.n_iters
        self.shift = np.array(res.shift[:])
        self.d, self.dx, self.max_x = res.d, res.dx, res.max_x
        self.num_x, self.num_b = res.num_x, res.num_b
        self.votes_cast = res.votes_cast
        self.iters = []
        for r in iters[: min(res.n_iters, len(iters))]:
            self.iters.append({
                "votes": r.votes, "status": r.status, "argmax": r.argmax,
                "a_raw": np.array(r.a_raw[:]), "b_raw": np.array(r.b_raw[:]),
                "a1": np.array(r.a1[:]), "b1": np.array(r.b1[:]), "rc1": r.rc1,
                "a2": np.array(r.a2[:]), "b2": np.array(r.b2[:]), "rc2": r.rc2,
                "inliers1": pool[r.off1:r.off1 + r.n1].copy(),
                "inliers2": pool[r.off2:r.off2 + r.n2].copy(),
            })


class Comm:
    """h3d_comm: NCCL communicator of the direction-sharded loop (one rank per GPU).  Rank 0 calls
    Comm.unique_id() and hands the 128 bytes to the other ranks; Comm(id, rank, world) is collective."""

    @staticmethod
    def unique_id():
        buf = (C.c_ubyte * 128)()
        check(lib().h3d_comm_unique_id(buf))
        return bytes(buf)

    def __init__(self, uid, rank, world):
        self.rank, self.world = rank, world
        self._c = C.c_void_p()
        check(lib().h3d_comm_create((C.c_ubyte * 128)(*uid), C.c_int(rank), C.c_int(world), C.byref(self._c)))

    def close(self):
        if self._c:
            lib().h3d_comm_destroy(self._c)
            self._c = C.c_void_p()


def hough3dlines(xyz, granularity, dx, minvotes, nlines, mode=0, max_iters=None, comm=None):
    """The whole loop on the device (wrap.h:167-350 without the segment post-tests).  comm: a Comm to
    shard the accumulator by direction over its ranks (h3d_hough3dlines_sharded; every rank gets the
    same Trace)."""
    a, p, n = _pts(xyz)
    if max_iters is None:
        max_iters = (nlines + 1) if nlines > 0 else 256
    iters = (Iter * max_iters)()
    pool_cap = max(1, 2 * max_iters * a.shape[0])
    pool = np.zeros(pool_cap, dtype=np.uint32)
    res = Result()
    check(lib().h3d_hough3dlines_sharded(p, n, C.c_int(granularity), C.c_double(dx), C.c_int(minvotes),
                                         C.c_int(nlines), C.c_int(mode), comm._c if comm is not None else None,
                                         iters, C.c_size_t(max_iters), pool.ctypes.data_as(C.POINTER(C.c_uint32)),
                                         C.c_size_t(pool_cap), C.byref(res)))
    return Trace(res, iters, pool)


def hough3dlines_timing():
    """[upload+prep, first vote, loop kernels, whole call] in microseconds on the device (CUDA events) for this
    thread's last hough3dlines() call."""
    us = (C.c_float * 4)()
    check(lib().h3d_hough3dlines_timing(us))
    return list(us)


def prefilter(xyz_f32, window, leaf):
    """cloudFiltering of node.cpp:386-421: PassThrough window + VoxelGrid centroids (f32)."""
    a = np.ascontiguousarray(xyz_f32, dtype=np.float32)
    a = a.reshape(-1, a.shape[-1] if a.ndim > 1 else 3)
    n, stride = a.shape
    out = np.empty((max(n, 1), 3), dtype=np.float32)
    n_out = C.c_size_t()
    win = (C.c_float * 6)(*[float(v) for v in window])
    check(lib().h3d_prefilter(a.ctypes.data_as(C.POINTER(C.c_float)), C.c_size_t(n), C.c_size_t(stride), win,
                              C.c_float(float(leaf)), out.ctypes.data_as(C.POINTER(C.c_float)),
                              C.c_size_t(out.shape[0]), C.byref(n_out)))
    return out[: n_out.value].copy()


class Pipeline:
    """Persistent per-frame pipeline: raw f32 frame -> K0 -> Hough loop (h3d_pipeline_*)."""

    def __init__(self, granularity, window, leaf, opt_dx, minvotes, nlines, max_raw_points, prefilter=True,
                 max_iters=0):
        cfg = PipelineConfig()
        cfg.granularity = granularity
        cfg.window = (C.c_float * 6)(*[float(v) for v in window])
        cfg.leaf = float(leaf)
        cfg.opt_dx = float(opt_dx)
        cfg.opt_minvotes, cfg.opt_nlines = int(minvotes), int(nlines)
        cfg.max_raw_points = int(max_raw_points)
        cfg.max_iters = int(max_iters)
        cfg.prefilter = 1 if prefilter else 0
        self._p = C.c_void_p()
        check(lib().h3d_pipeline_create(C.byref(cfg), C.byref(self._p)))
        self.max_iters = max_iters or ((nlines + 1) if nlines > 0 else 64)
        self.iters = (Iter * self.max_iters)()
        self.pool = np.zeros(2 * self.max_iters * int(max_raw_points), dtype=np.uint32)
        self.res = Result()

    def _run(self, fn, ptr, n, stride):
        check(fn(self._p, ptr, C.c_size_t(n), C.c_size_t(stride), self.iters,
                 self.pool.ctypes.data_as(C.POINTER(C.c_uint32)), C.c_size_t(self.pool.size), C.byref(self.res)))
        return self.res

    def run(self, raw_f32):
        """Host frame in, host results out.  Returns the ctypes Result (no unpacking)."""
        a = raw_f32 if (isinstance(raw_f32, np.ndarray) and raw_f32.dtype == np.float32 and raw_f32.flags.c_contiguous) \
            else np.ascontiguousarray(raw_f32, dtype=np.float32)
        n, stride = a.shape
        return self._run(lib().h3d_pipeline_run, a.ctypes.data_as(C.POINTER(C.c_float)), n, stride)

    def run_cloud2(self, data, width, height, point_step, row_step, off_x, off_y, off_z, is_bigendian=0):
        """A sensor_msgs/PointCloud2 byte buffer (bytes / uint8 array) as it comes off the wire."""
        buf = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data
        lay = Cloud2Layout(width, height, point_step, row_step, off_x, off_y, off_z, is_bigendian)
        check(lib().h3d_pipeline_run_cloud2(self._p, buf.ctypes.data_as(C.POINTER(C.c_uint8)), C.c_size_t(buf.size),
                                            C.byref(lay), self.iters, self.pool.ctypes.data_as(C.POINTER(C.c_uint32)),
                                            C.c_size_t(self.pool.size), C.byref(self.res)))
        return self.res

    def run_ptr(self, host_ptr, n, stride):
        """Same with a raw host address (e.g. a pinned torch tensor's data_ptr())."""
        return self._run(lib().h3d_pipeline_run, C.cast(C.c_void_p(host_ptr), C.POINTER(C.c_float)), n, stride)

    def run_dev(self, dev_ptr, n, stride):
        """Frame already resident in device memory (dev_ptr: integer device address)."""
        return self._run(lib().h3d_pipeline_run_dev, C.cast(C.c_void_p(dev_ptr), C.POINTER(C.c_float)), n, stride)

    def trace(self):
        return Trace(self.res, self.iters, self.pool)

    def filtered(self, cap):
        out = np.empty((cap, 3), dtype=np.float32)
        n_out = C.c_size_t()
        check(lib().h3d_pipeline_filtered(self._p, out.ctypes.data_as(C.POINTER(C.c_float)), C.c_size_t(cap),
                                          C.byref(n_out)))
        return out[: min(cap, n_out.value)].copy(), n_out.value

    def timing_us(self):
        us = (C.c_float * 4)()
        check(lib().h3d_pipeline_timing(self._p, us))
        return list(us)

    def enable_segments(self, cfg):
        """Run the segment post-tests (wrap.h:261-337) with every following frame; None switches them off."""
        check(lib().h3d_pipeline_enable_segments(self._p, C.byref(cfg) if cfg is not None else None))

    def set_pose(self, pose):
        """Sensor pose for the drone -> world step of the post-tests (node.cpp:429-470); None: off."""
        check(lib().h3d_pipeline_set_pose(self._p, C.byref(pose) if pose is not None else None))

    def segments(self):
        """(rc of the reference's hough3dlines, [Segment]) of the last frame."""
        n = C.c_size_t(0)
        rc_ref = C.c_int(0)
        cap = int(self.max_iters)
        segs = (Segment * cap)()
        check(lib().h3d_pipeline_segments(self._p, segs, C.c_size_t(cap), C.byref(n), C.byref(rc_ref)))
        return rc_ref.value, [segs[k] for k in range(min(n.value, cap))]

    def k0_stamps(self):
        """globaltimer ns of the K0 cluster kernel's phases for the last (fast-path) frame"""
        t = (C.c_ulonglong * 10)()
        check(lib().h3d_pipeline_k0_stamps(self._p, t))
        return list(t)

    def set_sm_share(self, n_ctas):
        """SMs the persistent loop kernel of this pipeline may use (0 = all)."""
        check(lib().h3d_pipeline_set_sm_share(self._p, C.c_int(n_ctas)))

    def profile_enable(self, iters):
        """Measurement hook: record phase stamps of the fused loop kernel for `iters` iterations."""
        check(lib().h3d_pipeline_profile(self._p, C.c_int(iters), None, C.c_size_t(0), None, None))
        self._prof_iters = iters

    def profile_read(self):
        """[iteration][cta][slot] globaltimer ns of the last frame (slot 8: re-voted planes + 1)."""
        n_cta, n_slots = C.c_int(), C.c_int()
        sm = device_info()["sm_count"]
        buf = np.zeros(self._prof_iters * sm * 16, dtype=np.uint64)
        check(lib().h3d_pipeline_profile(self._p, C.c_int(0), buf.ctypes.data_as(C.POINTER(C.c_ulonglong)),
                                         C.c_size_t(buf.size), C.byref(n_cta), C.byref(n_slots)))
        k = self._prof_iters * n_cta.value * n_slots.value
        return buf[:k].reshape(self._prof_iters, n_cta.value, n_slots.value)

    def close(self):
        if self._p:
            lib().h3d_pipeline_destroy(self._p)
            self._p = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- PointCloud entry points (lib/pointcloud.cpp) on host arrays --------------------------
class Replay:
    """Batched replay of independent frames over K pipelines (h3d_replay_run): K host threads in the
    library, each pipeline limited to `sm_share` SMs for its persistent loop kernel."""

    def __init__(self, n_pipes, sm_share, granularity, window, leaf, opt_dx, minvotes, nlines, max_raw_points,
                 max_frames):
        self.pipes = [Pipeline(granularity, window, leaf, opt_dx, minvotes, nlines, max_raw_points)
                      for _ in range(n_pipes)]
        for p in self.pipes:
            p.set_sm_share(sm_share)
        self.max_iters = self.pipes[0].max_iters
        self.pool_cap = 2 * self.max_iters * int(max_raw_points) // 8     # inlier lists of one frame
        self.max_frames = max_frames
        self.iters = (Iter * (self.max_iters * max_frames))()
        self.pool = np.zeros(self.pool_cap * max_frames, dtype=np.uint32)
        self.results = (Result * max_frames)()
        self._handles = (C.c_void_p * n_pipes)(*[p._p for p in self.pipes])
        self.segments = None

    def enable_segments(self, cfg, pose=None):
        """Segment post-tests (+ world transform with a pose) on every pipeline; run() then fills
        self.segments / self.n_segments / self.rc_ref per frame."""
        for p in self.pipes:
            p.enable_segments(cfg)
            p.set_pose(pose)
        if cfg is None:
            self.segments = None
        else:
            self.segments = (Segment * (self.max_iters * self.max_frames))()
            self.n_segments = (C.c_uint32 * self.max_frames)()
            self.rc_ref = (C.c_int * self.max_frames)()

    def frame_segments(self, i):
        return [self.segments[i * self.max_iters + k] for k in range(min(self.n_segments[i], self.max_iters))]

    def run(self, ptrs, n_points, stride, on_device):
        """ptrs: integer addresses of the frames (host or device); returns the ctypes Result array."""
        n = len(ptrs)
        assert n <= self.max_frames
        raw = (C.c_void_p * n)(*ptrs)
        npts = (C.c_size_t * n)(*n_points)
        if self.segments is not None:
            check(lib().h3d_replay_run_segments(self._handles, C.c_size_t(len(self.pipes)), raw, npts, C.c_size_t(stride),
                                                C.c_size_t(n), C.c_int(1 if on_device else 0), self.iters,
                                                C.c_size_t(self.max_iters), self.pool.ctypes.data_as(C.POINTER(C.c_uint32)),
                                                C.c_size_t(self.pool_cap), self.results, self.segments,
                                                C.c_size_t(self.max_iters), self.n_segments, self.rc_ref))
            return self.results
        check(lib().h3d_replay_run(self._handles, C.c_size_t(len(self.pipes)), raw, npts, C.c_size_t(stride),
                                   C.c_size_t(n), C.c_int(1 if on_device else 0), self.iters,
                                   C.c_size_t(self.max_iters), self.pool.ctypes.data_as(C.POINTER(C.c_uint32)),
                                   C.c_size_t(self.pool_cap), self.results))
        return self.results

    def run_world(self, ptrs, n_points, stride, on_device, poses, world):
        """node.cpp:286-306 without ROS (h3d_replay_run_world): frame i with sensor pose poses[i] (Pose), the
        kept world-frame segments of every frame fused into `world` (World) in frame order."""
        n = len(ptrs)
        assert n <= self.max_frames and len(poses) == n and self.segments is not None
        raw = (C.c_void_p * n)(*ptrs)
        npts = (C.c_size_t * n)(*n_points)
        check(lib().h3d_replay_run_world(self._handles, C.c_size_t(len(self.pipes)), raw, npts, C.c_size_t(stride),
                                         C.c_size_t(n), C.c_int(1 if on_device else 0), (Pose * n)(*poses),
                                         self.results, self.segments, C.c_size_t(self.max_iters), self.n_segments,
                                         self.rc_ref, world._w))
        return self.results

    def close(self):
        for p in self.pipes:
            p.close()


def min_max(xyz):
    a, p, n = _pts(xyz)
    mn, mx = (C.c_double * 3)(), (C.c_double * 3)()
    check(lib().h3d_min_max(p, n, mn, mx))
    return np.array(mn[:]), np.array(mx[:])


def shift_to_origin(xyz, shift=None):
    a = np.array(xyz, dtype=np.float64).reshape(-1, 3)
    sh = _d3(shift if shift is not None else (0.0, 0.0, 0.0))
    check(lib().h3d_shift_to_origin(a.ctypes.data_as(C.POINTER(C.c_double)), C.c_size_t(a.shape[0]), sh))
    return a, np.array(sh[:])


def mean_value(xyz):
    a, p, n = _pts(xyz)
    m = (C.c_double * 3)()
    check(lib().h3d_mean_value(p, n, m))
    return np.array(m[:])


def points_close_to_line(xyz, a_, b_, dx):
    a, p, n = _pts(xyz)
    idx = np.empty(max(a.shape[0], 1), dtype=np.uint32)
    n_out = C.c_size_t()
    check(lib().h3d_points_close_to_line(p, n, _d3(a_), _d3(b_), C.c_double(dx),
                                         idx.ctypes.data_as(C.POINTER(C.c_uint32)), C.byref(n_out)))
    return idx[: n_out.value].copy()


NOALLOC_MAX_POINTS = 1048576   # H3D_NOALLOC_MAX_POINTS (include/h3d.h)


def points_close_to_line_noalloc(xyz_dev_ptr, n, a_, b_, dx, idx_out_dev_ptr, n_out_dev_ptr, stream=0):
    """h3d_points_close_to_line_noalloc: DEVICE pointers (ints) in and out, launches on `stream`, no
    synchronisation and no allocation -- the caller owns every buffer (include/h3d.h)."""
    check(lib().h3d_points_close_to_line_noalloc(C.c_void_p(xyz_dev_ptr), C.c_size_t(n), _d3(a_), _d3(b_),
                                                 C.c_double(dx), C.c_void_p(idx_out_dev_ptr),
                                                 C.c_void_p(n_out_dev_ptr), C.c_void_p(stream)))


def remove_points(x, y):
    xa = np.array(x, dtype=np.float64).reshape(-1, 3)
    ya, yp, ny = _pts(y)
    n_out = C.c_size_t()
    check(lib().h3d_remove_points(xa.ctypes.data_as(C.POINTER(C.c_double)), C.c_size_t(xa.shape[0]), yp, ny,
                                  C.byref(n_out)))
    return xa[: n_out.value].copy()


def bench_lsq(xyz):
    """device time (us) of the phases of one orthogonal_LSQ call: means (incl. the sequential anchor), scatter sums, eig"""
    a, p, n = _pts(xyz)
    us = (C.c_float * 5)()
    check(lib().h3d_bench_lsq(p, n, us))
    return list(us)


def orthogonal_lsq(xyz):
    a, p, n = _pts(xyz)
    aa, bb, rc = (C.c_double * 3)(), (C.c_double * 3)(), C.c_double()
    check(lib().h3d_orthogonal_lsq(p, n, aa, bb, C.byref(rc)))
    return rc.value, np.array(aa[:]), np.array(bb[:])


def segment_test(y_centred, shift, a_, b_, cfg, pose=None):
    """wrap.h:261-337 on one line from host buffers (+ node.cpp:429-470 with a pose) -> Segment."""
    _, p, n = _pts(y_centred)
    seg = Segment()
    check(lib().h3d_segment_test_pose(p, n, _d3(shift), _d3(a_), _d3(b_), C.byref(cfg),
                                      C.byref(pose) if pose is not None else None, C.byref(seg)))
    return seg


class WorldConfig(C.Structure):
    """h3d_world_config: diag_voxel (node.cpp:242), min_weight (node.cpp:205)."""
    _fields_ = [("diag_voxel", C.c_double), ("min_weight", C.c_double)]


class WorldSegment(C.Structure):
    """h3d_world_segment: `segment` of wrap.h:21-29 without the point copies."""
    _fields_ = [("a", C.c_double * 3), ("b", C.c_double * 3), ("t_min", C.c_double), ("t_max", C.c_double),
                ("radius", C.c_double), ("points_size", C.c_int32), ("pad_", C.c_int32), ("pca_coeff", C.c_double),
                ("pca_eigenvalues", C.c_double * 3)]

    @classmethod
    def make(cls, a_, b_, t_min, t_max, radius, points_size, pca_coeff, pca_eigenvalues):
        s = cls()
        for k in range(3):
            s.a[k], s.b[k], s.pca_eigenvalues[k] = float(a_[k]), float(b_[k]), float(pca_eigenvalues[k])
        s.t_min, s.t_max, s.radius, s.pca_coeff = float(t_min), float(t_max), float(radius), float(pca_coeff)
        s.points_size = int(points_size)
        return s

    def key(self):
        return (tuple(self.a), tuple(self.b), self.t_min, self.t_max, self.radius, self.points_size, self.pca_coeff,
                tuple(self.pca_eigenvalues))


def check_similarity(drone, world, cfg, target_init=None):
    """checkSimilarity (node.cpp:594-667) -> (similar, target segment)."""
    out, sim = WorldSegment(), C.c_int(0)
    check(lib().h3d_check_similarity(C.byref(drone), C.byref(world), C.byref(cfg),
                                     C.byref(target_init) if target_init is not None else None, C.byref(out),
                                     C.byref(sim)))
    return bool(sim.value), out


def check_connections(s1, s2, cfg):
    """checkConnections (node.cpp:556-590) -> (connected, (t1, t2, distance))."""
    sol, con = (C.c_double * 3)(), C.c_int(0)
    check(lib().h3d_check_connections(C.byref(s1), C.byref(s2), C.byref(cfg), sol, C.byref(con)))
    return bool(con.value), tuple(sol)


def set_verbosity(level):
    """node.cpp:23 verbosity (0 NONE, 1 INFO, 2 WARN) of the stderr logger; returns the previous level"""
    return lib().h3d_set_verbosity(C.c_int(level))


class World:
    """world_segments + intersection_matrix kept across frames (node.cpp:136-138, segFiltering node.cpp:479-542)."""

    def __init__(self, diag_voxel, min_weight):
        self.cfg = WorldConfig(diag_voxel, min_weight)
        self._w = C.c_void_p()
        check(lib().h3d_world_create(C.byref(self.cfg), C.byref(self._w)))

    def close(self):
        if self._w:
            lib().h3d_world_destroy(self._w)
            self._w = C.c_void_p()

    __del__ = close

    def update(self, drone):
        """drone: sequence of WorldSegment already in the world frame."""
        arr = (WorldSegment * max(len(drone), 1))(*drone)
        check(lib().h3d_world_update(self._w, arr, C.c_size_t(len(drone))))

    def update_frame(self, segs):
        """segs: the Segment records of one frame (Pipeline.segments / Replay.frame_segments); returns how many were taken."""
        arr = (Segment * max(len(segs), 1))(*segs)
        used = C.c_size_t(0)
        check(lib().h3d_world_update_frame(self._w, arr, C.c_size_t(len(segs)), C.byref(used)))
        return used.value

    def segments(self):
        n = C.c_size_t(0)
        check(lib().h3d_world_segments(self._w, None, C.c_size_t(0), C.byref(n)))
        out = (WorldSegment * max(n.value, 1))()
        check(lib().h3d_world_segments(self._w, out, C.c_size_t(n.value), C.byref(n)))
        return [out[k] for k in range(n.value)]

    def intersections(self):
        """n x n x 2 array of (t1, t2); (-1, -1) = none."""
        n = C.c_size_t(0)
        check(lib().h3d_world_intersections(self._w, None, C.c_size_t(0), C.byref(n)))
        m = n.value
        buf = np.full((m, m, 2), -1.0)
        if m:
            check(lib().h3d_world_intersections(self._w, buf.ctypes.data_as(C.POINTER(C.c_double)), C.c_size_t(m * m),
                                                C.byref(n)))
        return buf

    def note_frame(self, wall_time_us, callback_time_us, nblines, hough_rc=0):
        """node.cpp:331-346: log by verbosity, append the frame's callback_info"""
        check(lib().h3d_world_note_frame(self._w, C.c_int(wall_time_us), C.c_int(callback_time_us), C.c_int(nblines), C.c_int(hough_rc)))

    def save_processing_time(self, path):
        """saveProcessingTimeToFile, node.cpp:905-919"""
        check(lib().h3d_world_save_processing_time(self._w, path.encode()))

    def save_csv(self, segments_path=None, intersections_path=None):
        enc = lambda p: p.encode() if p is not None else None
        check(lib().h3d_world_save_csv(self._w, enc(segments_path), enc(intersections_path)))


def set_replay_threads(n):
    """host threads of a replay (0 = one per pipeline); returns the previous setting"""
    return lib().h3d_set_replay_threads(C.c_int(n))


def set_wait_mode(mode):
    """0: host threads spin on their stream (lowest latency); 1: they sleep on a blocking event;
    2: they poll an event and yield the core between polls."""
    return lib().h3d_set_wait_mode(C.c_int(int(mode)))


def set_sliced_vote(on):
    """Point-sliced full vote for few directions x many points (default on); returns the previous setting."""
    return int(lib().h3d_set_sliced_vote(C.c_int(1 if on else 0)))


def set_big_kernels(on):
    """Chunked multi-CTA K3 / removal kernels for clouds >= 32 768 points (default on); returns the previous setting."""
    return int(lib().h3d_set_big_kernels(C.c_int(1 if on else 0)))


def set_fused(on):
    """Use (True) or bypass (False) the fused persistent loop kernel; returns the previous setting."""
    return bool(lib().h3d_set_fused(C.c_int(1 if on else 0)))


def set_debug_poison(on):
    """Start every frame from per-direction records of 0xFFFFFFFF votes (a record the first vote skips then
    wins every argmax); returns the previous setting."""
    return int(lib().h3d_set_debug_poison(C.c_int(1 if on else 0)))


def set_fast_frames(on):
    """Frame fast path of the pipelines (one cluster launch for K0 + prep + Hough ctor, one wait per frame);
    returns the previous setting."""
    return bool(lib().h3d_set_fast_frames(C.c_int(1 if on else 0)))


def set_stream_mode(mode):
    """Streaming vote kernel: 0 = register tiles fetched with LDGs (default), 1 = TMA-staged tiles through a producer
    warp, 2 = round-1 kernel; returns the previous mode."""
    return int(lib().h3d_set_stream_mode(C.c_int(int(mode))))


def set_vote_track(mode):
    """First vote of the frame fast path: 1 / True = track maxima through the atomics' return values, 0 / False =
    scan the planes, 2 = pick on the device by point count (default); returns the previous mode."""
    return int(lib().h3d_set_vote_track(C.c_int(int(mode))))
