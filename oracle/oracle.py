"""ctypes front-end for the CPU oracle -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Two libraries, same call shapes:
  * ``Oracle()``     -> oracle/_build/libh3d_oracle.so : the plain-C restatement
    (oracle/h3d_oracle.c), always buildable with gcc.
  * ``Reference()``  -> oracle/_ref/libhough_ref.so : the UNMODIFIED reference
    library compiled in place from /root/reference (oracle/ref_shim.cpp is the
    only file of ours in it).  Present only where `make -C oracle ref` ran.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "_build", "libh3d_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libhough_ref.so")
REFERENCE_ROOT = "/root/reference"


class IterRec(C.Structure):
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
    ]


class SegmentConfig(C.Structure):
    """wrap.h:167-170 arguments of the post-tests (layout of h3d_segment_config)."""
    _fields_ = [("diag_voxel", C.c_double), ("min_pca_coeff", C.c_double), ("rad_2_leaf_ratio", C.c_double),
                ("n_radius", C.c_uint32), ("pad_", C.c_uint32), ("radius_sizes", C.c_double * 8)]

    @classmethod
    def make(cls, diag_voxel, min_pca_coeff, rad_2_leaf_ratio, radius_sizes):
        c = cls()
        c.diag_voxel, c.min_pca_coeff, c.rad_2_leaf_ratio = diag_voxel, min_pca_coeff, rad_2_leaf_ratio
        c.n_radius = len(radius_sizes)
        for k, r in enumerate(radius_sizes):
            c.radius_sizes[k] = r
        return c


class Segment(C.Structure):
    """One tested line (layout of h3d_segment): status 0 = pushed to computed_lines (wrap.h:324-335)."""
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


@dataclass
class Iteration:
    votes: int
    status: int
    argmax: int
    a_raw: np.ndarray
    b_raw: np.ndarray
    a1: np.ndarray
    b1: np.ndarray
    rc1: float
    a2: np.ndarray
    b2: np.ndarray
    rc2: float
    inliers1: np.ndarray
    inliers2: np.ndarray


@dataclass
class Trace:
    rc: int
    n_lines: int
    n_input: int
    shift: np.ndarray
    d: float
    dx: float
    max_x: float
    num_x: int
    num_b: int
    iters: list = field(default_factory=list)
    timing_us: np.ndarray | None = None


def build(ref: bool | None = None) -> None:
    """Compile the C restatement and, when /root/reference is present, oracle/_ref."""
    targets = ["all"]
    if ref is None:
        ref = os.path.isdir(os.path.join(REFERENCE_ROOT, "libraries", "hough-3d-lines"))
    if ref:
        targets.append("ref")
    subprocess.run(["make", "-s", "-C", HERE] + targets, check=True)


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _pts(xyz):
    a = np.ascontiguousarray(xyz, dtype=np.float64).reshape(-1, 3)
    return a, a.shape[0]


class _Base:
    prefix = ""

    def _f(self, name):
        return getattr(self.lib, self.prefix + name)

    # -- shared-shape calls ---------------------------------------------------
    def min_max(self, xyz):
        a, n = _pts(xyz)
        mn = np.zeros(3); mx = np.zeros(3)
        self._f("min_max")(_dp(a), C.c_size_t(n), _dp(mn), _dp(mx))
        return mn, mx

    def shift_to_origin(self, xyz, shift=None):
        a, n = _pts(xyz)
        a = a.copy()
        sh = np.zeros(3) if shift is None else np.array(shift, dtype=np.float64)
        self._f("shift_to_origin")(_dp(a), C.c_size_t(n), _dp(sh))
        return a, sh

    def mean_value(self, xyz):
        a, n = _pts(xyz)
        m = np.zeros(3)
        self._f("mean_value")(_dp(a), C.c_size_t(n), _dp(m))
        return m

    def points_close_to_line(self, xyz, a, b, dx):
        p, n = _pts(xyz)
        idx = np.zeros(max(n, 1), dtype=np.uint32)
        av = np.array(a, dtype=np.float64); bv = np.array(b, dtype=np.float64)
        fn = self._f("points_close_to_line"); fn.restype = C.c_size_t
        k = fn(_dp(p), C.c_size_t(n), _dp(av), _dp(bv), C.c_double(dx), idx.ctypes.data_as(C.POINTER(C.c_uint32)))
        return idx[:k].copy()

    def remove_points(self, x, y):
        xa, nx = _pts(x); ya, ny = _pts(y)
        xa = xa.copy()
        fn = self._f("remove_points"); fn.restype = C.c_size_t
        k = fn(_dp(xa), C.c_size_t(nx), _dp(ya), C.c_size_t(ny))
        return xa[:k].copy()

    def _unpack(self, res, iters, pool, timing=None):
        tr = Trace(rc=res.rc, n_lines=res.n_lines, n_input=res.n_input, shift=np.array(res.shift[:]),
                   d=res.d, dx=res.dx, max_x=res.max_x, num_x=res.num_x, num_b=res.num_b, timing_us=timing)
        for i in range(res.n_iters):
            r = iters[i]
            tr.iters.append(Iteration(
                votes=r.votes, status=r.status, argmax=r.argmax,
                a_raw=np.array(r.a_raw[:]), b_raw=np.array(r.b_raw[:]),
                a1=np.array(r.a1[:]), b1=np.array(r.b1[:]), rc1=r.rc1,
                a2=np.array(r.a2[:]), b2=np.array(r.b2[:]), rc2=r.rc2,
                inliers1=pool[r.off1:r.off1 + r.n1].copy(),
                inliers2=pool[r.off2:r.off2 + r.n2].copy() if r.status != 1 else np.zeros(0, np.uint32)))
        return tr


class Oracle(_Base):
    """Plain-C restatement (oracle/h3d_oracle.c)."""
    prefix = "h3do_"

    def __init__(self):
        if not os.path.exists(ORACLE_SO):
            build(ref=False)
        self.lib = C.CDLL(ORACLE_SO)
        L = self.lib
        L.h3do_sphere_count.restype = C.c_size_t
        L.h3do_sphere_fill.restype = C.c_size_t
        L.h3do_sphere_triangles.restype = C.c_size_t
        L.h3do_hough_create.restype = C.c_void_p
        for f in ("num_x", "num_b", "cells"):
            getattr(L, "h3do_hough_" + f).restype = C.c_size_t
            getattr(L, "h3do_hough_" + f).argtypes = [C.c_void_p]
        for f in ("dx", "max_x"):
            getattr(L, "h3do_hough_" + f).restype = C.c_double
            getattr(L, "h3do_hough_" + f).argtypes = [C.c_void_p]
        L.h3do_hough_votes.restype = C.POINTER(C.c_uint32)
        L.h3do_hough_votes.argtypes = [C.c_void_p]
        L.h3do_hough_destroy.argtypes = [C.c_void_p]
        L.h3do_hough_get_line.restype = C.c_uint32
        L.h3do_hough_get_line_mt.restype = C.c_uint32
        L.h3do_orthogonal_lsq.restype = C.c_double
        L.h3do_prefilter.restype = C.c_size_t
        self._dirs = {}

    def sphere(self, g):
        if g not in self._dirs:
            n = self.lib.h3do_sphere_count(C.c_int(g))
            a = np.zeros((n, 3))
            self.lib.h3do_sphere_fill(C.c_int(g), _dp(a), C.c_size_t(n))
            self._dirs[g] = a
        return self._dirs[g]

    def sphere_triangles(self, g):
        n = self.lib.h3do_sphere_triangles(C.c_int(g), None, C.c_size_t(0))
        t = np.zeros(n, dtype=np.uint32)
        self.lib.h3do_sphere_triangles(C.c_int(g), t.ctypes.data_as(C.POINTER(C.c_uint32)), C.c_size_t(n))
        return t

    def hough(self, g, minP, maxP, dx):
        return OracleHough(self, g, minP, maxP, dx)

    def orthogonal_lsq(self, xyz):
        p, n = _pts(xyz)
        a = np.zeros(3); b = np.zeros(3)
        rc = self.lib.h3do_orthogonal_lsq(_dp(p), C.c_size_t(n), _dp(a), _dp(b))
        return rc, a, b

    def segment_test(self, y_centred, shift, a, b, cfg):
        """wrap.h:261-337 on one line.  Returns (rc, Segment); rc 1 = find_t had no solution."""
        p, n = _pts(y_centred)
        seg = Segment()
        sh = np.ascontiguousarray(shift, dtype=np.float64)
        aa = np.ascontiguousarray(a, dtype=np.float64); bb = np.ascontiguousarray(b, dtype=np.float64)
        rc = self.lib.h3do_segment_test(_dp(p), C.c_size_t(n), _dp(sh), _dp(aa), _dp(bb), C.byref(cfg), C.byref(seg))
        return rc, seg

    def segments(self, xyz, trace, cfg, pose=None):
        """Post-tests of every accepted line of a trace of hough3dlines(xyz, ...): (rc, [Segment]).
        rc 1 = the reference would return 1 at the last segment of the list (status 3).
        pose: also drone2WorldSeg + heighSegmentCutoff (node.cpp:429-470) on the kept segments."""
        p, _ = _pts(xyz)
        p = p.reshape(-1, 3)
        centred = p - trace.shift            # PointCloud::shiftToOrigin, lib/pointcloud.cpp:18-26
        out = []
        for k, it in enumerate(trace.iters):
            if it.status != 0:
                continue
            rc, seg = self.segment_test(centred[it.inliers2], trace.shift, it.a2, it.b2, cfg)
            seg.iter = k
            if pose is not None:
                self.lib.h3do_segment_to_world(C.byref(pose), C.byref(seg))
            out.append(seg)
            if rc:
                return 1, out
        return 0, out

    def eig3f_fixed(self, m):
        m = np.ascontiguousarray(m, dtype=np.float32).reshape(9)
        ev = np.zeros(3, np.float32)
        fp = C.POINTER(C.c_float)
        rc = self.lib.h3do_eig3f_fixed(m.ctypes.data_as(fp), ev.ctypes.data_as(fp))
        return rc, ev

    def eig3f(self, m):
        m = np.ascontiguousarray(m, dtype=np.float32).reshape(9)
        ev = np.zeros(3, np.float32); evec = np.zeros(9, np.float32)
        fp = C.POINTER(C.c_float)
        rc = self.lib.h3do_eig3f(m.ctypes.data_as(fp), ev.ctypes.data_as(fp), evec.ctypes.data_as(fp))
        return rc, ev, evec.reshape(3, 3).T.copy()   # evec[:, k] = k-th eigenvector

    def prefilter(self, xyz_f32, window, leaf, stride=None):
        a = np.ascontiguousarray(xyz_f32, dtype=np.float32)
        if a.ndim == 1:
            stride = stride or 3
            n = a.size // stride
        else:
            n, stride = a.shape
        out = np.zeros((max(n, 1), 3), dtype=np.float32)
        win = np.array(window, dtype=np.float32)
        fp = C.POINTER(C.c_float)
        k = self.lib.h3do_prefilter(a.ctypes.data_as(fp), C.c_size_t(n), C.c_size_t(stride), win.ctypes.data_as(fp),
                                    C.c_float(leaf), out.ctypes.data_as(fp), C.c_size_t(max(n, 1)))
        if k == C.c_size_t(-1).value:
            raise OverflowError("voxel index overflow")
        return out[:k].copy()

    def hough3dlines(self, xyz, granularity, dx, minvotes, nlines, mode=0, max_iters=None):
        p, n = _pts(xyz)
        cap_it = max_iters or (nlines + 2 if nlines else n + 2)
        iters = (IterRec * cap_it)()
        pool = np.zeros(2 * max(n, 1) * min(cap_it, 64) + 16, dtype=np.uint32)
        res = Result()
        self.lib.h3do_hough3dlines(_dp(p), C.c_size_t(n), C.c_int(granularity), C.c_double(dx), C.c_int(minvotes),
                                   C.c_int(nlines), C.c_int(mode), iters, C.c_size_t(cap_it),
                                   pool.ctypes.data_as(C.POINTER(C.c_uint32)), C.c_size_t(pool.size), C.byref(res))
        return self._unpack(res, iters, pool)


class OracleHough:
    def __init__(self, orc, g, minP, maxP, dx):
        self.o = orc
        self.dirs = orc.sphere(g)
        mn = np.array(minP, dtype=np.float64); mx = np.array(maxP, dtype=np.float64)
        self.h = orc.lib.h3do_hough_create(_dp(self.dirs), C.c_size_t(len(self.dirs)), _dp(mn), _dp(mx), C.c_double(dx))
        if not self.h:
            raise MemoryError("h3do_hough_create failed")
        L = orc.lib
        self.num_x = L.h3do_hough_num_x(self.h); self.num_b = L.h3do_hough_num_b(self.h)
        self.dx = L.h3do_hough_dx(self.h); self.max_x = L.h3do_hough_max_x(self.h)
        self.cells = L.h3do_hough_cells(self.h)

    def add(self, xyz, threads=1):
        p, n = _pts(xyz)
        if threads > 1:
            self.o.lib.h3do_hough_add_mt(C.c_void_p(self.h), _dp(p), C.c_size_t(n), C.c_int(threads))
        else:
            self.o.lib.h3do_hough_add(C.c_void_p(self.h), _dp(p), C.c_size_t(n))

    def subtract(self, xyz, threads=1):
        p, n = _pts(xyz)
        if threads > 1:
            self.o.lib.h3do_hough_subtract_mt(C.c_void_p(self.h), _dp(p), C.c_size_t(n), C.c_int(threads))
        else:
            self.o.lib.h3do_hough_subtract(C.c_void_p(self.h), _dp(p), C.c_size_t(n))

    def get_line(self, threads=1):
        a = np.zeros(3); b = np.zeros(3); idx = C.c_uint64(0)
        if threads > 1:
            v = self.o.lib.h3do_hough_get_line_mt(C.c_void_p(self.h), _dp(a), _dp(b), C.byref(idx), C.c_int(threads))
        else:
            v = self.o.lib.h3do_hough_get_line(C.c_void_p(self.h), _dp(a), _dp(b), C.byref(idx))
        return v, idx.value, a, b

    def votes(self):
        ptr = self.o.lib.h3do_hough_votes(C.c_void_p(self.h))
        return np.ctypeslib.as_array(ptr, shape=(self.cells,)).copy()

    def close(self):
        if self.h:
            self.o.lib.h3do_hough_destroy(C.c_void_p(self.h)); self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Reference(_Base):
    """The real reference library (compiled in place) behind oracle/ref_shim.cpp."""
    prefix = "ref_"

    @staticmethod
    def available():
        return os.path.exists(REF_SO)

    def __init__(self):
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO + " (run `make -C oracle ref` where /root/reference exists)")
        self.lib = C.CDLL(REF_SO)
        L = self.lib
        L.ref_sphere.restype = C.c_size_t
        L.ref_sphere_triangles.restype = C.c_size_t
        L.ref_hough_create.restype = C.c_void_p
        for f in ("num_x", "num_b", "cells"):
            getattr(L, "ref_hough_" + f).restype = C.c_size_t
            getattr(L, "ref_hough_" + f).argtypes = [C.c_void_p]
        for f in ("dx", "max_x"):
            getattr(L, "ref_hough_" + f).restype = C.c_double
            getattr(L, "ref_hough_" + f).argtypes = [C.c_void_p]
        L.ref_hough_votes.restype = C.POINTER(C.c_uint32)
        L.ref_hough_votes.argtypes = [C.c_void_p]
        L.ref_hough_destroy.argtypes = [C.c_void_p]
        L.ref_hough_get_line.restype = C.c_uint32
        self._dirs = {}

    def sphere(self, g):
        if g not in self._dirs:
            n = self.lib.ref_sphere(C.c_int(g), None, C.c_size_t(0))
            a = np.zeros((n, 3))
            self.lib.ref_sphere(C.c_int(g), _dp(a), C.c_size_t(n))
            self._dirs[g] = a
        return self._dirs[g]

    def sphere_triangles(self, g):
        n = self.lib.ref_sphere_triangles(C.c_int(g), None, C.c_size_t(0))
        t = np.zeros(n, dtype=np.uint32)
        self.lib.ref_sphere_triangles(C.c_int(g), t.ctypes.data_as(C.POINTER(C.c_uint32)), C.c_size_t(n))
        return t

    def hough(self, g, minP, maxP, dx):
        return ReferenceHough(self, g, minP, maxP, dx)

    def read_from_file(self, path, delim=","):
        cap = 1 << 20
        a = np.zeros((cap, 3)); n = C.c_size_t(0)
        rc = self.lib.ref_read_from_file(path.encode(), C.c_char(delim.encode()), _dp(a), C.c_size_t(cap), C.byref(n))
        return rc, a[:n.value].copy()

    def set_index_recovery(self, on):
        """on = False: the loop driver runs the reference calls only (no flat-index recovery scan after getLine,
        no inlier index lists) -- what bench.py times; records then carry argmax = 0 and empty index lists."""
        return bool(self.lib.ref_set_index_recovery(C.c_int(1 if on else 0)))

    def hough3dlines(self, xyz, granularity, dx, minvotes, nlines, mode=0, max_iters=None):
        p, n = _pts(xyz)
        cap_it = max_iters or (nlines + 2 if nlines else n + 2)
        iters = (IterRec * cap_it)()
        pool = np.zeros(2 * max(n, 1) * min(cap_it, 64) + 16, dtype=np.uint32)
        res = Result()
        timing = np.zeros(6)
        self.lib.ref_hough3dlines(_dp(p), C.c_size_t(n), C.c_int(granularity), C.c_double(dx), C.c_int(minvotes),
                                  C.c_int(nlines), C.c_int(mode), iters, C.c_size_t(cap_it),
                                  pool.ctypes.data_as(C.POINTER(C.c_uint32)), C.c_size_t(pool.size), C.byref(res),
                                  _dp(timing))
        return self._unpack(res, iters, pool, timing)


class ReferenceHough:
    def __init__(self, ref, g, minP, maxP, dx):
        self.r = ref
        ref.lib.ref_init_space(C.c_int(g))
        mn = np.array(minP, dtype=np.float64); mx = np.array(maxP, dtype=np.float64)
        self.h = ref.lib.ref_hough_create(_dp(mn), _dp(mx), C.c_double(dx))
        if not self.h:
            raise MemoryError("ref_hough_create failed")
        L = ref.lib
        self.num_x = L.ref_hough_num_x(self.h); self.num_b = L.ref_hough_num_b(self.h)
        self.dx = L.ref_hough_dx(self.h); self.max_x = L.ref_hough_max_x(self.h)
        self.cells = L.ref_hough_cells(self.h)

    def add(self, xyz, threads=1):
        p, n = _pts(xyz)
        self.r.lib.ref_hough_add(C.c_void_p(self.h), _dp(p), C.c_size_t(n))

    def subtract(self, xyz, threads=1):
        p, n = _pts(xyz)
        self.r.lib.ref_hough_subtract(C.c_void_p(self.h), _dp(p), C.c_size_t(n))

    def get_line(self, threads=1):
        a = np.zeros(3); b = np.zeros(3); idx = C.c_uint64(0)
        v = self.r.lib.ref_hough_get_line(C.c_void_p(self.h), _dp(a), _dp(b), C.byref(idx))
        return v, idx.value, a, b

    def votes(self):
        ptr = self.r.lib.ref_hough_votes(C.c_void_p(self.h))
        return np.ctypeslib.as_array(ptr, shape=(self.cells,)).copy()

    def close(self):
        if self.h:
            self.r.lib.ref_hough_destroy(C.c_void_p(self.h)); self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def read_dat(path, delim=","):
    """Pure-python reader of the reference .dat format (lib/pointcloud.cpp:63-91): '#' comments, 3 fields."""
    pts = []
    with open(path) as f:
        for line in f:
            if line.startswith("#"):
                continue
            parts = line.strip().split(delim)
            if len(parts) < 3:
                raise ValueError("bad line: " + line)
            pts.append([float(parts[0]), float(parts[1]), float(parts[2])])
    return np.array(pts, dtype=np.float64).reshape(-1, 3)
