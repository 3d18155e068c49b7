"""Deterministic synthetic inputs of the BASELINE.json configurations.

No reference data exists for them (tower.stl is a missing blob, the ToF sensor has no
resolution in starling.proto:598-606), so the shapes are synthesised here: steel
beams are cylinders of radius 0.05 m, the ToF frame is a CPU ray cast of a
224x172 pinhole (hFOV 2.04 rad, range 0.1-4 m, sigma 0.01 m), x forward / y left /
z up as in the node's PassThrough window (node.cpp:394-407).
"""
from __future__ import annotations

import numpy as np

TOF_W, TOF_H = 224, 172
TOF_HFOV = 2.04
TOF_RANGE = (0.1, 4.0)
TOF_SIGMA = 0.01
BEAM_RADIUS = 0.05
# node.cpp:394-407 window, config.yaml: rad_2_leaf_ratio 1.5, radius 0.05
WINDOW = (0.0, 1.5, -1.5, 1.5, -1.5, 1.5)
LEAF = np.float32(BEAM_RADIUS / 1.5)
OPT_DX = float(np.sqrt(3.0) * float(LEAF))   # node.cpp:241-243


def tower_beams(height=3.0, side=1.0):
    """12 beam axes (p0, p1): 4 verticals, 4 top horizontals, 4 face diagonals."""
    h, s = height, side / 2.0
    corners = [(-s, -s), (s, -s), (s, s), (-s, s)]
    beams = []
    for (x, y) in corners:
        beams.append(((x, y, 0.0), (x, y, h)))
    for k in range(4):
        (x0, y0), (x1, y1) = corners[k], corners[(k + 1) % 4]
        beams.append(((x0, y0, h * 0.66), (x1, y1, h * 0.66)))
        beams.append(((x0, y0, 0.1 * h), (x1, y1, 0.6 * h)))
    return np.array(beams, dtype=np.float64)


def beam_cloud(beams, n_points, radius=BEAM_RADIUS, sigma=0.01, seed=1, clutter=0.0, box=None):
    """Points uniform in (t, theta) on the cylinder surfaces + N(0, sigma^2) per axis."""
    rng = np.random.default_rng(seed)
    beams = np.asarray(beams, dtype=np.float64)
    nb = len(beams)
    n_clutter = int(round(n_points * clutter))
    n_surf = n_points - n_clutter
    which = rng.integers(0, nb, size=n_surf)
    t = rng.random(n_surf)
    th = rng.random(n_surf) * 2 * np.pi
    p0, p1 = beams[which, 0], beams[which, 1]
    ax = p1 - p0
    ax /= np.linalg.norm(ax, axis=1, keepdims=True)
    helper = np.where(np.abs(ax[:, 2:3]) < 0.9, np.array([[0.0, 0.0, 1.0]]), np.array([[1.0, 0.0, 0.0]]))
    u = np.cross(ax, helper)
    u /= np.linalg.norm(u, axis=1, keepdims=True)
    v = np.cross(ax, u)
    rad = np.broadcast_to(np.asarray(radius, dtype=np.float64), (nb,))[which][:, None]
    pts = p0 + (p1 - p0) * t[:, None] + rad * (np.cos(th)[:, None] * u + np.sin(th)[:, None] * v)
    pts += rng.standard_normal(pts.shape) * sigma
    if n_clutter:
        lo, hi = (pts.min(0), pts.max(0)) if box is None else box
        pts = np.vstack([pts, lo + rng.random((n_clutter, 3)) * (np.asarray(hi) - np.asarray(lo))])
    return np.ascontiguousarray(pts)


def steel_tower(n_points=10000, seed=1):
    """Config 1: 10k-point steel tower, 12 beams r = 0.05, noise 1 cm."""
    return beam_cloud(tower_beams(), n_points, seed=seed)


def infrastructure(n_points=1_000_000, n_beams=200, seed=4):
    """Config 4: dense cloud of 200 random beams in a 40x40x60 m volume + 5 % clutter."""
    rng = np.random.default_rng(seed)
    lo, hi = np.array([-20.0, -20.0, 0.0]), np.array([20.0, 20.0, 60.0])
    p0 = lo + rng.random((n_beams, 3)) * (hi - lo)
    d = rng.standard_normal((n_beams, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    length = 2.0 + rng.random(n_beams) * 18.0
    p1 = np.clip(p0 + d * length[:, None], lo, hi)
    radius = 0.05 + rng.random(n_beams) * 0.2
    beams = np.stack([p0, p1], axis=1)
    return beam_cloud(beams, n_points, radius=radius, sigma=0.01, seed=seed + 1, clutter=0.05, box=(lo, hi))


def _ray_cylinders(o, d, beams, radius):
    """Nearest positive hit distance of rays (o + s d) with finite cylinders; inf on miss."""
    best = np.full(d.shape[0], np.inf)
    for (p0, p1) in beams:
        ax = p1 - p0
        length = np.linalg.norm(ax)
        ax = ax / length
        oc = o - p0
        d_par = d @ ax
        oc_par = oc @ ax
        dp = d - d_par[:, None] * ax
        op = oc - oc_par * ax
        a = np.einsum("ij,ij->i", dp, dp)
        b = 2.0 * (dp @ op)
        c = op @ op - radius * radius
        disc = b * b - 4 * a * c
        ok = (disc > 0) & (a > 1e-12)
        s = np.where(ok, (-b - np.sqrt(np.where(ok, disc, 0.0))) / np.where(ok, 2 * a, 1.0), np.inf)
        tt = oc_par + s * d_par
        hit = ok & (s > 0) & (tt >= 0) & (tt <= length)
        best = np.where(hit & (s < best), s, best)
    return best


def tof_frame(pose=(-1.0, -0.25, 1.0, 0.12), seed=2, beams=None, floor=True, stride=4):
    """Config 2: one 224x172 ToF frame in the SENSOR frame, float32, `stride` floats per
    point (x, y, z, pad) like pcl::PointXYZ.  pose = (x, y, z, yaw) of the sensor in the
    world; the tower stands on the origin.  All 38 528 pixels are returned; rays without a
    return in range are NaN points (is_dense = false), which the PassThrough drops."""
    rng = np.random.default_rng(seed)
    beams = tower_beams() if beams is None else beams
    px, py, pz, yaw = pose
    f = (TOF_W / 2.0) / np.tan(TOF_HFOV / 2.0)
    u = (np.arange(TOF_W) + 0.5 - TOF_W / 2.0) / f
    v = (np.arange(TOF_H) + 0.5 - TOF_H / 2.0) / f
    uu, vv = np.meshgrid(u, v)
    d_s = np.stack([np.ones_like(uu), -uu, -vv], axis=-1).reshape(-1, 3)   # x fwd, y left, z up
    d_s /= np.linalg.norm(d_s, axis=1, keepdims=True)
    cy, sy = np.cos(yaw), np.sin(yaw)
    rot = np.array([[cy, -sy, 0.0], [sy, cy, 0.0], [0.0, 0.0, 1.0]])
    d_w = d_s @ rot.T
    o = np.array([px, py, pz], dtype=np.float64)
    rng_hit = _ray_cylinders(o, d_w, beams, BEAM_RADIUS)
    if floor:
        with np.errstate(divide="ignore", invalid="ignore"):
            s_floor = np.where(d_w[:, 2] < -1e-9, -pz / d_w[:, 2], np.inf)
        rng_hit = np.minimum(rng_hit, s_floor)
    ok = (rng_hit >= TOF_RANGE[0]) & (rng_hit <= TOF_RANGE[1])
    r = np.where(ok, rng_hit, np.nan) + rng.standard_normal(ok.shape[0]) * TOF_SIGMA
    pts = (d_s * r[:, None]).astype(np.float32)       # misses stay in the frame as NaN points
    out = np.zeros((pts.shape[0], stride), dtype=np.float32)
    out[:, :3] = pts
    if stride > 3:
        out[:, 3] = 1.0
    return out


def tower_trajectory(n_frames=1000):
    """Config 3: poses circling the tower at 1 m (stand-in for config_auto_pilot/wp_tower.csv,
    whose z range needs the 15 m tower that tower.stl would have provided)."""
    k = np.arange(n_frames)
    ang = 2 * np.pi * k / max(n_frames, 1) * 3.0
    rad = 1.15 + 0.1 * np.sin(7 * ang)
    z = 0.6 + 1.6 * (k / max(n_frames - 1, 1))
    x, y = rad * np.cos(ang), rad * np.sin(ang)
    yaw = np.arctan2(-y, -x) + 0.1 * np.sin(5 * ang)
    return np.stack([x, y, z, yaw], axis=1)
