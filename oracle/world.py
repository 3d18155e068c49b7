"""oracle/world.py -- TEST INFRASTRUCTURE ONLY (never imported by the product path).

CPU restatement, in scalar Python floats (IEEE double, no FMA), of the reference's
world model kept across frames -- SURVEY 8(f) rank 4:

  segFiltering      src/pointcloud_segmentation_node.cpp:479-542
  checkConnections  src/pointcloud_segmentation_node.cpp:556-590
  checkSimilarity   src/pointcloud_segmentation_node.cpp:594-667
  find_proj         include/pointcloud_segmentation/hough_3d_lines.h:78-85
  save*ToFile       src/pointcloud_segmentation_node.cpp:850-897

Third party: Eigen3 (version unpinned, not vendored, absent here): fixed-size
Vector3d expressions (3-term reductions x0 + (x1 + x2)) and
ColPivHouseholderQR<Matrix3d>::solve, restated from Eigen 3.3's published
algorithm with its small inner sums in index order.  The reference holds no test
or golden vector for these functions and cannot be built here (ROS/PCL/Eigen):
PARITY UNPINNED.  `qr_solve3` is checked against numpy.linalg.solve in
tests/test_world_cpu.py.

Where the reference reads indeterminate values (`segment target_seg;` at
node.cpp:494 is uninitialised when node.cpp:651-654 add to its fields) the
restatement starts `target_seg` from the world segment; int 0 / 0 gives weight 0.
"""
import math
import sys

EPS = sys.float_info.epsilon
DBL_MIN = sys.float_info.min


class Seg:
    """`segment` (hough_3d_lines.h:21-29) without the point copies."""
    __slots__ = ("a", "b", "t_min", "t_max", "radius", "points_size", "pca_coeff", "pca_eigenvalues")

    def __init__(self, a, b, t_min, t_max, radius, points_size, pca_coeff, pca_eigenvalues):
        self.a = [float(v) for v in a]
        self.b = [float(v) for v in b]
        self.t_min, self.t_max, self.radius = float(t_min), float(t_max), float(radius)
        self.points_size = int(points_size)
        self.pca_coeff = float(pca_coeff)
        self.pca_eigenvalues = [float(v) for v in pca_eigenvalues]

    def copy(self):
        return Seg(self.a, self.b, self.t_min, self.t_max, self.radius, self.points_size, self.pca_coeff,
                   self.pca_eigenvalues)

    def key(self):
        return (tuple(self.a), tuple(self.b), self.t_min, self.t_max, self.radius, self.points_size, self.pca_coeff,
                tuple(self.pca_eigenvalues))


def _dot(u, v):
    return u[0] * v[0] + (u[1] * v[1] + u[2] * v[2])


def _sub(u, v):
    return [u[0] - v[0], u[1] - v[1], u[2] - v[2]]


def _add(u, v):
    return [u[0] + v[0], u[1] + v[1], u[2] + v[2]]


def _scale(s, v):
    return [s * v[0], s * v[1], s * v[2]]


def _norm(u):
    return math.sqrt(_dot(u, u))


def _div(x, y):
    """IEEE division (Python raises on a zero divisor)."""
    if y == 0.0:
        if x == 0.0 or x != x:
            return math.nan
        return math.copysign(math.inf, x) * math.copysign(1.0, y)
    return x / y


def _point(s, t):
    return _add(_scale(t, s.b), s.a)                       # t * b + a


def find_proj(a, b, p):
    """hough_3d_lines.h:78-85."""
    d = _sub(_add(a, b), a)
    num, den = _dot(_sub(p, a), d), _dot(d, d)
    return [a[k] + _div(d[k] * num, den) for k in range(3)]


def _cmin(x, y):                                           # std::min
    return y if y < x else x


def _cmax(x, y):                                           # std::max
    return y if x < y else x


def _wrap32(v):
    v &= 0xFFFFFFFF
    return v - (1 << 32) if v & 0x80000000 else v


def _cdiv(a, b):                                           # C int division truncates toward zero
    q = abs(a) // abs(b)
    return q if (a < 0) == (b < 0) else -q


def check_similarity(drone, world, diag_voxel, min_weight, target_init=None):
    """node.cpp:594-667 -> (similar, target_seg)."""
    target = (target_init if target_init is not None else world).copy()
    world_p1, world_p2 = _point(world, world.t_min), _point(world, world.t_max)
    test_p1, test_p2 = _point(drone, drone.t_min), _point(drone, drone.t_max)
    proj_p1 = find_proj(world.a, world.b, test_p1)
    proj_p2 = find_proj(world.a, world.b, test_p2)
    epsilon = (drone.radius + world.radius) + 2 * (2 * diag_voxel)
    if (_norm(_sub(proj_p1, test_p1)) < epsilon and _norm(_sub(proj_p2, test_p2)) < epsilon
            and drone.radius == world.radius):
        total = _wrap32(world.points_size + drone.points_size)
        weight = 0.0
        if total != 0 and not (drone.points_size == -(1 << 31) and total == -1):
            weight = float(_cdiv(drone.points_size, total))
        weight = _cmax(min_weight, weight)
        coeff = _div(drone.pca_coeff * weight, world.pca_coeff * (1 - weight) + drone.pca_coeff * weight)
        new_a = _add(proj_p1, _scale(coeff, _sub(test_p1, proj_p1)))
        new_b = _add(_sub(proj_p2, proj_p1), _scale(coeff, _sub(_sub(test_p2, proj_p2), _sub(test_p1, proj_p1))))
        ts = [_div(find_proj(new_a, new_b, p)[0] - new_a[0], new_b[0]) for p in (test_p1, test_p2, world_p1, world_p2)]
        if not (_cmin(ts[0], ts[1]) > _cmax(ts[2], ts[3]) or _cmax(ts[0], ts[1]) < _cmin(ts[2], ts[3])):
            hi = lo = ts[0]
            for t in ts[1:]:
                if hi < t:
                    hi = t
                if t < lo:
                    lo = t
            target.a, target.b, target.t_max, target.t_min = new_a, new_b, hi, lo
            target.radius = drone.radius
            target.points_size = _wrap32(target.points_size + drone.points_size)
            target.pca_coeff = target.pca_coeff * (1 - weight) + drone.pca_coeff * weight
            target.pca_eigenvalues = [target.pca_eigenvalues[k] * (1 - weight) + drone.pca_eigenvalues[k] * weight
                                      for k in range(3)]
            return True, target
    return False, drone.copy()


def qr_solve3(cols, rhs):
    """Eigen 3.3 ColPivHouseholderQR<Matrix3d>(M).solve(rhs); cols = the three columns of M."""
    n = 3
    m = [[float(v) for v in c] for c in cols]              # m[column][row]
    direct = [0.0] * n
    for k in range(n):
        s = 0.0
        for r in range(n):
            s += m[k][r] * m[k][r]
        direct[k] = math.sqrt(s)
    upd = list(direct)
    th = max(upd) * EPS
    helper = (th * th) / float(n)
    down = math.sqrt(EPS)
    rank = n
    taus, trans = [0.0] * n, [0] * n
    for k in range(n):
        big = k
        for j in range(k + 1, n):
            if upd[j] > upd[big]:
                big = j
        if rank == n and upd[big] * upd[big] < helper * float(n - k):
            rank = k
        trans[k] = big
        if big != k:
            m[k], m[big] = m[big], m[k]
            upd[k], upd[big] = upd[big], upd[k]
            direct[k], direct[big] = direct[big], direct[k]
        tail = 0.0
        for r in range(k + 1, n):
            tail += m[k][r] * m[k][r]
        c0 = m[k][k]
        if tail <= DBL_MIN:
            tau, beta = 0.0, c0
            for r in range(k + 1, n):
                m[k][r] = 0.0
        else:
            beta = math.sqrt(c0 * c0 + tail)
            if c0 >= 0.0:
                beta = -beta
            for r in range(k + 1, n):
                m[k][r] = m[k][r] / (c0 - beta)
            tau = (beta - c0) / beta
        taus[k] = tau
        m[k][k] = beta
        for j in range(k + 1, n):
            if n - k == 1:
                m[j][k] *= (1.0 - tau)
            elif tau != 0.0:
                tmp = 0.0
                for r in range(k + 1, n):
                    tmp += m[k][r] * m[j][r]
                tmp += m[j][k]
                m[j][k] -= tau * tmp
                for r in range(k + 1, n):
                    m[j][r] -= (tau * m[k][r]) * tmp
        for j in range(k + 1, n):
            if upd[j] != 0.0:
                temp = abs(m[j][k]) / upd[j]
                temp = (1.0 + temp) * (1.0 - temp)
                temp = 0.0 if temp < 0.0 else temp
                q = upd[j] / direct[j]
                if temp * (q * q) <= down:
                    s = 0.0
                    for r in range(k + 1, n):
                        s += m[j][r] * m[j][r]
                    direct[j] = upd[j] = math.sqrt(s)
                else:
                    upd[j] *= math.sqrt(temp)
    perm = [0, 1, 2]
    for k in range(n):
        perm[k], perm[trans[k]] = perm[trans[k]], perm[k]
    x = [0.0, 0.0, 0.0]
    if rank == 0:
        return x
    c = [float(v) for v in rhs]
    for k in range(rank):
        tau = taus[k]
        if n - k == 1:
            c[k] *= (1.0 - tau)
        elif tau != 0.0:
            tmp = 0.0
            for r in range(k + 1, n):
                tmp += m[k][r] * c[r]
            tmp += c[k]
            c[k] -= tau * tmp
            for r in range(k + 1, n):
                c[r] -= (tau * m[k][r]) * tmp
    for i in range(rank - 1, -1, -1):
        s = c[i]
        for j in range(i + 1, rank):
            s -= m[j][i] * c[j]
        c[i] = _div(s, m[i][i])
    for i in range(rank):
        x[perm[i]] = c[i]
    return x


def check_connections(s1, s2, diag_voxel):
    """node.cpp:556-590 (s1 = drone_seg, s2 = world_seg) -> (connected, sol)."""
    p1, q1 = _point(s1, s1.t_min), _point(s2, s2.t_min)
    b1, b2 = s1.b, s2.b
    cr = [b2[1] * b1[2] - b2[2] * b1[1], b2[2] * b1[0] - b2[0] * b1[2], b2[0] * b1[1] - b2[1] * b1[0]]
    if _norm(cr) < 1e-2:
        return False, [0.0, 0.0, 0.0]
    z = _dot(cr, cr)
    if z > 0.0:
        s = math.sqrt(z)
        cr = [cr[0] / s, cr[1] / s, cr[2] / s]
    sol = qr_solve3([b1, [-b2[0], -b2[1], -b2[2]], cr], _sub(q1, p1))
    eps = (2 * diag_voxel + s1.radius) + s2.radius
    ok = (s1.t_min <= sol[0] + s1.t_min <= s1.t_max) and (s2.t_min <= sol[1] + s2.t_min <= s2.t_max) \
        and abs(sol[2]) < eps
    return ok, sol


class World:
    """world_segments + intersection_matrix of PtCdProcessing (node.cpp:136-138)."""

    def __init__(self, diag_voxel, min_weight):
        self.diag_voxel, self.min_weight = float(diag_voxel), float(min_weight)
        self.segments = []
        self.inter = []                                    # inter[i][j] = (t1, t2)

    def update(self, drone):
        """segFiltering, node.cpp:479-542."""
        new_idx, mod_idx = [], []
        nw = [s.copy() for s in self.segments]
        if not self.segments:
            nw = [s.copy() for s in drone]
        else:
            for i, d in enumerate(drone):
                found = False
                for j, ws in enumerate(self.segments):
                    similar, target = check_similarity(d, ws, self.diag_voxel, self.min_weight)
                    if similar:
                        nw[j] = target
                        mod_idx.append(j)
                        found = True
                        break
                if not found:
                    nw.append(d.copy())
                    new_idx.append(len(nw) + i)            # node.cpp:508 as written
        n = len(nw)
        inter = [list(row) + [(-1.0, -1.0)] * (n - len(row)) for row in self.inter]
        inter += [[(-1.0, -1.0)] * n for _ in range(n - len(inter))]
        targets = set(mod_idx) | set(new_idx)
        for i in range(n):
            for j in range(i):
                if i in targets or j in targets:
                    ok, sol = check_connections(nw[i], nw[j], self.diag_voxel)
                    if ok:
                        inter[i][j] = (nw[i].t_min + sol[0], nw[j].t_min + sol[1])
        self.segments, self.inter = nw, inter

    def csv(self):
        """saveSegmentsToFile / saveIntersectionsToFile (node.cpp:850-897): default ostream formatting = %g."""
        seg = ["segment,a_x,a_y,a_z,b_x,b_y,b_z,t_min,t_max"]
        for i, s in enumerate(self.segments):
            seg.append(",".join([str(i)] + ["%g" % v for v in (*s.a, *s.b, s.t_min, s.t_max)]))
        it = ["seg1,t1,seg2,t2"]
        for i, row in enumerate(self.inter):
            for j in range(i):
                t1, t2 = row[j]
                if t1 != -1.0 and t2 != -1.0:
                    it.append("%d,%g,%d,%g" % (i, t1, j, t2))
        return "\n".join(seg) + "\n", "\n".join(it) + "\n"
