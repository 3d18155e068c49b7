/*
 * h3d_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see h3d_oracle.h).
 *
 * Plain-C restatement of the reference algorithm, one function per reference
 * function, each citing the file:line it follows.  Build with
 *   gcc -O2 -ffp-contract=off   (no -march=native, no -ffast-math)
 * to match the reference build (CMakeLists.txt:5, lib/Makefile:19: no FMA).
 *
 * Parity: everything except h3do_orthogonal_lsq's float eigen pipeline and
 * h3do_prefilter is checked bit-for-bit against the reference library compiled
 * in place (oracle/_ref).  Those two restate un-vendored third parties (Eigen3,
 * PCL 1.10): PARITY UNPINNED.
 */
#include "h3d_oracle.h"

#include <float.h>
#include <limits.h>
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ======================================================================== */
/* Direction sphere -- lib/sphere.cpp                                        */
/* ======================================================================== */

typedef struct { double x, y, z; } v3;

typedef struct {
    v3 *v; size_t nv, capv;
    unsigned *t; size_t nt, capt;   /* 3 indices per triangle, FIFO order */
} sphere_t;

static void sph_push_v(sphere_t *s, v3 p) {
    if (s->nv == s->capv) { s->capv = s->capv ? 2 * s->capv : 64; s->v = (v3 *)realloc(s->v, s->capv * sizeof(v3)); }
    s->v[s->nv++] = p;
}

/* lib/sphere.cpp:27-145 getIcosahedron: constants are computed in FLOAT */
static void sph_icosahedron(sphere_t *s) {
    float tau = 1.61803399f;
    float norm = sqrtf(1 + tau * tau);
    float v = 1 / norm;
    tau = tau / norm;
    const double V = v, T = tau;
    const v3 verts[12] = {
        {-V, T, 0}, {V, T, 0}, {0, V, -T}, {0, V, T}, {-T, 0, -V}, {T, 0, -V},
        {-T, 0, V}, {T, 0, V}, {0, -V, -T}, {0, -V, T}, {-V, -T, 0}, {V, -T, 0}};
    static const unsigned tris[60] = {
        0, 1, 2,  0, 1, 3,  0, 2, 4,  0, 4, 6,  0, 3, 6,  1, 2, 5,  1, 3, 7,
        1, 5, 7,  2, 4, 8,  2, 5, 8,  3, 6, 9,  3, 7, 9,  4, 8, 10, 8, 10, 11,
        5, 8, 11, 5, 7, 11, 7, 9, 11, 9, 10, 11, 6, 9, 10, 4, 6, 10};
    s->nv = 0;
    for (int i = 0; i < 12; i++) sph_push_v(s, verts[i]);
    s->capt = 60; s->t = (unsigned *)malloc(60 * sizeof(unsigned)); s->nt = 60;
    memcpy(s->t, tris, sizeof(tris));
}

/* edge -> midpoint-vertex map.  The reference finds an existing midpoint by a
 * linear search for an exactly equal vertex among the vertices created in this
 * subdivision (lib/sphere.cpp:197-214).  Two triangles sharing an edge compute
 * the same sums (FP add commutes), so "equal value" == "same edge"; keying on
 * the edge gives the same first-seen order in O(1). */
typedef struct { uint64_t key; unsigned val; } edge_slot;
typedef struct { edge_slot *s; size_t cap; } edge_map;

static unsigned *edge_find(edge_map *m, unsigned a, unsigned b) {
    uint64_t lo = a < b ? a : b, hi = a < b ? b : a;
    uint64_t key = (hi << 32) | lo;
    uint64_t h = key * 0x9E3779B97F4A7C15ull;
    size_t i = (size_t)(h >> 20) & (m->cap - 1);
    for (;;) {
        if (m->s[i].key == key) return &m->s[i].val;
        if (m->s[i].key == UINT64_MAX) { m->s[i].key = key; m->s[i].val = UINT_MAX; return &m->s[i].val; }
        i = (i + 1) & (m->cap - 1);
    }
}

static v3 sph_mid(v3 p, v3 q) {          /* lib/sphere.cpp:166-192 */
    v3 d;
    d.x = (p.x + q.x); d.y = (p.y + q.y); d.z = (p.z + q.z);
    double norm = sqrt(d.x * d.x + d.y * d.y + d.z * d.z);
    d.x /= norm; d.y /= norm; d.z /= norm;
    return d;
}

/* lib/sphere.cpp:148-244 subDivide */
static void sph_subdivide(sphere_t *s) {
    size_t num = s->nt / 3;
    unsigned *nt = (unsigned *)malloc(num * 12 * sizeof(unsigned));
    size_t k = 0;
    edge_map m; m.cap = 1; while (m.cap < num * 4) m.cap <<= 1;
    m.s = (edge_slot *)malloc(m.cap * sizeof(edge_slot));
    for (size_t i = 0; i < m.cap; i++) m.s[i].key = UINT64_MAX;
    for (size_t i = 0; i < num; i++) {
        unsigned ai = s->t[3 * i], bi = s->t[3 * i + 1], ci = s->t[3 * i + 2];
        v3 a = s->v[ai], b = s->v[bi], c = s->v[ci];
        v3 d = sph_mid(a, b);            /* d = a+b */
        v3 e = sph_mid(c, b);            /* e = c+b (reference operand order) */
        v3 f = sph_mid(a, c);            /* f = a+c */
        unsigned *pd = edge_find(&m, ai, bi);
        unsigned *pe = edge_find(&m, bi, ci);
        unsigned *pf = edge_find(&m, ci, ai);
        /* new vertices are appended in the order d, e, f (sphere.cpp:215-226) */
        unsigned di = *pd, ei = *pe, fi = *pf;
        if (di == UINT_MAX) { di = (unsigned)s->nv; sph_push_v(s, d); *pd = di; }
        if (ei == UINT_MAX) { ei = (unsigned)s->nv; sph_push_v(s, e); *pe = ei; }
        if (fi == UINT_MAX) { fi = (unsigned)s->nv; sph_push_v(s, f); *pf = fi; }
        nt[k++] = ai; nt[k++] = di; nt[k++] = fi;
        nt[k++] = di; nt[k++] = bi; nt[k++] = ei;
        nt[k++] = fi; nt[k++] = ei; nt[k++] = ci;
        nt[k++] = fi; nt[k++] = di; nt[k++] = ei;
    }
    free(m.s); free(s->t);
    s->t = nt; s->nt = k; s->capt = k;
}

/* lib/sphere.cpp:247-330 makeUnique: drop z<0; on the equator drop x<0 and
 * (0,-1,0); erase order == stable filter; triangles touching a dropped vertex
 * are erased and the remaining indices shifted down. */
static void sph_make_unique(sphere_t *s) {
    unsigned *remap = (unsigned *)malloc(s->nv * sizeof(unsigned));
    size_t k = 0;
    for (size_t i = 0; i < s->nv; i++) {
        v3 p = s->v[i];
        int drop = 0;
        if (p.z < 0) drop = 1;
        else if (p.z == 0) {
            if (p.x < 0) drop = 1;
            else if (p.x == 0 && p.y == -1) drop = 1;
        }
        if (drop) remap[i] = UINT_MAX; else { remap[i] = (unsigned)k; s->v[k++] = p; }
    }
    s->nv = k;
    size_t kt = 0;
    for (size_t i = 0; i + 2 < s->nt; i += 3) {
        unsigned a = remap[s->t[i]], b = remap[s->t[i + 1]], c = remap[s->t[i + 2]];
        if (a == UINT_MAX || b == UINT_MAX || c == UINT_MAX) continue;
        s->t[kt++] = a; s->t[kt++] = b; s->t[kt++] = c;
    }
    s->nt = kt;
    free(remap);
}

/* lib/sphere.cpp:18-24 fromIcosahedron */
static int sph_build(sphere_t *s, int g) {
    if (g < 0 || g > 8) return 1;
    memset(s, 0, sizeof(*s));
    sph_icosahedron(s);
    for (int i = 0; i < g; i++) sph_subdivide(s);
    sph_make_unique(s);
    return 0;
}

size_t h3do_sphere_count(int g) {
    sphere_t s; if (sph_build(&s, g)) return 0;
    size_t n = s.nv; free(s.v); free(s.t); return n;
}

size_t h3do_sphere_fill(int g, double *xyz, size_t cap) {
    sphere_t s; if (sph_build(&s, g)) return 0;
    size_t n = s.nv;
    if (n > cap) { free(s.v); free(s.t); return 0; }
    for (size_t i = 0; i < n; i++) { xyz[3 * i] = s.v[i].x; xyz[3 * i + 1] = s.v[i].y; xyz[3 * i + 2] = s.v[i].z; }
    free(s.v); free(s.t); return n;
}

size_t h3do_sphere_triangles(int g, unsigned *tri, size_t cap) {
    sphere_t s; if (sph_build(&s, g)) return 0;
    size_t n = s.nt;
    if (tri) { if (n > cap) n = 0; else memcpy(tri, s.t, n * sizeof(unsigned)); }
    free(s.v); free(s.t); return n;
}

/* ======================================================================== */
/* Hough accumulator -- lib/hough.cpp                                        */
/* ======================================================================== */

struct h3do_hough {
    const double *dirs; size_t num_b;
    double dx, max_x; size_t num_x;
    unsigned *votes; size_t cells;
};

static double round_to_nearest(double num) {       /* lib/hough.cpp:22-24 */
    return (num > 0.0) ? floor(num + 0.5) : ceil(num - 0.5);
}

static double norm3(double x, double y, double z) { /* lib/vector3d.cpp:40-42 */
    return sqrt((x * x) + (y * y) + (z * z));
}

/* lib/hough.cpp:26-43 */
h3do_hough *h3do_hough_create(const double *dirs, size_t n_dir, const double minP[3],
                              const double maxP[3], double var_dx) {
    h3do_hough *h = (h3do_hough *)calloc(1, sizeof(*h));
    if (!h) return NULL;
    h->dirs = dirs; h->num_b = n_dir;
    double nmax = norm3(maxP[0], maxP[1], maxP[2]), nmin = norm3(minP[0], minP[1], minP[2]);
    h->max_x = nmax < nmin ? nmin : nmax;            /* std::max(maxP.norm(), minP.norm()) */
    double range_x = 2 * h->max_x;
    h->dx = var_dx;
    if (h->dx == 0.0) h->dx = range_x / 64.0;
    h->num_x = (size_t)round_to_nearest(range_x / h->dx);
    h->cells = h->num_x * h->num_x * h->num_b;
    h->votes = (unsigned *)calloc(h->cells ? h->cells : 1, sizeof(unsigned));
    if (!h->votes) { free(h); return NULL; }
    return h;
}

void h3do_hough_destroy(h3do_hough *h) { if (h) { free(h->votes); free(h); } }
size_t h3do_hough_num_x(const h3do_hough *h) { return h->num_x; }
size_t h3do_hough_num_b(const h3do_hough *h) { return h->num_b; }
double h3do_hough_dx(const h3do_hough *h) { return h->dx; }
double h3do_hough_max_x(const h3do_hough *h) { return h->max_x; }
size_t h3do_hough_cells(const h3do_hough *h) { return h->cells; }
const unsigned *h3do_hough_votes(const h3do_hough *h) { return h->votes; }

/* lib/hough.cpp:67-100 pointVote, restricted to directions [j0, j1) */
static void point_vote(h3do_hough *h, double px, double py, double pz, int add, size_t j0, size_t j1) {
    const size_t num_x = h->num_x, num_b = h->num_b;
    const double max_x = h->max_x, dx = h->dx;
    for (size_t j = j0; j < j1; j++) {
        double bx = h->dirs[3 * j], by = h->dirs[3 * j + 1], bz = h->dirs[3 * j + 2];
        double beta = 1 / (1 + bz);
        double x_new = ((1 - (beta * (bx * bx))) * px) - ((beta * (bx * by)) * py) - (bx * pz);
        double y_new = ((-beta * (bx * by)) * px) + ((1 - (beta * (by * by))) * py) - (by * pz);
        size_t x_i = round_to_nearest((x_new + max_x) / dx);
        size_t y_i = round_to_nearest((y_new + max_x) / dx);
        size_t index = (x_i * num_x * num_b) + (y_i * num_b) + j;
        if (index < h->cells) {          /* flat bound only: rows may alias (hough.cpp:92) */
            if (add) h->votes[index]++; else h->votes[index]--;
        }
    }
}

void h3do_hough_add(h3do_hough *h, const double *xyz, size_t n) {        /* hough.cpp:50-55 */
    for (size_t i = 0; i < n; i++) point_vote(h, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], 1, 0, h->num_b);
}
void h3do_hough_subtract(h3do_hough *h, const double *xyz, size_t n) {   /* hough.cpp:58-63 */
    for (size_t i = 0; i < n; i++) point_vote(h, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], 0, 0, h->num_b);
}

/* inverse map of lib/hough.cpp:114-132 (same integer/double conversions) */
static void inverse_map(const h3do_hough *h, unsigned int index, double a[3], double b[3]) {
    const size_t num_x = h->num_x, num_b = h->num_b;
    const double dx = h->dx, max_x = h->max_x;
    double x = (int)(index / (num_x * num_b));
    index -= (int)x * num_x * num_b;
    x = x * dx - max_x;
    double y = (int)index / num_b;
    index -= (int)y * num_b;
    y = y * dx - max_x;
    double bx = h->dirs[3 * index], by = h->dirs[3 * index + 1], bz = h->dirs[3 * index + 2];
    b[0] = bx; b[1] = by; b[2] = bz;
    a[0] = x * (1 - ((bx * bx) / (1 + bz))) - y * ((bx * by) / (1 + bz));
    a[1] = x * (-((bx * by) / (1 + bz))) + y * (1 - ((by * by) / (1 + bz)));
    a[2] = -x * bx - y * by;
}

/* lib/hough.cpp:103-135 getLine: strict > scan => first maximum wins */
unsigned h3do_hough_get_line(const h3do_hough *h, double a[3], double b[3], uint64_t *flat_index) {
    unsigned int votes = 0, index = 0;
    for (size_t i = 0; i < h->cells; i++) {
        if (h->votes[i] > votes) { votes = h->votes[i]; index = (unsigned)i; }
    }
    if (flat_index) *flat_index = index;
    inverse_map(h, index, a, b);
    return votes;
}

/* ---- multi-threaded variants (CPU arm of bench.py only) ------------------ */
typedef struct { h3do_hough *h; const double *xyz; size_t n; int add; size_t j0, j1;
                 size_t c0, c1; unsigned best; size_t best_i; } mt_job;

static void *mt_vote_fn(void *p) {
    mt_job *j = (mt_job *)p;
    for (size_t i = 0; i < j->n; i++)
        point_vote(j->h, j->xyz[3 * i], j->xyz[3 * i + 1], j->xyz[3 * i + 2], j->add, j->j0, j->j1);
    return NULL;
}

static void mt_vote(h3do_hough *h, const double *xyz, size_t n, int add, int threads) {
    if (threads < 1) threads = 1;
    if ((size_t)threads > h->num_b) threads = (int)h->num_b;
    if (threads <= 1) { if (add) h3do_hough_add(h, xyz, n); else h3do_hough_subtract(h, xyz, n); return; }
    pthread_t *th = (pthread_t *)malloc(threads * sizeof(pthread_t));
    mt_job *jobs = (mt_job *)calloc(threads, sizeof(mt_job));
    for (int t = 0; t < threads; t++) {
        jobs[t].h = h; jobs[t].xyz = xyz; jobs[t].n = n; jobs[t].add = add;
        jobs[t].j0 = h->num_b * t / threads; jobs[t].j1 = h->num_b * (t + 1) / threads;
        pthread_create(&th[t], NULL, mt_vote_fn, &jobs[t]);
    }
    for (int t = 0; t < threads; t++) pthread_join(th[t], NULL);
    free(th); free(jobs);
}
void h3do_hough_add_mt(h3do_hough *h, const double *xyz, size_t n, int threads) { mt_vote(h, xyz, n, 1, threads); }
void h3do_hough_subtract_mt(h3do_hough *h, const double *xyz, size_t n, int threads) { mt_vote(h, xyz, n, 0, threads); }

static void *mt_max_fn(void *p) {
    mt_job *j = (mt_job *)p;
    unsigned best = 0; size_t bi = j->c0;
    int found = 0;
    for (size_t i = j->c0; i < j->c1; i++)
        if (j->h->votes[i] > best) { best = j->h->votes[i]; bi = i; found = 1; }
    j->best = best; j->best_i = found ? bi : SIZE_MAX;
    return NULL;
}

unsigned h3do_hough_get_line_mt(const h3do_hough *h, double a[3], double b[3], uint64_t *flat_index, int threads) {
    if (threads <= 1 || h->cells < 1u << 16) return h3do_hough_get_line(h, a, b, flat_index);
    pthread_t *th = (pthread_t *)malloc(threads * sizeof(pthread_t));
    mt_job *jobs = (mt_job *)calloc(threads, sizeof(mt_job));
    for (int t = 0; t < threads; t++) {
        jobs[t].h = (h3do_hough *)h;
        jobs[t].c0 = h->cells * t / threads; jobs[t].c1 = h->cells * (t + 1) / threads;
        pthread_create(&th[t], NULL, mt_max_fn, &jobs[t]);
    }
    unsigned votes = 0; size_t index = 0;
    for (int t = 0; t < threads; t++) {
        pthread_join(th[t], NULL);
        if (jobs[t].best_i != SIZE_MAX && jobs[t].best > votes) { votes = jobs[t].best; index = jobs[t].best_i; }
    }
    free(th); free(jobs);
    if (flat_index) *flat_index = index;
    inverse_map(h, (unsigned)index, a, b);
    return votes;
}

/* ======================================================================== */
/* PointCloud -- lib/pointcloud.cpp                                          */
/* ======================================================================== */

void h3do_min_max(const double *p, size_t n, double mn[3], double mx[3]) {  /* pointcloud.cpp:41-59 */
    if (n == 0) { for (int c = 0; c < 3; c++) mn[c] = mx[c] = 0; return; }
    for (int c = 0; c < 3; c++) mn[c] = mx[c] = p[c];
    for (size_t i = 0; i < n; i++)
        for (int c = 0; c < 3; c++) {
            if (mn[c] > p[3 * i + c]) mn[c] = p[3 * i + c];
            if (mx[c] < p[3 * i + c]) mx[c] = p[3 * i + c];
        }
}

void h3do_shift_to_origin(double *p, size_t n, double shift[3]) {          /* pointcloud.cpp:18-26 */
    double p1[3], p2[3], ns[3];
    h3do_min_max(p, n, p1, p2);
    for (int c = 0; c < 3; c++) ns[c] = (p1[c] + p2[c]) / 2.0;
    for (size_t i = 0; i < n; i++) for (int c = 0; c < 3; c++) p[3 * i + c] = p[3 * i + c] - ns[c];
    for (int c = 0; c < 3; c++) shift[c] = shift[c] + ns[c];
}

void h3do_mean_value(const double *p, size_t n, double mean[3]) {          /* pointcloud.cpp:29-38 */
    double s[3] = {0, 0, 0};
    for (size_t i = 0; i < n; i++) for (int c = 0; c < 3; c++) s[c] = s[c] + p[3 * i + c];
    if (n > 0) for (int c = 0; c < 3; c++) s[c] = s[c] / (double)n;
    mean[0] = s[0]; mean[1] = s[1]; mean[2] = s[2];
}

size_t h3do_points_close_to_line(const double *p, size_t n, const double a[3], const double b[3],
                                 double dx, uint32_t *idx_out) {           /* pointcloud.cpp:94-105 */
    size_t k = 0;
    for (size_t i = 0; i < n; i++) {
        double px = p[3 * i], py = p[3 * i + 1], pz = p[3 * i + 2];
        double ex = px - a[0], ey = py - a[1], ez = pz - a[2];
        double t = (b[0] * ex + b[1] * ey + b[2] * ez);        /* b * (p - a), vector3d.cpp:57-59 */
        double tx = t * b[0], ty = t * b[1], tz = t * b[2];    /* t*b */
        double qx = a[0] + tx, qy = a[1] + ty, qz = a[2] + tz; /* a + t*b */
        double dxv = px - qx, dyv = py - qy, dzv = pz - qz;
        if (norm3(dxv, dyv, dzv) <= dx) idx_out[k++] = (uint32_t)i;
    }
    return k;
}

size_t h3do_remove_points(double *x, size_t nx, const double *y, size_t ny) {  /* pointcloud.cpp:109-128 */
    if (ny == 0) return nx;
    size_t i, j, k = 0;
    for (i = 0, j = 0; i < nx && j < ny; i++) {
        if (x[3 * i] == y[3 * j] && x[3 * i + 1] == y[3 * j + 1] && x[3 * i + 2] == y[3 * j + 2]) j++;
        else { x[3 * k] = x[3 * i]; x[3 * k + 1] = x[3 * i + 1]; x[3 * k + 2] = x[3 * i + 2]; k++; }
    }
    for (; i < nx; i++) { x[3 * k] = x[3 * i]; x[3 * k + 1] = x[3 * i + 1]; x[3 * k + 2] = x[3 * i + 2]; k++; }
    return k;
}

/* ======================================================================== */
/* orthogonal_LSQ -- wrap.h:121-150 (== lib/hough3dlines.cpp:49-77)          */
/* Third party: Eigen3 (un-vendored, version unpinned; Ubuntu 20.04 = 3.3.7).*/
/* Restated: MatrixXf copy, colwise().mean(), centred^T * centred in float,  */
/* every float sum in the canonical TREE32 order (32 interleaved partial     */
/* sums, partial l = elements l, l+32, ... in index order, then a halving    */
/* tree l += l+16, +8, +4, +2, +1 -- the shape of a vectorised reduction such */
/* as Eigen's), SelfAdjointEigenSolver<MatrixXf>::compute =                  */
/* scale by max|coeff| -> Householder tridiagonalisation -> implicit         */
/* symmetric QR with Wilkinson shift -> ascending sort.  PARITY UNPINNED.    */
/* ======================================================================== */

static void givens_f(float p, float q, float *c, float *s) {  /* Eigen JacobiRotation::makeGivens (real) */
    if (q == 0.0f) { *c = p < 0.0f ? -1.0f : 1.0f; *s = 0.0f; }
    else if (p == 0.0f) { *c = 0.0f; *s = q < 0.0f ? 1.0f : -1.0f; }
    else if (fabsf(p) > fabsf(q)) {
        float t = q / p; float u = sqrtf(1.0f + t * t); if (p < 0.0f) u = -u;
        *c = 1.0f / u; *s = -t * *c;
    } else {
        float t = p / q; float u = sqrtf(1.0f + t * t); if (q < 0.0f) u = -u;
        *s = -1.0f / u; *c = -t * *s;
    }
}

static float hypot_f(float x, float y) {                      /* Eigen numext::hypot */
    float ax = fabsf(x), ay = fabsf(y);
    float p = ax > ay ? ax : ay;
    if (p == 0.0f) return 0.0f;
    float q = ax > ay ? ay : ax;
    float qp = q / p;
    return p * sqrtf(1.0f + qp * qp);
}

static int eig3f_impl(const float m_in[9], float evals[3], float evecs[9], int fixed3) {
    /* lower triangle of the (column-major, symmetric) input */
    float a00 = m_in[0], a10 = m_in[1], a20 = m_in[2], a11 = m_in[4], a21 = m_in[5], a22 = m_in[8];
    /* scale to [-1,1] */
    float scale = fabsf(a00);
    if (fabsf(a10) > scale) scale = fabsf(a10);
    if (fabsf(a20) > scale) scale = fabsf(a20);
    if (fabsf(a11) > scale) scale = fabsf(a11);
    if (fabsf(a21) > scale) scale = fabsf(a21);
    if (fabsf(a22) > scale) scale = fabsf(a22);
    if (scale == 0.0f) scale = 1.0f;
    a00 /= scale; a10 /= scale; a20 /= scale; a11 /= scale; a21 /= scale; a22 /= scale;

    /* Householder step on column 0 (the second step of a 3x3 is the identity) */
    float diag[3], sub[2];
    float q[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};          /* column-major Q */
    if (fixed3) {
        /* Eigen tridiagonalization_inplace_selector<MatrixType,3,false> (fixed-size Matrix3f, as
         * pcl::PCA uses): closed form, eigenvalues only here (Q stays the identity) */
        float v1norm2 = a20 * a20;
        diag[0] = a00;
        if (v1norm2 <= FLT_MIN) { diag[1] = a11; diag[2] = a22; sub[0] = a10; sub[1] = a21; }
        else {
            float beta = sqrtf(a10 * a10 + v1norm2);
            float inv_beta = 1.0f / beta;
            float m01 = a10 * inv_beta, m02 = a20 * inv_beta;
            float qq = 2.0f * m01 * a21 + m02 * (a22 - a11);
            diag[1] = a11 + m02 * qq;
            diag[2] = a22 - m02 * qq;
            sub[0] = beta;
            sub[1] = a21 - m01 * qq;
        }
    } else {
    float tail_sq = a20 * a20;
    if (tail_sq <= FLT_MIN) {
        diag[0] = a00; diag[1] = a11; diag[2] = a22; sub[0] = a10; sub[1] = a21;
    } else {
        float beta = sqrtf(a10 * a10 + tail_sq);
        if (a10 >= 0.0f) beta = -beta;
        float v1 = a20 / (a10 - beta);
        float tau = (beta - a10) / beta;
        /* p = A_sub * (tau * v), v = (1, v1) */
        float u0 = tau, u1 = tau * v1;
        float p0 = a11 * u0 + a21 * u1;
        float p1 = a21 * u0 + a22 * u1;
        float dot = p0 + p1 * v1;
        float alpha = tau * -0.5f * dot;
        float w0 = p0 + alpha, w1 = p1 + alpha * v1;
        /* A_sub -= v w^T + w v^T (lower) */
        a11 = a11 - (w0 + w0);
        a21 = a21 - (v1 * w0 + w1);
        a22 = a22 - (v1 * w1 + w1 * v1);
        diag[0] = a00; diag[1] = a11; diag[2] = a22; sub[0] = beta; sub[1] = a21;
        /* Q = I - tau * (0,1,v1)(0,1,v1)^T */
        float tv = tau * v1;
        q[4] = 1.0f - tau;  q[7] = 0.0f - tau * v1;     /* row 1: (.., 1-tau, -tau*v1) */
        q[5] = 0.0f - tv;   q[8] = 1.0f - tv * v1;      /* row 2 */
    }
    }

    /* implicit symmetric QR on the tridiagonal (Eigen computeFromTridiagonal_impl) */
    const int n = 3;
    int end = n - 1, start = 0, iter = 0;
    const float precision = 2.0f * FLT_EPSILON, consider_zero = FLT_MIN;
    int ok = 1;
    while (end > 0) {
        for (int i = start; i < end; i++)
            if (fabsf(sub[i]) <= (fabsf(diag[i]) + fabsf(diag[i + 1])) * precision || fabsf(sub[i]) <= consider_zero)
                sub[i] = 0.0f;
        while (end > 0 && sub[end - 1] == 0.0f) end--;
        if (end <= 0) break;
        iter++;
        if (iter > 30 * n) { ok = 0; break; }
        start = end - 1;
        while (start > 0 && sub[start - 1] != 0.0f) start--;

        /* tridiagonal_qr_step */
        float td = (diag[end - 1] - diag[end]) * 0.5f;
        float e = sub[end - 1];
        float mu = diag[end];
        if (td == 0.0f) mu -= fabsf(e);
        else {
            float e2 = e * e;
            float h = hypot_f(td, e);
            if (e2 == 0.0f) mu -= (e / (td + (td > 0.0f ? 1.0f : -1.0f))) * (e / h);
            else mu -= e2 / (td + (td > 0.0f ? h : -h));
        }
        float x = diag[start] - mu;
        float z = sub[start];
        for (int k = start; k < end; k++) {
            float c, s;
            givens_f(x, z, &c, &s);
            float sdk = s * diag[k] + c * sub[k];
            float dkp1 = s * sub[k] + c * diag[k + 1];
            diag[k] = c * (c * diag[k] - s * sub[k]) - s * (c * sub[k] - s * diag[k + 1]);
            diag[k + 1] = s * sdk + c * dkp1;
            sub[k] = c * sdk - s * dkp1;
            if (k > start) sub[k - 1] = c * sub[k - 1] - s * z;
            x = sub[k];
            if (k < end - 1) { z = -s * sub[k + 1]; sub[k + 1] = c * sub[k + 1]; }
            for (int r = 0; r < 3; r++) {            /* Q = Q * G on columns k, k+1 */
                float xi = q[k * 3 + r], yi = q[(k + 1) * 3 + r];
                q[k * 3 + r] = c * xi - s * yi;
                q[(k + 1) * 3 + r] = s * xi + c * yi;
            }
        }
    }
    /* ascending selection sort with column swaps */
    if (ok)
        for (int i = 0; i < n - 1; i++) {
            int k = i;
            for (int j = i + 1; j < n; j++) if (diag[j] < diag[k]) k = j;
            if (k > i) {
                float t = diag[i]; diag[i] = diag[k]; diag[k] = t;
                for (int r = 0; r < 3; r++) { float u = q[i * 3 + r]; q[i * 3 + r] = q[k * 3 + r]; q[k * 3 + r] = u; }
            }
        }
    for (int i = 0; i < 3; i++) evals[i] = diag[i] * scale;
    memcpy(evecs, q, sizeof(q));
    return ok ? 0 : 1;
}

int h3do_eig3f(const float m_in[9], float evals[3], float evecs[9]) { return eig3f_impl(m_in, evals, evecs, 0); }
/* SelfAdjointEigenSolver<Matrix3f>: eigenvalues ascending (eigenvectors not produced) */
int h3do_eig3f_fixed(const float m_in[9], float evals[3]) {
    float evecs[9];
    return eig3f_impl(m_in, evals, evecs, 1);
}

/* halving tree over the 32 interleaved partial sums (destroys v) */
static float tree32_fold(float v[32]) {
    for (int o = 16; o; o >>= 1) for (int l = 0; l < o; l++) v[l] = v[l] + v[l + o];
    return v[0];
}

double h3do_orthogonal_lsq(const double *p, size_t n, double a[3], double b[3]) {
    h3do_mean_value(p, n, a);                         /* wrap.h:126 */
    /* wrap.h:129-134: float copy; wrap.h:137: column means in float */
    float part[6][32];
    float mean[3];
    memset(part, 0, sizeof part);
    for (size_t i = 0; i < n; i++) for (int c = 0; c < 3; c++) part[c][i & 31] = part[c][i & 31] + (float)p[3 * i + c];
    for (int c = 0; c < 3; c++) mean[c] = tree32_fold(part[c]) / (float)n;   /* n==0 -> NaN, unused below */
    /* wrap.h:138: scatter = centred^T centred, float, TREE32 over k */
    float s[6];                                       /* xx xy xz yy yz zz */
    memset(part, 0, sizeof part);
    for (size_t i = 0; i < n; i++) {
        float cx = (float)p[3 * i] - mean[0], cy = (float)p[3 * i + 1] - mean[1], cz = (float)p[3 * i + 2] - mean[2];
        float *q0 = &part[0][i & 31];
        q0[0] = q0[0] + cx * cx; q0[32] = q0[32] + cx * cy; q0[64] = q0[64] + cx * cz;
        q0[96] = q0[96] + cy * cy; q0[128] = q0[128] + cy * cz; q0[160] = q0[160] + cz * cz;
    }
    for (int c = 0; c < 6; c++) s[c] = tree32_fold(part[c]);
    float m[9] = {s[0], s[1], s[2], s[1], s[3], s[4], s[2], s[4], s[5]};
    float ev[3], evec[9];
    h3do_eig3f(m, ev, evec);                          /* wrap.h:141-142 */
    b[0] = evec[6]; b[1] = evec[7]; b[2] = evec[8];   /* last column, wrap.h:146 */
    return (double)ev[2];                             /* wrap.h:147 */
}

/* ======================================================================== */
/* Segment post-tests on one accepted line -- wrap.h:261-337 with find_t     */
/* (wrap.h:41-68), find_proj (wrap.h:78-85), segPCA (wrap.h:94-110).          */
/* Third party: Eigen3 fixed-size Vector3d arithmetic (3-term sums reduce as  */
/* x0 + (x1 + x2), Eigen's unrolled redux) and pcl::PCA<PointXYZ> (PCL 1.10,  */
/* un-vendored): float centroid (sequential sum / n), demeaned float cloud,   */
/* covariance = (1/(n-1)) * sum of products -- restated with the canonical    */
/* TREE32 order for the sums (PCL leaves the order to Eigen's GEMM) -- and    */
/* SelfAdjointEigenSolver<Matrix3f>.  PARITY UNPINNED for the PCA part.       */
/* y: the line's final inliers in the CENTRED frame (Y.points), in Y order;   */
/* shift = X.shift; a = anchor (already + shift), b = direction.              */
/* returns 1 when find_t has no solution (b.x == 0): the reference then       */
/* returns 1 from hough3dlines (wrap.h:276-279); out->status = 3.             */
/* ======================================================================== */

static double dot3_eigen(const double u[3], const double v[3]) {   /* Eigen redux, 3 terms */
    return u[0] * v[0] + (u[1] * v[1] + u[2] * v[2]);
}

int h3do_segment_test(const double *y, size_t n, const double shift[3], const double a[3],
                      const double b[3], const h3do_segment_config *cfg, h3do_segment *out) {
    memset(out, 0, sizeof *out);
    for (int c = 0; c < 3; c++) { out->a[c] = a[c]; out->b[c] = b[c]; }
    out->n_points = (uint32_t)n;
    if (n == 0 || cfg->n_radius == 0) { out->status = 1; return 0; }   /* not reachable from the loop */
    double *t_sorted = (double *)malloc(sizeof(double) * n);
    double *norm_sorted = (double *)malloc(sizeof(double) * n);
    float *pf = (float *)malloc(sizeof(float) * 3 * n);
    size_t m = 0;
    double r_front = 0.0, r_back = 0.0;
    double d[3], dd;
    for (int c = 0; c < 3; c++) d[c] = (a[c] + b[c]) - a[c];           /* p_B - p_A, wrap.h:81-82 */
    dd = dot3_eigen(d, d);
    for (size_t i = 0; i < n; i++) {
        double p[3], pa[3], proj[3], e[3], q[3];
        for (int c = 0; c < 3; c++) { p[c] = y[3 * i + c] + shift[c]; pa[c] = p[c] - a[c]; }   /* wrap.h:271 */
        const double num = dot3_eigen(pa, d);
        for (int c = 0; c < 3; c++) { proj[c] = a[c] + (d[c] * num) / dd; e[c] = proj[c] - p[c]; }
        const double radius = sqrt(dot3_eigen(e, e));                  /* wrap.h:275 */
        if (i == 0) r_front = radius;
        r_back = radius;
        if (b[0] == 0.0) {                                             /* wrap.h:43-45, 276-279 */
            out->status = 3;
            free(t_sorted); free(norm_sorted); free(pf);
            return 1;
        }
        const double t = (proj[0] - a[0]) / b[0];                      /* wrap.h:48 */
        for (int c = 0; c < 3; c++) q[c] = a[c] + t * b[c];
        const double nrm = sqrt(dot3_eigen(q, q));
        /* std::upper_bound: first element greater than t (wrap.h:56) */
        size_t lo = 0, hi = m;
        while (lo < hi) { size_t mid = lo + (hi - lo) / 2; if (t < t_sorted[mid]) hi = mid; else lo = mid + 1; }
        memmove(t_sorted + lo + 1, t_sorted + lo, sizeof(double) * (m - lo));
        memmove(norm_sorted + lo + 1, norm_sorted + lo, sizeof(double) * (m - lo));
        t_sorted[lo] = t; norm_sorted[lo] = nrm; m++;
        for (int c = 0; c < 3; c++) pf[3 * i + c] = (float)p[c];       /* wrap.h:99: pcl::PointXYZ(x, y, z) */
    }
    double max_dist = 0.0;                                             /* wrap.h:286-292 */
    for (size_t i = 0; i + 1 < n; i++) {
        const double dist = fabs(norm_sorted[i + 1] - norm_sorted[i]);
        if (dist > max_dist) max_dist = dist;
    }
    const double radius = r_front > r_back ? r_front : r_back;         /* std::max(front, back), wrap.h:295 */
    double closest = cfg->radius_sizes[0];
    double min_diff = fabs(radius - cfg->radius_sizes[0]);
    double max_radius = fabs(radius - cfg->radius_sizes[0]);           /* sic, wrap.h:298 */
    for (uint32_t k = 0; k < cfg->n_radius; k++) {
        const double r = cfg->radius_sizes[k];
        const double cur = fabs(radius - r);
        if (cur < min_diff) { min_diff = cur; closest = r; }
        if (r > max_radius) max_radius = r;
    }
    out->t_min = t_sorted[0]; out->t_max = t_sorted[n - 1];
    out->radius_raw = radius; out->radius = closest; out->max_dist = max_dist;
    out->status = 1;
    if (min_diff < cfg->diag_voxel && max_radius <= closest && max_dist < 2 * cfg->diag_voxel) {   /* wrap.h:312 */
        /* segPCA: pcl::PCA float */
        float cen[3] = {0.0f, 0.0f, 0.0f};
        for (size_t i = 0; i < n; i++) for (int c = 0; c < 3; c++) cen[c] = cen[c] + pf[3 * i + c];
        for (int c = 0; c < 3; c++) cen[c] = cen[c] / (float)n;
        float part[6][32];
        memset(part, 0, sizeof part);
        for (size_t i = 0; i < n; i++) {
            const float cx = pf[3 * i] - cen[0], cy = pf[3 * i + 1] - cen[1], cz = pf[3 * i + 2] - cen[2];
            float *q0 = &part[0][i & 31];
            q0[0] = q0[0] + cx * cx; q0[32] = q0[32] + cx * cy; q0[64] = q0[64] + cx * cz;
            q0[96] = q0[96] + cy * cy; q0[128] = q0[128] + cy * cz; q0[160] = q0[160] + cz * cz;
        }
        const float inv = 1.0f / ((float)n - 1.0f);
        float sc[6];
        for (int c = 0; c < 6; c++) sc[c] = inv * tree32_fold(part[c]);
        float mat[9] = {sc[0], sc[1], sc[2], sc[1], sc[3], sc[4], sc[2], sc[4], sc[5]};
        float ev[3] = {NAN, NAN, NAN};
        int finite = 1;
        for (int c = 0; c < 6; c++) if (!isfinite(sc[c])) finite = 0;
        if (finite) h3do_eig3f_fixed(mat, ev);
        const double e0 = (double)ev[2], e1 = (double)ev[1], e2 = (double)ev[0];   /* descending, wrap.h:107 */
        out->pca_eigenvalues[0] = e0; out->pca_eigenvalues[1] = e1; out->pca_eigenvalues[2] = e2;
        out->pca_coeff = e0 / (e0 + e1 + e2);                          /* wrap.h:315 */
        double p1[3], p2[3], dl[3];
        for (int c = 0; c < 3; c++) {                                  /* wrap.h:317-319 */
            p1[c] = t_sorted[0] * b[c] + a[c]; p2[c] = t_sorted[n - 1] * b[c] + a[c]; dl[c] = p2[c] - p1[c];
        }
        out->length = sqrt(dot3_eigen(dl, dl));
        const double want = 2.0 * closest * out->length /
                            (cfg->rad_2_leaf_ratio * 2 * cfg->diag_voxel * 2 * cfg->diag_voxel);   /* wrap.h:320 */
        /* static_cast<int>: the x86 conversion (cvttsd2si) gives INT_MIN when out of range or NaN */
        int min_nb = (want >= 2147483648.0 || want <= -2147483649.0 || want != want) ? INT_MIN : (int)want;
        out->min_nb_points = min_nb;
        /* points.size() > min_nb_points_seg compares size_t with int: the int converts to unsigned */
        out->status = (out->pca_coeff > cfg->min_pca_coeff && (uint64_t)n > (uint64_t)(int64_t)min_nb) ? 0 : 2;
    }
    free(t_sorted); free(norm_sorted); free(pf);
    return 0;
}

/* drone2WorldSeg (node.cpp:429-447) + heighSegmentCutoff (node.cpp:455-470).  Third party: Eigen3
 * Quaterniond::toRotationMatrix and the fixed-size 3x3 * 3x1 product (each row reduces as
 * x0 + (x1 + x2)).  No reference test pins it: PARITY UNPINNED. */
void h3do_segment_to_world(const h3do_pose *P, h3do_segment *s) {
    const double w = P->q[0], x = P->q[1], y = P->q[2], z = P->q[3];
    const double tx = 2.0 * x, ty = 2.0 * y, tz = 2.0 * z;
    const double twx = tx * w, twy = ty * w, twz = tz * w;
    const double txx = tx * x, txy = ty * x, txz = tz * x;
    const double tyy = ty * y, tyz = tz * y, tzz = tz * z;
    const double R[3][3] = {{1.0 - (tyy + tzz), txy - twz, txz + twy},
                            {txy + twz, 1.0 - (txx + tzz), tyz - twx},
                            {txz - twy, tyz + twx, 1.0 - (txx + tyy)}};
    if (s->status != 0) return;
    for (int i = 0; i < 3; i++) {
        s->a_world[i] = dot3_eigen(R[i], s->a) + P->position[i];
        s->b_world[i] = dot3_eigen(R[i], s->b);
    }
    const double z1 = s->t_min * s->b_world[2] + s->a_world[2];
    const double z2 = s->t_max * s->b_world[2] + s->a_world[2];
    s->world_status = (z1 > P->floor_trim_height || z2 > P->floor_trim_height) ? 1u : 2u;
}

/* ======================================================================== */
/* Pre-filter -- node.cpp:386-421.  Third party: PCL 1.10 (un-vendored):     */
/* PassThrough<PCLPointCloud2> x,y,z (inclusive, drops non-finite field) then */
/* VoxelGrid<PCLPointCloud2> (bbox -> ijk = floor(p*inv_leaf) - min_b ->      */
/* linear index -> sort -> float centroid per voxel, ascending index).        */
/* PCL uses std::sort (order inside a voxel unspecified); the restatement     */
/* fixes it to the stable (input) order.  PARITY UNPINNED.                    */
/* ======================================================================== */

typedef struct { int idx; uint32_t pt; } vox_key;

static void vox_merge_sort(vox_key *a, vox_key *tmp, size_t n) {   /* stable */
    if (n < 2) return;
    size_t h = n / 2;
    vox_merge_sort(a, tmp, h); vox_merge_sort(a + h, tmp, n - h);
    size_t i = 0, j = h, k = 0;
    while (i < h && j < n) tmp[k++] = (a[j].idx < a[i].idx) ? a[j++] : a[i++];
    while (i < h) tmp[k++] = a[i++];
    while (j < n) tmp[k++] = a[j++];
    memcpy(a, tmp, n * sizeof(vox_key));
}

size_t h3do_prefilter(const float *in, size_t n, size_t stride, const float win[6], float leaf,
                      float *out, size_t cap) {
    vox_key *keys = (vox_key *)malloc((n ? n : 1) * sizeof(vox_key));
    size_t m = 0;
    /* PassThrough x then y then z == one conjunctive test, order preserved */
    for (size_t i = 0; i < n; i++) {
        float x = in[i * stride], y = in[i * stride + 1], z = in[i * stride + 2];
        if (!isfinite(x) || !isfinite(y) || !isfinite(z)) continue;
        if (x > win[1] || x < win[0]) continue;
        if (y > win[3] || y < win[2]) continue;
        if (z > win[5] || z < win[4]) continue;
        keys[m++].pt = (uint32_t)i;
    }
    if (m == 0) { free(keys); return 0; }
    float mn[3], mx[3];
    for (int c = 0; c < 3; c++) { mn[c] = FLT_MAX; mx[c] = -FLT_MAX; }
    for (size_t k = 0; k < m; k++)
        for (int c = 0; c < 3; c++) {
            float v = in[keys[k].pt * stride + c];
            if (v < mn[c]) mn[c] = v;
            if (v > mx[c]) mx[c] = v;
        }
    float inv = 1.0f / leaf;
    int min_b[3], max_b[3]; long long div_b[3];
    for (int c = 0; c < 3; c++) {
        min_b[c] = (int)floorf(mn[c] * inv);
        max_b[c] = (int)floorf(mx[c] * inv);
        div_b[c] = (long long)max_b[c] - min_b[c] + 1;
    }
    if (div_b[0] * div_b[1] * div_b[2] > (long long)INT_MAX) { free(keys); return (size_t)-1; }
    int mul[3] = {1, (int)div_b[0], (int)(div_b[0] * div_b[1])};
    for (size_t k = 0; k < m; k++) {
        const float *p = in + keys[k].pt * stride;
        int i0 = (int)(floorf(p[0] * inv) - (float)min_b[0]);
        int i1 = (int)(floorf(p[1] * inv) - (float)min_b[1]);
        int i2 = (int)(floorf(p[2] * inv) - (float)min_b[2]);
        keys[k].idx = i0 * mul[0] + i1 * mul[1] + i2 * mul[2];
    }
    vox_key *tmp = (vox_key *)malloc(m * sizeof(vox_key));
    vox_merge_sort(keys, tmp, m);
    free(tmp);
    size_t nout = 0, k = 0;
    while (k < m) {
        size_t e = k + 1;
        while (e < m && keys[e].idx == keys[k].idx) e++;
        float cx = 0.0f, cy = 0.0f, cz = 0.0f;
        for (size_t t = k; t < e; t++) {
            const float *p = in + keys[t].pt * stride;
            cx = cx + p[0]; cy = cy + p[1]; cz = cz + p[2];
        }
        float cnt = (float)(e - k);
        if (nout < cap) { out[3 * nout] = cx / cnt; out[3 * nout + 1] = cy / cnt; out[3 * nout + 2] = cz / cnt; }
        nout++;
        k = e;
    }
    free(keys);
    return nout;
}

/* ======================================================================== */
/* Iterative loop -- wrap.h:172-259,339-347 / lib/hough3dlines.cpp:200-317   */
/* ======================================================================== */

int h3do_hough3dlines(const double *xyz_in, size_t n_in, int granularity, double opt_dx,
                      int opt_minvotes, int opt_nlines, int mode,
                      h3do_iter *iters, size_t max_iters, uint32_t *pool, size_t pool_cap,
                      h3do_result *res) {
    memset(res, 0, sizeof(*res));
    double *X = (double *)malloc((n_in ? n_in : 1) * 3 * sizeof(double));
    uint32_t *orig = (uint32_t *)malloc((n_in ? n_in : 1) * sizeof(uint32_t));
    size_t n = 0;
    for (size_t i = 0; i < n_in; i++) {               /* wrap.h:175-189 (inputs there are float) */
        double x = xyz_in[3 * i], y = xyz_in[3 * i + 1], z = xyz_in[3 * i + 2];
        if (!(isnan(x) || isinf(x)) && !(isnan(y) || isinf(y)) && !(isnan(z) || isinf(z))) {
            X[3 * n] = x; X[3 * n + 1] = y; X[3 * n + 2] = z; orig[n] = (uint32_t)i; n++;
        }
    }
    res->n_input = (uint32_t)n;
    if (mode == 1) {                                   /* hough3dlines.cpp:170-173,196-199 */
        if (opt_minvotes < 2) opt_minvotes = 2;
        if (n < 2) { res->rc = 1; free(X); free(orig); return 1; }
    }
    double minP[3], maxP[3], minPs[3], maxPs[3];
    h3do_min_max(X, n, minP, maxP);                    /* wrap.h:200 */
    double d = norm3(maxP[0] - minP[0], maxP[1] - minP[1], maxP[2] - minP[2]);
    res->d = d;
    if (d == 0.0) { res->rc = 1; free(X); free(orig); return 1; }   /* wrap.h:202-205 */
    h3do_shift_to_origin(X, n, res->shift);            /* wrap.h:206 */
    h3do_min_max(X, n, minPs, maxPs);                  /* wrap.h:207 */
    if (mode == 1 && opt_dx == 0.0) opt_dx = d / 64.0; /* hough3dlines.cpp:211-213 */
    else if (opt_dx >= d) { res->rc = 1; free(X); free(orig); return 1; }  /* wrap.h:209-212 */

    size_t nb = h3do_sphere_count(granularity);
    double *dirs = (double *)malloc(nb * 3 * sizeof(double));
    h3do_sphere_fill(granularity, dirs, nb);
    h3do_hough *h = h3do_hough_create(dirs, nb, minPs, maxPs, opt_dx);   /* wrap.h:220 */
    if (!h) { res->rc = 1; free(dirs); free(X); free(orig); return 1; }
    res->dx = h->dx; res->max_x = h->max_x; res->num_x = h->num_x; res->num_b = h->num_b;
    h3do_hough_add(h, X, n);                           /* wrap.h:228 */

    double *Y = (double *)malloc((n ? n : 1) * 3 * sizeof(double));
    uint32_t *yi = (uint32_t *)malloc((n ? n : 1) * sizeof(uint32_t));
    size_t ny = 0, used = 0;
    uint32_t nlines = 0, it = 0;
    do {                                               /* wrap.h:237 */
        double a[3], b[3];
        h3do_iter rec; memset(&rec, 0, sizeof(rec));
        h3do_hough_subtract(h, Y, ny);                 /* wrap.h:241 */
        rec.votes = h3do_hough_get_line(h, a, b, &rec.argmax);   /* wrap.h:243 */
        memcpy(rec.a_raw, a, sizeof(a)); memcpy(rec.b_raw, b, sizeof(b));

        ny = h3do_points_close_to_line(X, n, a, b, opt_dx, yi);  /* wrap.h:245 */
        for (size_t k = 0; k < ny; k++) memcpy(Y + 3 * k, X + 3 * yi[k], 3 * sizeof(double));
        rec.n1 = (uint32_t)ny; rec.off1 = (uint32_t)used;
        for (size_t k = 0; k < ny && used < pool_cap; k++) pool[used++] = orig[yi[k]];
        rec.rc1 = h3do_orthogonal_lsq(Y, ny, a, b);    /* wrap.h:247 */
        memcpy(rec.a1, a, sizeof(a)); memcpy(rec.b1, b, sizeof(b));
        rec.off2 = (uint32_t)used;
        if (rec.rc1 == 0.0) { rec.status = 1; if (it < max_iters) iters[it] = rec; it++; break; }

        ny = h3do_points_close_to_line(X, n, a, b, opt_dx, yi);  /* wrap.h:250 */
        for (size_t k = 0; k < ny; k++) memcpy(Y + 3 * k, X + 3 * yi[k], 3 * sizeof(double));
        rec.n2 = (uint32_t)ny;
        for (size_t k = 0; k < ny && used < pool_cap; k++) pool[used++] = orig[yi[k]];
        if ((unsigned)ny < (unsigned int)opt_minvotes) { rec.status = 2; if (it < max_iters) iters[it] = rec; it++; break; }

        rec.rc2 = h3do_orthogonal_lsq(Y, ny, a, b);    /* wrap.h:254 */
        for (int c = 0; c < 3; c++) { rec.a2[c] = a[c]; rec.b2[c] = b[c]; }
        if (rec.rc2 == 0.0) { rec.status = 3; if (it < max_iters) iters[it] = rec; it++; break; }
        for (int c = 0; c < 3; c++) rec.a2[c] = a[c] + res->shift[c];   /* wrap.h:257 */
        nlines++;
        if (it < max_iters) iters[it] = rec;
        it++;

        /* wrap.h:339 removePoints; keep the original numbering alongside */
        {
            size_t i, j, k = 0;
            for (i = 0, j = 0; i < n && j < ny; i++) {
                if (X[3 * i] == Y[3 * j] && X[3 * i + 1] == Y[3 * j + 1] && X[3 * i + 2] == Y[3 * j + 2]) j++;
                else { memmove(X + 3 * k, X + 3 * i, 3 * sizeof(double)); orig[k] = orig[i]; k++; }
            }
            for (; i < n; i++) { memmove(X + 3 * k, X + 3 * i, 3 * sizeof(double)); orig[k] = orig[i]; k++; }
            n = k;
        }
    } while ((n > 1) && ((opt_nlines == 0) || (opt_nlines > (int)nlines)));   /* wrap.h:341-342 */

    res->n_iters = it; res->n_lines = nlines; res->pool_used = used;
    h3do_hough_destroy(h);
    free(dirs); free(X); free(orig); free(Y); free(yi);
    return 0;
}
