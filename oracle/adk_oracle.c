/* TEST INFRASTRUCTURE — see adk_oracle.h.  Plain C99, IEEE binary64, no FMA contraction (-ffp-contract=off).
 *
 * CPU restatement of the reference's area-bounded (Eppstein) and area+diameter-bounded (AntipodalOptimized)
 * largest-convex-subset solvers.  All citations are file:line relative to /root/reference/src/common/.
 * Differences from the reference that do not change results:
 *   - double instead of long double (the reference author's platform has long double == double; the
 *     x87 build in oracle/_ref is the arbiter and tests/test_oracle_cpu.py compares the two);
 *   - the slope list of (anchor, centre) is computed once instead of once per level m (it does not depend on m);
 *   - per-layer open-addressing tables instead of std::unordered_map (iteration order only matters on exact ties);
 *   - stable merge sorts instead of std::sort (identical for strict orders).
 */
#include "adk_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define ADKO_EPS 1e-10       /* CustomMath.h:46-47 */
#define ADKO_ANGLE_TOL 1e-6  /* Utils.cpp:301 */
#define ADKO_TWO_PI (2.0 * M_PI)

typedef struct { const double* x; const double* y; uint32_t n; } pts_t;

/* ---------------------------------------------------------------- scalar predicates (CustomMath.cpp) */

static inline double sarea_xy(double ax, double ay, double bx, double by, double cx, double cy)
{   /* CustomMath.cpp:32-36 */
    return 0.5 * ((bx - ax) * (cy - ay) - (cx - ax) * (by - ay));
}
static inline double sarea(const pts_t* P, int a, int b, int c)
{
    return sarea_xy(P->x[a], P->y[a], P->x[b], P->y[b], P->x[c], P->y[c]);
}
static inline int near_zero(double v) { return fabs(v) < ADKO_EPS; }               /* CustomMath.cpp:22-25 */
static inline int collinear(const pts_t* P, int a, int b, int c) { return near_zero(sarea(P, a, b, c)); } /* :38-41 */
static inline int orient_xy(double ax, double ay, double bx, double by, double cx, double cy)
{   /* CustomMath.cpp:43-54: -1 right turn, 0 collinear, +1 left turn */
    const double s = sarea_xy(ax, ay, bx, by, cx, cy);
    if (near_zero(s)) return 0;
    return s > 0 ? 1 : -1;
}
static inline int orient(const pts_t* P, int a, int b, int c)
{
    return orient_xy(P->x[a], P->y[a], P->x[b], P->y[b], P->x[c], P->y[c]);
}
static inline double dist2_xy(double ax, double ay, double bx, double by)
{   /* CustomMath.cpp:15-20 */
    const double dx = bx - ax, dy = by - ay;
    return dx * dx + dy * dy;
}
static inline double dist2(const pts_t* P, int a, int b) { return dist2_xy(P->x[a], P->y[a], P->x[b], P->y[b]); }
static inline double angle_xy(double rx, double ry, double px, double py)
{   /* CustomMath.cpp:56-65 */
    const double a = atan2(py - ry, px - rx);
    return a < 0 ? a + ADKO_TWO_PI : a;
}
static inline double pseudo_dist(const pts_t* P, int s, int e, int p)
{   /* CustomMath.cpp:89-93 */
    return fabs((P->x[p] - P->x[s]) * (-P->y[e] + P->y[s]) + (P->y[p] - P->y[s]) * (P->x[e] - P->x[s]));
}

/* ArePointsClockwise (Utils.cpp:295-309) on explicit coordinates with precomputed angles. */
static inline int cw_before_xy(double rx, double ry, double ax, double ay, double aang,
                               double bx, double by, double bang)
{
    if (fabs(aang - bang) < ADKO_ANGLE_TOL && near_zero(sarea_xy(rx, ry, ax, ay, bx, by)))
        return dist2_xy(rx, ry, ax, ay) < dist2_xy(rx, ry, bx, by);
    return aang > bang;
}

/* ---------------------------------------------------------------- generic stable merge sort of int32 keys */

typedef int (*less_fn)(const void* ctx, int32_t a, int32_t b);

static void msort(int32_t* v, uint32_t n, less_fn less, const void* ctx)
{
    if (n < 2) return;
    int32_t* tmp = (int32_t*) malloc(sizeof(int32_t) * n);
    int32_t *src = v, *dst = tmp;
    for (uint32_t w = 1; w < n; w *= 2)
    {
        for (uint32_t lo = 0; lo < n; lo += 2 * w)
        {
            uint32_t mid = lo + w < n ? lo + w : n, hi = lo + 2 * w < n ? lo + 2 * w : n;
            uint32_t a = lo, b = mid, o = lo;
            while (a < mid && b < hi) dst[o++] = less(ctx, src[b], src[a]) ? src[b++] : src[a++];
            while (a < mid) dst[o++] = src[a++];
            while (b < hi) dst[o++] = src[b++];
        }
        int32_t* t = src; src = dst; dst = t;
    }
    if (src != v) memcpy(v, src, sizeof(int32_t) * n);
    free(tmp);
}

/* ---------------------------------------------------------------- clockwise neighbour lists (Utils.cpp:311-316) */

typedef struct { const pts_t* P; int c; const double* ang; } cw_ctx;

static int cw_less(const void* vctx, int32_t a, int32_t b)
{
    const cw_ctx* c = (const cw_ctx*) vctx;
    const pts_t* P = c->P;
    return cw_before_xy(P->x[c->c], P->y[c->c], P->x[a], P->y[a], c->ang[a], P->x[b], P->y[b], c->ang[b]);
}

/* order: n*(n-1) indices; ang (optional): n*n angles, ang[c*n+p] = Angle(c, p). */
static void build_cw(const pts_t* P, int32_t* order, double* ang)
{
    const uint32_t n = P->n;
    double* a = (double*) malloc(sizeof(double) * n);
    for (uint32_t c = 0; c < n; ++c)
    {
        int32_t* row = order + (size_t) c * (n - 1);
        uint32_t k = 0;
        for (uint32_t p = 0; p < n; ++p)
        {
            a[p] = p == c ? 0.0 : angle_xy(P->x[c], P->y[c], P->x[p], P->y[p]);
            if (p != c) row[k++] = (int32_t) p;
        }
        cw_ctx ctx = { P, (int) c, a };
        msort(row, n - 1, cw_less, &ctx);
        if (ang) memcpy(ang + (size_t) c * n, a, sizeof(double) * n);
    }
    free(a);
}

int adko_cw_order(const double* x, const double* y, uint32_t n, int32_t* order)
{
    pts_t P = { x, y, n };
    if (n < 2) return 0;
    build_cw(&P, order, NULL);
    return 0;
}

/* ---------------------------------------------------------------- count tables (Utils.cpp:74-143) */

typedef struct { int32_t* below; int32_t* coll; int32_t* rb; uint32_t n; } tables_t;
#define BELOW(T, a, b) ((T)->below[(size_t) (a) * (T)->n + (b)])
#define COLL(T, a, b) ((T)->coll[(size_t) (a) * (T)->n + (b)])

static int xy_less(const void* ctx, int32_t a, int32_t b)
{   /* SortPointsHorizontally, CustomMath.h:21-23 */
    const pts_t* P = (const pts_t*) ctx;
    return P->x[a] < P->x[b] || (P->x[a] == P->x[b] && P->y[a] < P->y[b]);
}
static int y_less(const void* ctx, int32_t a, int32_t b)
{   /* SortPointsVertically, CustomMath.h:25-27 */
    const pts_t* P = (const pts_t*) ctx;
    return P->y[a] < P->y[b];
}

/* The O(n^2) sweep.  For the i-th point in (x,y) order, walk the i points to its left in clockwise
 * order starting straight below it, deriving every count from the previous neighbour (Utils.cpp:86-144). */
static void sweep_tables(const pts_t* P, const int32_t* cw, tables_t* T)
{
    const uint32_t n = P->n, len = n - 1;
    memset(T->below, 0, sizeof(int32_t) * (size_t) n * n);
    memset(T->coll, 0, sizeof(int32_t) * (size_t) n * n);
    memset(T->rb, 0, sizeof(int32_t) * n);
    int32_t* byx = (int32_t*) malloc(sizeof(int32_t) * n);
    for (uint32_t i = 0; i < n; ++i) byx[i] = (int32_t) i;
    msort(byx, n, xy_less, P);
    for (uint32_t i = 1; i < n; ++i)
    {
        const int pi = byx[i];
        const int32_t* ring = cw + (size_t) pi * len;
        /* locGetFirstClockWiseLeftIndex, Utils.cpp:55-65 */
        uint32_t start = 0;
        for (uint32_t t = 0; t + 1 < len; ++t)
            if (P->x[ring[t]] > P->x[pi] && P->x[ring[t + 1]] <= P->x[pi]) { start = t + 1; break; }
        if (P->x[pi] == P->x[byx[i - 1]]) T->rb[pi] = 1 + T->rb[byx[i - 1]];   /* :92-95 */
        const uint32_t stop = (start + i) % len;
        for (uint32_t j = (start + 1) % len, jp = start; j != stop; jp = j, j = (j + 1) % len)
        {
            const int q = ring[j], qp = ring[jp];
            if (collinear(P, pi, qp, q))
            {   /* :106-115 */
                COLL(T, pi, q) = 1 + COLL(T, pi, qp);
                if (P->x[q] == P->x[pi]) continue;
                BELOW(T, q, pi) = BELOW(T, pi, qp) + BELOW(T, qp, q) + T->rb[qp];
            }
            else if (P->x[q] > P->x[qp])
            {   /* :116-125 */
                if (P->x[q] == P->x[pi]) continue;
                BELOW(T, q, pi) = (BELOW(T, pi, qp) + COLL(T, pi, qp)) - BELOW(T, q, qp) - T->rb[q];
            }
            else if (P->x[q] < P->x[qp])
            {   /* :126-136 */
                int v = BELOW(T, pi, qp) + COLL(T, pi, qp) + BELOW(T, qp, q) + COLL(T, qp, q) + T->rb[qp];
                if (P->x[qp] != P->x[pi]) v += 1;
                BELOW(T, q, pi) = v;
            }
            else
            {   /* :137-140 */
                BELOW(T, q, pi) = BELOW(T, qp, pi);
            }
            COLL(T, q, pi) = COLL(T, pi, q);      /* :142 */
            BELOW(T, pi, q) = BELOW(T, q, pi);    /* :143 */
        }
    }
    free(byx);
}

/* Definitions spelled out by the reference's disabled self-check (Utils.cpp:145-223). */
static void brute_tables(const pts_t* P, tables_t* T)
{
    const uint32_t n = P->n;
    memset(T->below, 0, sizeof(int32_t) * (size_t) n * n);
    memset(T->coll, 0, sizeof(int32_t) * (size_t) n * n);
    memset(T->rb, 0, sizeof(int32_t) * n);
    for (uint32_t a = 0; a < n; ++a)
    {
        /* rightBelow: same-x points that precede a in (x,y) order (Utils.cpp:92-95) */
        int rb = 0;
        for (uint32_t p = 0; p < n; ++p)
            if (p != a && P->x[p] == P->x[a] && P->y[p] < P->y[a]) ++rb;
        T->rb[a] = rb;
        for (uint32_t b = 0; b < n; ++b)
        {
            if (a == b) continue;
            /* r = the later point in (x,y) order, l = the other (the sweep's pi / pj) */
            const int r = xy_less(P, (int32_t) a, (int32_t) b) ? (int) b : (int) a;
            const int l = r == (int) a ? (int) b : (int) a;
            int cc = 0, bb = 0;
            const double d2 = dist2(P, r, l);
            for (uint32_t p = 0; p < n; ++p)
            {
                if ((int) p == r || (int) p == l) continue;
                if (collinear(P, r, l, (int) p))
                {   /* :148-162 */
                    const double dot = (P->x[l] - P->x[r]) * (P->x[p] - P->x[r]) + (P->y[l] - P->y[r]) * (P->y[p] - P->y[r]);
                    if (dot > 0 && dot < d2) ++cc;
                }
                if (P->x[p] > P->x[l] && P->x[p] < P->x[r] && orient(P, r, l, (int) p) == 1) ++bb;   /* :176-187 */
            }
            COLL(T, a, b) = cc;
            BELOW(T, a, b) = P->x[a] == P->x[b] ? 0 : bb;
        }
    }
}

/* PointsInTriangle (Utils.cpp:251-293) */
static int pit(const pts_t* P, const tables_t* T, int a, int b, int c)
{
    if (collinear(P, a, b, c)) return 0;
    int l = a, m = b, r = c, t;
    /* three-element sort by (x, y): CustomMath.h:21-23 */
    if (xy_less(P, m, l)) { t = l; l = m; m = t; }
    if (xy_less(P, r, m)) { t = m; m = r; r = t; }
    if (xy_less(P, m, l)) { t = l; l = m; m = t; }
    if (P->x[l] == P->x[m]) return BELOW(T, m, r) - BELOW(T, l, r) - COLL(T, l, r);              /* :269-274 */
    if (P->x[m] == P->x[r]) return BELOW(T, l, r) - BELOW(T, l, m) - COLL(T, l, m);              /* :275-280 */
    if (orient(P, r, l, m) == 1)                                                                  /* :281-286 */
        return BELOW(T, l, r) - BELOW(T, l, m) - BELOW(T, m, r) - COLL(T, l, m) - COLL(T, m, r) - T->rb[m] - 1;
    return BELOW(T, l, m) + BELOW(T, m, r) + T->rb[m] - BELOW(T, l, r) - COLL(T, l, r);          /* :287-292 */
}

static tables_t alloc_tables(uint32_t n)
{
    tables_t T;
    T.n = n;
    T.below = (int32_t*) malloc(sizeof(int32_t) * (size_t) n * n);
    T.coll = (int32_t*) malloc(sizeof(int32_t) * (size_t) n * n);
    T.rb = (int32_t*) malloc(sizeof(int32_t) * n);
    return T;
}
static void free_tables(tables_t* T) { free(T->below); free(T->coll); free(T->rb); }

int adko_count_tables(const double* x, const double* y, uint32_t n, int32_t* below, int32_t* coll, int32_t* right_below)
{
    pts_t P = { x, y, n };
    tables_t T = { below, coll, right_below, n };
    if (n < 2) return 0;
    int32_t* cw = (int32_t*) malloc(sizeof(int32_t) * (size_t) n * (n - 1));
    build_cw(&P, cw, NULL);
    sweep_tables(&P, cw, &T);
    free(cw);
    return 0;
}

int adko_count_tables_bruteforce(const double* x, const double* y, uint32_t n, int32_t* below, int32_t* coll, int32_t* right_below)
{
    pts_t P = { x, y, n };
    tables_t T = { below, coll, right_below, n };
    brute_tables(&P, &T);
    return 0;
}

int adko_points_in_triangle(const double* x, const double* y, uint32_t n, const uint32_t* tri, uint32_t m, int32_t* out)
{
    pts_t P = { x, y, n };
    tables_t T = alloc_tables(n);
    int32_t* cw = (int32_t*) malloc(sizeof(int32_t) * (size_t) n * (n - 1));
    build_cw(&P, cw, NULL);
    sweep_tables(&P, cw, &T);
    for (uint32_t t = 0; t < m; ++t) out[t] = pit(&P, &T, (int) tri[3 * t], (int) tri[3 * t + 1], (int) tri[3 * t + 2]);
    free(cw);
    free_tables(&T);
    return 0;
}

/* ---------------------------------------------------------------- rotating calipers (Utils.h:57-109, Utils.cpp:318-338) */

typedef struct { double best; int a, b; const pts_t* P; const int32_t* hull; } diam_acc;

static void diam_visit(diam_acc* acc, uint32_t a, uint32_t b)
{   /* the lambda of ComputeDiameter, Utils.cpp:328-335 */
    const double d = dist2(acc->P, acc->hull[a], acc->hull[b]);
    if (d > acc->best) { acc->best = d; acc->a = (int) a; acc->b = (int) b; }
}

static void hull_diameter(const pts_t* P, const int32_t* hull, uint32_t cnt, uint32_t out[2])
{
#define HA(p, q, r) sarea(P, hull[p], hull[q], hull[r])
#define NX(v) (((v) + 1) % cnt)
    diam_acc acc = { 0.0, 0, 0, P, hull };
    /* Step budget: on a degenerate cycle (collinear vertices, exactly equal areas in binary64 where the reference's 80-bit
     * arithmetic happens to break the ties) the walk can circle forever; it then falls back to the plain farthest pair.
     * Same rule as the CUDA path (adk_area.cu, hull_diameter_dev). */
    uint32_t budget = 8 * cnt + 32;
    int gave_up = 0;
    if (cnt > 2)
    {
        uint32_t pi = cnt - 1, qi = 0;
        while (HA(pi, NX(pi), NX(qi)) > HA(pi, NX(pi), qi)) { qi = NX(qi); if (--budget == 0) { gave_up = 1; break; } }
        const uint32_t q0 = qi;
        while (qi != 0 && !gave_up)
        {
            if (--budget == 0) { gave_up = 1; break; }
            pi = NX(pi);
            diam_visit(&acc, pi, qi);
            while (HA(pi, NX(pi), NX(qi)) > HA(pi, NX(pi), qi))
            {
                if (--budget == 0) { gave_up = 1; break; }
                qi = NX(qi);
                if (pi != q0 || qi != 0) diam_visit(&acc, pi, qi);
                else break;
            }
            if (gave_up) break;
            const double diff = HA(pi, NX(pi), NX(qi)) - HA(pi, NX(pi), qi);
            if (near_zero(diff))
            {
                if (pi != q0 || qi != cnt - 1) diam_visit(&acc, pi, NX(qi));
                else break;
            }
        }
    }
    if (gave_up)
    {
        acc.best = 0.0; acc.a = 0; acc.b = 0;
        for (uint32_t a = 0; a < cnt; ++a) for (uint32_t b = a + 1; b < cnt; ++b) diam_visit(&acc, a, b);
    }
    out[0] = (uint32_t) hull[acc.a];
    out[1] = (uint32_t) hull[acc.b];
#undef HA
#undef NX
}

/* ---------------------------------------------------------------- path A: slope order (Eppstein.cpp:45-114) */

typedef struct { double ang, d2; int32_t pos; } wrap_t;   /* the reference's Wrapper tuple: (angle, dist, position in the cw list) */

static int wrap_less(const wrap_t* a, const wrap_t* b)
{   /* std::less<> on the tuple */
    if (a->ang != b->ang) return a->ang < b->ang;
    if (a->d2 != b->d2) return a->d2 < b->d2;
    return a->pos < b->pos;
}

/* Slope order of the points around centre j seen from anchor i.  ring = clockwise list around j.
 * scratch: 2*n wrap_t + 2*n int32.  Writes POSITIONS in ring to out; returns the length. */
static uint32_t slope_positions(const pts_t* P, const int32_t* ring, int i, int j, wrap_t* wl, wrap_t* wr,
                                int32_t* sl, int32_t* sr, int32_t* out)
{
    const uint32_t len = P->n - 1;
    const double jx = P->x[j], jy = P->y[j];
    uint32_t nl = 0, nr = 0;
    for (uint32_t p = 0; p < len; ++p)
    {   /* :69-92 */
        const int q = ring[p];
        if (P->y[q] < P->y[i]) continue;
        wrap_t w = { angle_xy(jx, jy, P->x[q], P->y[q]), dist2_xy(jx, jy, P->x[q], P->y[q]), (int32_t) p };
        const int o = orient(P, i, j, q);
        if (o == 0) { if (P->y[q] > jy) wl[nl++] = w; }
        else if (o < 0) wr[nr++] = w;
        else wl[nl++] = w;
    }
    /* locCollectSortedPoints (:45-60): rotate each list so that it starts at its maximum */
    uint32_t cl = 0, cr = 0;
    if (nl)
    {
        uint32_t best = 0;
        for (uint32_t t = 1; t < nl; ++t) if (wrap_less(&wl[best], &wl[t])) best = t;
        for (uint32_t t = 0; t < nl; ++t) sl[cl++] = wl[(best + t) % nl].pos;
    }
    if (nr)
    {   /* comparator of :97-101: angle + 2*pi for points strictly above the centre, then distance */
        uint32_t best = 0;
        double ba = wr[0].ang + (P->y[ring[wr[0].pos]] > jy ? ADKO_TWO_PI : 0.0);
        for (uint32_t t = 1; t < nr; ++t)
        {
            const double ta = wr[t].ang + (P->y[ring[wr[t].pos]] > jy ? ADKO_TWO_PI : 0.0);
            if (ba < ta || (ba == ta && wr[best].d2 < wr[t].d2)) { best = t; ba = ta; }
        }
        for (uint32_t t = 0; t < nr; ++t) sr[cr++] = wr[(best + t) % nr].pos;
    }
    /* std::merge(left, right, comp(right, left)) with the reflected right point (:103-113) */
    uint32_t a = 0, b = 0, o = 0;
    while (a < cl && b < cr)
    {
        const int qr = ring[sr[b]], ql = ring[sl[a]];
        const double ox = 2.0 * jx - P->x[qr], oy = 2.0 * jy - P->y[qr];
        const int take_right = cw_before_xy(jx, jy, ox, oy, angle_xy(jx, jy, ox, oy),
                                            P->x[ql], P->y[ql], angle_xy(jx, jy, P->x[ql], P->y[ql]));
        out[o++] = take_right ? sr[b++] : sl[a++];
    }
    while (a < cl) out[o++] = sl[a++];
    while (b < cr) out[o++] = sr[b++];
    return o;
}

int adko_slope_order(const double* x, const double* y, uint32_t n, uint32_t i, uint32_t j, int32_t* out)
{
    pts_t P = { x, y, n };
    if (n < 3 || i >= n || j >= n || i == j) return -1;
    int32_t* cw = (int32_t*) malloc(sizeof(int32_t) * (size_t) n * (n - 1));
    build_cw(&P, cw, NULL);
    wrap_t* w = (wrap_t*) malloc(sizeof(wrap_t) * 2 * n);
    int32_t* s = (int32_t*) malloc(sizeof(int32_t) * 3 * n);
    const int32_t* ring = cw + (size_t) j * (n - 1);
    const uint32_t cnt = slope_positions(&P, ring, (int) i, (int) j, w, w + n, s, s + n, s + 2 * n);
    for (uint32_t t = 0; t < cnt; ++t) out[t] = ring[s[2 * n + t]];
    free(s); free(w); free(cw);
    return (int) cnt;
}

/* ---------------------------------------------------------------- path A: the chain DP (Eppstein.cpp:128-277) */

static inline int tri_count_A(const pts_t* P, const tables_t* T, uint64_t m, int i, int j, int l)
{   /* locPointsWithinTriangleCount, Eppstein.cpp:14-30 */
    int c = COLL(T, i, j) + COLL(T, j, l) + pit(P, T, i, j, l);
    if (m == 3 && !collinear(P, i, j, l)) c += COLL(T, l, i);
    return c;
}

static double now_ms(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

/* The anchors of one solve are independent (Eppstein.cpp:177-238 is a loop over them with a running best that only a strictly
 * better (count, area) replaces), so a solve can be dealt over processes: shard s of nshards takes every nshards-th anchor of the
 * bottom-up order, keeps its own best and the table of its own best anchor, and reports the position of that anchor in the order so
 * that the merge can keep the first of equals as the loop does.  tests/golden/make_oracle_sharded.py --area. */
static int area_solve_impl(const double* x, const double* y, uint32_t n, uint64_t kmax, double amax,
                           int reconstruct, adko_result* out, uint32_t* hull, uint32_t shard, uint32_t nshards, uint32_t* best_pos);

int adko_solve_area(const double* x, const double* y, uint32_t n, uint64_t kmax, double amax,
                    int reconstruct, adko_result* out, uint32_t* hull)
{
    return area_solve_impl(x, y, n, kmax, amax, reconstruct, out, hull, 0, 1, NULL);
}

int adko_solve_area_shard(const double* x, const double* y, uint32_t n, uint64_t kmax, double amax, int reconstruct,
                          uint32_t shard, uint32_t nshards, adko_result* out, uint32_t* hull, uint32_t* best_anchor_pos)
{
    if (nshards == 0 || shard >= nshards) return 1;
    return area_solve_impl(x, y, n, kmax, amax, reconstruct, out, hull, shard, nshards, best_anchor_pos);
}

static int area_solve_impl(const double* x, const double* y, uint32_t n, uint64_t kmax, double amax,
                           int reconstruct, adko_result* out, uint32_t* hull, uint32_t shard, uint32_t nshards, uint32_t* best_pos)
{
    memset(out, 0, sizeof(*out));
    out->entries_alloc = -1; out->entries_required = -1;
    if (n < 3 || kmax < 3) return 0;   /* :137-140 */
    const double t0 = now_ms();
    pts_t P = { x, y, n };
    const uint32_t len = n - 1;
    int32_t* cw = (int32_t*) malloc(sizeof(int32_t) * (size_t) n * len);
    build_cw(&P, cw, NULL);                       /* :143-151 */
    tables_t T = alloc_tables(n);
    sweep_tables(&P, cw, &T);                     /* :154-155 */
    int32_t* byy = (int32_t*) malloc(sizeof(int32_t) * n);
    for (uint32_t i = 0; i < n; ++i) byy[i] = (int32_t) i;
    msort(byy, n, y_less, &P);                    /* :158-159 */

    const uint64_t K = kmax < n ? kmax : n;       /* :162 */
    const size_t plane = (size_t) n * n;
    const size_t total = (size_t) (K + 1) * 2 * plane;
    double* tab = (double*) malloc(sizeof(double) * total);
    if (!tab) return 2;
    for (size_t t = 0; t < total; ++t) tab[t] = INFINITY;
#define TAB(m, buf, a, b) tab[(size_t) (m) * 2 * plane + (size_t) (buf) * plane + (size_t) (a) * n + (b)]
    out->entries_alloc = (int64_t) total;         /* :171-175 */
    out->entries_required = 0;

    /* per-anchor hoisted slope lists: for the p-th centre in j order, lists[p*n ..] holds POINT indices */
    int32_t* lists = (int32_t*) malloc(sizeof(int32_t) * (size_t) n * n);
    uint32_t* llen = (uint32_t*) malloc(sizeof(uint32_t) * n);
    int32_t* jorder = (int32_t*) malloc(sizeof(int32_t) * n);
    wrap_t* w = (wrap_t*) malloc(sizeof(wrap_t) * 2 * n);
    int32_t* s = (int32_t*) malloc(sizeof(int32_t) * 3 * n);

    int have = 0, flip = 0, buf = 0;
    uint64_t best_m = 0;
    double best_area = INFINITY;
    int best_i = -1, best_j = -1;

    uint32_t best_a = 0;
    for (uint32_t a = 0; a < n; ++a)
    {
        if (a % nshards != shard) continue;       /* another shard's anchor */
        const int i = byy[a];
        if (flip) { buf ^= 1; flip = 0; }         /* :180-184 */
        for (size_t t = 0; t < plane; ++t) TAB(2, buf, 0, t) = 0.0;   /* :191 */
        const int32_t* ring_i = cw + (size_t) i * len;
        /* locGetFirstClockWiseUpIndex (:32-42) and the j loop header (:196) */
        uint32_t start = 0;
        for (uint32_t t = 0; t + 1 < len; ++t)
            if (P.y[ring_i[t]] < P.y[i] && P.y[ring_i[t + 1]] >= P.y[i]) { start = t + 1; break; }
        uint32_t nj = 0;
        for (uint32_t t = start, c = 0; c < len && P.y[ring_i[t]] >= P.y[i]; ++c, t = (t + 1) % len) jorder[nj++] = ring_i[t];
        for (uint32_t p = 0; p < nj; ++p)
        {
            const int j = jorder[p];
            const int32_t* ring_j = cw + (size_t) j * len;
            llen[p] = slope_positions(&P, ring_j, i, j, w, w + n, s, s + n, s + 2 * n);
            for (uint32_t t = 0; t < llen[p]; ++t) lists[(size_t) p * n + t] = ring_j[s[2 * n + t]];
        }
        for (uint64_t m = 3; m <= K; ++m)         /* :192 */
        {
            for (uint32_t p = 0; p < nj; ++p)
            {
                const int j = jorder[p];
                double run = INFINITY, last = -INFINITY;   /* :202-203 */
                for (uint32_t t = 0; t < llen[p]; ++t)
                {
                    const int l = lists[(size_t) p * n + t];
                    ++out->cells;
                    const uint64_t c = 1 + (uint64_t) tri_count_A(&P, &T, m, i, j, l);     /* :209 */
                    if (c <= m && orient(&P, i, j, l) >= 0)                                /* :210 */
                    {
                        const double cur = TAB(m - c, buf, l, j) + fabs(sarea(&P, i, j, l));   /* :212-214 */
                        if (cur < run && cur <= amax)                                      /* :216 */
                        {
                            run = cur;
                            if (!have || m > best_m || (m == best_m && run < best_area))   /* :219 */
                            {
                                flip |= !have || best_i != i;
                                have = 1; best_m = m; best_area = run; best_i = i; best_j = j; best_a = a;
                            }
                        }
                    }
                    if (run != last) { last = run; out->entries_required += 1; }           /* :228-232 */
                    TAB(m, buf, j, l) = run;                                               /* :234 */
                }
            }
        }
    }

    if (best_pos) *best_pos = best_a;
    if (have)
    {
        out->found = 1;
        out->count = (uint32_t) best_m;
        out->area = best_area;
        if (reconstruct)
        {   /* :240-274 */
            buf = flip ? buf : (buf ^ 1);
            uint64_t m = best_m;
            int j = best_j;
            uint32_t h = 0;
            int32_t* hl = (int32_t*) malloc(sizeof(int32_t) * (n + 2));
            hl[h++] = best_i;
            while (m > 2 && h < n)          /* h < n: on exactly degenerate inputs a step can consume no point (count 0) */
            {
                const int lastp = hl[h - 1];
                const int32_t* ring_j = cw + (size_t) j * len;
                const uint32_t cnt = slope_positions(&P, ring_j, best_i, j, w, w + n, s, s + n, s + 2 * n);
                int l = -1;
                for (uint32_t k = cnt; k-- > 0;)
                {
                    const int cur = ring_j[s[2 * n + k]];
                    const int changed = k == 0 || TAB(m, buf, j, cur) != TAB(m, buf, j, ring_j[s[2 * n + k - 1]]);
                    if (changed && orient(&P, lastp, j, cur) >= 0) { l = cur; break; }
                }
                if (l < 0) break;   /* cannot happen on a table this DP produced */
                m -= 1 + (uint64_t) tri_count_A(&P, &T, m, best_i, j, l);
                hl[h++] = j;
                j = l;
            }
            hl[h++] = j;
            out->hull_len = h;
            for (uint32_t t = 0; t < h; ++t) hull[t] = (uint32_t) hl[t];
            hull_diameter(&P, hl, h, out->diam);
            out->diam_valid = 1;
            free(hl);
        }
    }
#undef TAB
    free(s); free(w); free(jorder); free(llen); free(lists); free(tab); free(byy); free_tables(&T); free(cw);
    out->solve_ms = now_ms() - t0;
    return 0;
}

/* ---------------------------------------------------------------- path AD: lune lists (AntipodalOptimized.cpp:39-109) */

static double proj_dist2(const pts_t* P, int s, int t, int p)
{   /* CustomMath.cpp:67-87 */
    if (p == s) return 0.0;
    const double d2 = dist2(P, s, t);
    if (p == t) return d2;
    const double dot = (P->x[t] - P->x[s]) * (P->x[p] - P->x[s]) + (P->y[t] - P->y[s]) * (P->y[p] - P->y[s]);
    double r = (dot * dot) / d2;
    if (r < 0.0) r = 0.0;
    if (r > d2) r = d2;
    return r;
}

typedef struct { const pts_t* P; int s, t; } proj_ctx;
static int proj_asc(const void* v, int32_t a, int32_t b)
{
    const proj_ctx* c = (const proj_ctx*) v;
    return proj_dist2(c->P, c->s, c->t, a) < proj_dist2(c->P, c->s, c->t, b);
}
static int proj_desc(const void* v, int32_t a, int32_t b)
{
    const proj_ctx* c = (const proj_ctx*) v;
    return proj_dist2(c->P, c->s, c->t, a) > proj_dist2(c->P, c->s, c->t, b);
}

static void lune_lists(const pts_t* P, int s, int t, int32_t* L, uint32_t* nl, int32_t* R, uint32_t* nr)
{
    const double d2 = dist2(P, s, t);
    const double stx = P->x[t] - P->x[s], sty = P->y[t] - P->y[s];
    uint32_t a = 0, b = 0;
    for (uint32_t p = 0; p < P->n; ++p)
    {   /* :48-70 */
        if ((int) p == s || (int) p == t) continue;
        if (!(dist2(P, s, (int) p) <= d2 && dist2(P, t, (int) p) <= d2)) continue;
        const double spx = P->x[p] - P->x[s], spy = P->y[p] - P->y[s];
        const double tpx = P->x[p] - P->x[t], tpy = P->y[p] - P->y[t];
        if (!(stx * spx + sty * spy >= 0 && (-stx) * tpx + (-sty) * tpy >= 0)) continue;
        if (orient(P, s, t, (int) p) <= 0) R[b++] = (int32_t) p; else L[a++] = (int32_t) p;
    }
    L[a++] = s; L[a++] = t; R[b++] = s; R[b++] = t;   /* :102-105 */
    proj_ctx ctx = { P, s, t };
    msort(L, a, proj_asc, &ctx);                      /* :107 */
    msort(R, b, proj_desc, &ctx);                     /* :108 */
    *nl = a; *nr = b;
}

int adko_lune_lists(const double* x, const double* y, uint32_t n, uint32_t s, uint32_t t,
                    int32_t* left, uint32_t* nleft, int32_t* right, uint32_t* nright)
{
    pts_t P = { x, y, n };
    lune_lists(&P, (int) s, (int) t, left, nleft, right, nright);
    return 0;
}

/* ---------------------------------------------------------------- path AD: layered state store (AntipodalOptimized.cpp:3,111-127) */

enum { SIDE_LL = 0, SIDE_LR = 1, SIDE_RR = 2, SIDE_RL = 3 };   /* :11-17 */

typedef struct { uint64_t key; uint32_t o, s, ss, prev; double area; int side; } entry_t;

typedef struct
{
    entry_t* e; size_t cnt, cap;
    uint32_t* slot; size_t hcap;   /* open addressing: slot holds entry index + 1 */
} layer_t;

static inline uint64_t mix64(uint64_t k)
{
    k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33;
    return k;
}

static void layer_rehash(layer_t* Ly, size_t hcap)
{
    free(Ly->slot);
    Ly->slot = (uint32_t*) calloc(hcap, sizeof(uint32_t));
    Ly->hcap = hcap;
    for (size_t t = 0; t < Ly->cnt; ++t)
    {
        size_t h = mix64(Ly->e[t].key) & (hcap - 1);
        while (Ly->slot[h]) h = (h + 1) & (hcap - 1);
        Ly->slot[h] = (uint32_t) t + 1;
    }
}

static entry_t* layer_find(layer_t* Ly, uint64_t key)
{
    if (!Ly->hcap) return NULL;
    size_t h = mix64(key) & (Ly->hcap - 1);
    while (Ly->slot[h])
    {
        entry_t* e = &Ly->e[Ly->slot[h] - 1];
        if (e->key == key) return e;
        h = (h + 1) & (Ly->hcap - 1);
    }
    return NULL;
}

static void layer_insert_new(layer_t* Ly, const entry_t* ne)
{
    if (Ly->cnt == Ly->cap)
    {
        Ly->cap = Ly->cap ? Ly->cap * 2 : 64;
        Ly->e = (entry_t*) realloc(Ly->e, sizeof(entry_t) * Ly->cap);
    }
    Ly->e[Ly->cnt++] = *ne;
    if (Ly->cnt * 2 > Ly->hcap) layer_rehash(Ly, Ly->hcap ? Ly->hcap * 2 : 128);
    else
    {
        size_t h = mix64(ne->key) & (Ly->hcap - 1);
        while (Ly->slot[h]) h = (h + 1) & (Ly->hcap - 1);
        Ly->slot[h] = (uint32_t) Ly->cnt;
    }
}

/* "insert if absent, else overwrite area/prev/side only if strictly smaller" (:207-217 and its three clones) */
#ifdef ADK_ORACLE_DEVSTATS
static uint64_t adk_dev_relax;   /* development statistic (tools/ad_cell_stats.c): transitions that pass every test */
#define DEVSTAT_RELAX() (++adk_dev_relax)
#else
#define DEVSTAT_RELAX() ((void) 0)
#endif
static void layer_relax(layer_t* Ly, uint64_t key, uint32_t o, uint32_t s, uint32_t ss, uint32_t prev, double area, int side)
{
    DEVSTAT_RELAX();
    entry_t* e = layer_find(Ly, key);
    if (!e)
    {
        entry_t ne = { key, o, s, ss, prev, area, side };
        layer_insert_new(Ly, &ne);
    }
    else if (area < e->area) { e->area = area; e->prev = prev; e->side = side; }
}

static size_t layers_clear(layer_t* Ls, size_t nlayers)
{   /* locClearCache, :118-127 */
    size_t total = 0;
    for (size_t k = 0; k < nlayers; ++k)
    {
        total += Ls[k].cnt;
        if (Ls[k].cnt) { Ls[k].cnt = 0; if (Ls[k].hcap) memset(Ls[k].slot, 0, sizeof(uint32_t) * Ls[k].hcap); }
    }
    return total;
}

static inline int tri_count_AD(const pts_t* P, const tables_t* T, size_t k, int p, int pp, int q)
{   /* locPointsWithinTriangleCount, AntipodalOptimized.cpp:21-37 */
    int c = COLL(T, p, pp) + COLL(T, pp, q) + pit(P, T, p, pp, q);
    if (k == 0) c += COLL(T, p, q);
    return c;
}

#define KEY3(a, b, c, n) ((uint64_t) (a) * (n) * (n) + (uint64_t) (b) * (n) + (uint64_t) (c))   /* :3 */

/* locProcessSegment (:129-333).  Returns 1 and (k, area) if the pair admits a polygon. */
static int process_segment(const pts_t* P, const tables_t* T, const int32_t* L, uint32_t nl, const int32_t* R, uint32_t nr,
                           uint64_t kmax, double amax, layer_t* Ls, size_t nlayers, uint64_t* cells,
                           uint64_t* out_k, double* out_area)
{
    const uint64_t n = P->n;
    uint64_t lim = (uint64_t) nl + nr - 4;
    if (kmax - 2 < lim) lim = kmax - 2;         /* :137 */
    if (lim == 0) return 0;
    const int S = L[0], E = L[nl - 1];          /* start / end point of the candidate diameter */
    const double seg2 = dist2(P, S, E);
    int have = 0; uint64_t bk = 0; double ba = INFINITY;

    for (uint32_t i = 1; i < nl; ++i) layer_relax(&Ls[0], KEY3(S, E, L[i], n), 0, 0, i, UINT32_MAX, 0.0, SIDE_LL);   /* :149-154 */
    for (uint32_t i = 1; i < nr; ++i) layer_relax(&Ls[0], KEY3(E, S, R[i], n), 0, 0, i, UINT32_MAX, 0.0, SIDE_RR);   /* :155-160 */

    for (uint64_t k = 0; k <= lim && k < nlayers; ++k)
    {
        for (size_t t = 0; t < Ls[k].cnt; ++t)
        {
            const entry_t en = Ls[k].e[t];   /* copy: deeper layers may realloc, this one cannot */
            if ((en.side >= SIDE_RR && R[en.ss] == S) || (en.side <= SIDE_LR && L[en.ss] == E))
            {   /* :167-175 */
                if (k > 0 && (!have || k > bk || (bk == k && en.area < ba))) { have = 1; bk = k; ba = en.area; }
                continue;
            }
            if (en.side >= SIDE_RR)
            {   /* pending edge r -> pr on the right chain, l = current left vertex (:177-249) */
                const int l = L[en.o], r = R[en.s], pr = R[en.ss];
                if (dist2(P, l, pr) > seg2) continue;
                const double tri = fabs(sarea(P, S, r, pr));
                const uint64_t c = 1 + (uint64_t) tri_count_AD(P, T, k, r, pr, S);
                const double area = en.area + tri;
                if (c + k >= nlayers) continue;
                layer_t* dst = &Ls[k + c];
                for (uint32_t i = en.ss + 1; i < nr; ++i)
                {
                    ++*cells;
                    if (orient(P, R[i], pr, r) >= 0 && orient(P, R[i], pr, S) >= 0 && area <= amax)
                        layer_relax(dst, KEY3(l, pr, R[i], n), en.o, en.ss, i, en.s, area, SIDE_RR);
                }
                const double refd = pseudo_dist(P, r, pr, l);
                for (uint32_t i = en.o + 1; i < nl; ++i)
                {
                    ++*cells;
                    const double d = pseudo_dist(P, r, pr, L[i]);
                    if (d <= refd && (l == S || (orient(P, l, E, L[i]) >= 0 && orient(P, L[i], l, S) >= 0)) && area <= amax)
                        layer_relax(dst, KEY3(pr, l, L[i], n), en.ss, en.o, i, en.s, area, SIDE_LR);
                }
            }
            else
            {   /* mirror image (:250-325) */
                const int r = R[en.o], l = L[en.s], pl = L[en.ss];
                if (dist2(P, r, pl) > seg2) continue;
                const double tri = fabs(sarea(P, E, l, pl));
                const uint64_t c = 1 + (uint64_t) tri_count_AD(P, T, k, l, pl, E);
                const double area = en.area + tri;
                if (c + k >= nlayers) continue;
                layer_t* dst = &Ls[k + c];
                for (uint32_t i = en.ss + 1; i < nl; ++i)
                {
                    ++*cells;
                    if (orient(P, L[i], pl, l) >= 0 && orient(P, L[i], pl, E) >= 0 && area <= amax)
                        layer_relax(dst, KEY3(r, pl, L[i], n), en.o, en.ss, i, en.s, area, SIDE_LL);
                }
                const double refd = pseudo_dist(P, l, pl, r);
                for (uint32_t i = en.o + 1; i < nr; ++i)
                {
                    ++*cells;
                    const double d = pseudo_dist(P, l, pl, R[i]);
                    if (d <= refd && (r == E || (orient(P, r, S, R[i]) >= 0 && orient(P, R[i], r, E) >= 0)) && area <= amax)
                        layer_relax(dst, KEY3(pl, r, R[i], n), en.ss, en.o, i, en.s, area, SIDE_RL);
                }
            }
        }
    }
    if (have) { *out_k = bk + 2; *out_area = ba; }   /* :328-331 */
    return have;
}

/* locReconstructHull (:335-436) */
static uint32_t reconstruct_AD(const pts_t* P, const tables_t* T, const int32_t* L, uint32_t nl, const int32_t* R, uint32_t nr,
                               layer_t* Ls, uint64_t kopt, uint32_t* hull)
{
    const uint64_t n = P->n;
    const int S = L[0], E = L[nl - 1];
    uint64_t k = kopt;
    (void) nr;
    entry_t en; memset(&en, 0, sizeof(en)); en.area = INFINITY;
    for (size_t t = 0; t < Ls[k].cnt; ++t)
    {
        const entry_t* v = &Ls[k].e[t];
        if ((v->side >= SIDE_RR && R[v->ss] == S) || (v->side <= SIDE_LR && L[v->ss] == E))
            if (v->area < en.area) en = *v;
    }
    int32_t* left = (int32_t*) malloc(sizeof(int32_t) * (n + 4));
    int32_t* right = (int32_t*) malloc(sizeof(int32_t) * (n + 4));
    uint32_t a = 0, b = 0;
    if (en.side >= SIDE_RR) left[a++] = E; else right[b++] = S;
    while (k > 0 && a + b < n + 2)     /* the bound only matters on exactly degenerate inputs (a step that consumes no point) */
    {
        uint64_t key = 0; int done = 0;
        switch (en.side)
        {
            case SIDE_LL:
                left[a++] = L[en.ss];
                k -= 1 + (uint64_t) pit(P, T, E, L[en.prev], L[en.s]);
                if (k == 0) { left[a++] = L[en.s]; done = 1; break; }
                key = KEY3(R[en.o], L[en.prev], L[en.s], n);
                break;
            case SIDE_LR:
                left[a++] = L[en.ss];
                k -= 1 + (uint64_t) pit(P, T, S, R[en.prev], R[en.o]);
                if (k == 0) { right[b++] = R[en.o]; done = 1; break; }
                key = KEY3(L[en.s], R[en.prev], R[en.o], n);
                break;
            case SIDE_RR:
                right[b++] = R[en.ss];
                k -= 1 + (uint64_t) pit(P, T, S, R[en.prev], R[en.s]);
                if (k == 0) { right[b++] = R[en.s]; done = 1; break; }
                key = KEY3(L[en.o], R[en.prev], R[en.s], n);
                break;
            default:
                right[b++] = R[en.ss];
                k -= 1 + (uint64_t) pit(P, T, E, L[en.prev], L[en.o]);
                if (k == 0) { left[a++] = L[en.o]; done = 1; break; }
                key = KEY3(R[en.s], L[en.prev], L[en.o], n);
                break;
        }
        if (done) break;
        entry_t* nx = layer_find(&Ls[k], key);
        if (!nx) break;   /* cannot happen on tables this DP produced */
        en = *nx;
    }
    for (uint32_t t = 0; t < a; ++t) hull[t] = (uint32_t) left[t];
    for (uint32_t t = 0; t < b; ++t) hull[a + t] = (uint32_t) right[t];
    free(left); free(right);
    return a + b;
}

/* The candidate pairs of one solve are independent (AntipodalOptimized.cpp:495-527 is a loop over them with a running best), so a
 * long solve can be cut into shards: shard s of nshards takes every nshards-th pair of the enumeration order.  The shards' bests
 * merge by (count desc, area asc, enumeration order asc) — the strict comparisons of :512-517 keep the first of equals.  Used by
 * tests/golden/make_oracle_sharded.py to afford instances one core cannot (the sigma = 0.5 bench set); shard 0 of 1 is the solve. */
static int ad_solve_impl(const double* x, const double* y, uint32_t n, uint64_t kmax, double amax, double dmax,
                         int reconstruct, uint32_t shard, uint32_t nshards, adko_result* out, uint32_t* hull);

int adko_solve_area_diameter(const double* x, const double* y, uint32_t n, uint64_t kmax, double amax, double dmax,
                             int reconstruct, adko_result* out, uint32_t* hull)
{
    return ad_solve_impl(x, y, n, kmax, amax, dmax, reconstruct, 0, 1, out, hull);
}

int adko_solve_area_diameter_shard(const double* x, const double* y, uint32_t n, uint64_t kmax, double amax, double dmax,
                                   uint32_t shard, uint32_t nshards, adko_result* out)
{
    if (nshards == 0 || shard >= nshards) return 1;
    return ad_solve_impl(x, y, n, kmax, amax, dmax, 0, shard, nshards, out, NULL);
}

static int ad_solve_impl(const double* x, const double* y, uint32_t n, uint64_t kmax, double amax, double dmax,
                         int reconstruct, uint32_t shard, uint32_t nshards, adko_result* out, uint32_t* hull)
{
    memset(out, 0, sizeof(*out));
    out->entries_alloc = -1; out->entries_required = -1;
    const uint64_t cap = kmax < n ? kmax : n;
    if (cap < 3) return 0;                         /* :461-464 */
    const double t0 = now_ms();
    pts_t P = { x, y, n };
    int32_t* cw = (int32_t*) malloc(sizeof(int32_t) * (size_t) n * (n - 1));
    build_cw(&P, cw, NULL);                        /* :469-479 */
    tables_t T = alloc_tables(n);
    sweep_tables(&P, cw, &T);                      /* :480 */
    free(cw);
    const size_t nlayers = (size_t) cap - 1;       /* :483 */
    layer_t* Ls = (layer_t*) calloc(nlayers, sizeof(layer_t));
    const double dmax2 = dmax * dmax;              /* :486 */
    int32_t* L = (int32_t*) malloc(sizeof(int32_t) * (n + 2));
    int32_t* R = (int32_t*) malloc(sizeof(int32_t) * (n + 2));
    out->entries_alloc = 0; out->entries_required = 0;
    int have = 0; uint64_t bk = 0; double ba = INFINITY; int bs = -1, bt = -1;

    for (uint32_t i = 0; i < n; ++i)
        for (uint32_t j = i + 1; j < n; ++j)
        {   /* :495-527 */
            if (!(dist2(&P, (int) i, (int) j) <= dmax2)) continue;
            if ((out->pairs++ % nshards) != shard) continue;
            uint32_t nl, nr;
            lune_lists(&P, (int) i, (int) j, L, &nl, R, &nr);
            uint64_t pk = 0; double pa = INFINITY;
            const int ok = process_segment(&P, &T, L, nl, R, nr, kmax, amax, Ls, nlayers, &out->cells, &pk, &pa);
            if (ok && pk <= kmax && (!have || pk > bk || (pk == bk && pa < ba)))
            {
                have = 1; bk = pk; ba = pa; bs = (int) i; bt = (int) j;
            }
            const size_t cnt = layers_clear(Ls, nlayers);
            out->states += cnt;
            out->entries_alloc += (int64_t) cnt;
            out->entries_required += (int64_t) cnt;
        }

    if (have)
    {
        out->found = 1;
        out->count = (uint32_t) bk;
        out->area = ba;
        out->diam_valid = 1;
        out->diam[0] = (uint32_t) bs; out->diam[1] = (uint32_t) bt;   /* :516 */
        if (reconstruct)
        {   /* :529-541 */
            uint32_t nl, nr; uint64_t pk; double pa; uint64_t dummy = 0;
            lune_lists(&P, bs, bt, L, &nl, R, &nr);
            process_segment(&P, &T, L, nl, R, nr, kmax, amax, Ls, nlayers, &dummy, &pk, &pa);
            out->hull_len = reconstruct_AD(&P, &T, L, nl, R, nr, Ls, bk - 2, hull);
            const size_t cnt = layers_clear(Ls, nlayers);
            out->entries_alloc += (int64_t) cnt;
            out->entries_required += (int64_t) cnt;
        }
    }
    for (size_t k = 0; k < nlayers; ++k) { free(Ls[k].e); free(Ls[k].slot); }
    free(Ls); free(L); free(R); free_tables(&T);
    out->solve_ms = now_ms() - t0;
    return 0;
}
