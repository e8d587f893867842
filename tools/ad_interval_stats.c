/* Development tool (not product, not test): how often is the set of source edges that pass the transition tests of one
 * destination an INTERVAL of the sources sorted by angle around the chain head?  (Sizes the sorted-sweep accumulate.)
 * Build: gcc -O2 -o /tmp/ad_interval_stats tools/ad_interval_stats.c -lm ; run: /tmp/ad_interval_stats points.txt amax dmax stride */
#define ADK_ORACLE_DEVSTATS 1
#include "../oracle/adk_oracle.c"
#include <stdio.h>

enum { K_EMPTY, K_FULL, K_PREFIX, K_SUFFIX, K_MIDDLE, K_ENDS, K_OTHER, K_N };
static const char* kname[K_N] = { "empty", "full", "prefix", "suffix", "middle", "prefix+suffix", "other" };
static uint64_t cls[4][K_N];      /* [R stay, R switch, L stay, L switch] */
static uint64_t passes[4], evals[4], runs_hist[4][8];

static const pts_t* gP; static int gH;
static double gang[4096];
static double gfx, gfy;   /* unit-less frame axis: direction s_pt -> t_pt */
static double pseudo_angle(double hx, double hy, double px, double py, double sgn)
{
    const double dx = px - hx, dy = py - hy;
    const double u = sgn * (dx * gfx + dy * gfy), v = sgn * (gfx * dy - gfy * dx);
    const double den = fabs(u) + fabs(v);
    if (den == 0.0) return 0.0;
    const double r = v / den;
    return u >= 0.0 ? r : (v >= 0.0 ? 2.0 - r : -2.0 - r);
}
static int ang_less(const void* a, const void* b)
{
    const double x = gang[*(const int*) a], y = gang[*(const int*) b];
    return x < y ? -1 : x > y ? 1 : (*(const int*) a - *(const int*) b);
}

static int g_lastlo; static uint64_t g_ne, g_nd, g_dd;
static void classify(int kind, const unsigned char* m, int P)
{
    { int o = 0; for (int p = 0; p < P; ++p) o += m[p]; g_lastlo = P - o; }
    int runs = 0, ones = 0;
    for (int p = 0; p < P; ++p) { ones += m[p]; if (m[p] && (p == 0 || !m[p - 1])) ++runs; }
    passes[kind] += ones; evals[kind] += P;
    ++runs_hist[kind][runs > 7 ? 7 : runs];
    int c;
    if (ones == 0) c = K_EMPTY; else if (ones == P) c = K_FULL;
    else if (runs == 1 && m[0]) c = K_PREFIX; else if (runs == 1 && m[P - 1]) c = K_SUFFIX; else if (runs == 1) c = K_MIDDLE;
    else if (runs == 2 && m[0] && m[P - 1]) c = K_ENDS; else c = K_OTHER;
    ++cls[kind][c];
}

int main(int argc, char** argv)
{
    FILE* f = fopen(argv[1], "r");
    const double amax = atof(argv[2]), dmax = atof(argv[3]);
    const int stride = atoi(argv[4]);
    static double x[4096], y[4096];
    uint32_t n = 0;
    while (fscanf(f, "%lf %lf", &x[n], &y[n]) == 2) ++n;
    pts_t P = { x, y, n }; gP = &P;
    int32_t* cw = (int32_t*) malloc(sizeof(int32_t) * (size_t) n * (n - 1));
    build_cw(&P, cw, NULL);
    tables_t T = alloc_tables(n);
    sweep_tables(&P, cw, &T);
    const size_t nlayers = n - 1;
    layer_t* Ls = (layer_t*) calloc(nlayers, sizeof(layer_t));
    int32_t* L = malloc(4 * (n + 2)); int32_t* R = malloc(4 * (n + 2));
    uint64_t pairs = 0, used = 0, cells = 0;
    for (uint32_t i = 0; i < n; ++i) for (uint32_t j = i + 1; j < n; ++j)
    {
        if (!(dist2(&P, i, j) <= dmax * dmax)) continue;
        if ((pairs++ % stride) != 0) continue;
        uint32_t nl, nr; lune_lists(&P, i, j, L, &nl, R, &nr);
        uint64_t pk = 0; double pa = 0;
        process_segment(&P, &T, L, nl, R, nr, UINT64_MAX, amax, Ls, nlayers, &cells, &pk, &pa);
        ++used;
        const int S = L[0], E = L[nl - 1];
        const double seg2 = dist2(&P, S, E);
        gfx = x[E] - x[S]; gfy = y[E] - y[S];
        const size_t NC = (size_t) nl * nr, W = nl + nr;
        unsigned char* alive = calloc(NC * W, 1);
        for (size_t k = 0; k < nlayers; ++k) for (size_t t = 0; t < Ls[k].cnt; ++t)
        {
            const entry_t* e = &Ls[k].e[t];
            uint32_t a, b, edge;
            if (e->side >= SIDE_RR) { a = e->o; b = e->ss; edge = e->s; } else { a = e->ss; b = e->o; edge = nr + e->s; }
            alive[((size_t) a * nr + b) * W + edge] = 1;
        }
        static int srcs[4096]; static unsigned char m[4096]; static int loR[8192], loL[8192];
        for (uint32_t a = 0; a < nl; ++a) for (uint32_t b = 0; b < nr; ++b)
        {
            const unsigned char* al = alive + ((size_t) a * nr + b) * W;
            const int l = L[a], r = R[b];
            if (dist2(&P, l, r) > seg2) continue;
            if (b == nr - 1 || a == nl - 1) { /* terminal on one side: those sources do not expand */ }
            /* R sources: pending edge R[s]->R[b] */
            const int nDR = nr - 1 - b, nDL = nl - 1 - a, DD = nDR + nDL;
            for (int d = 0; d < DD; ++d) { loR[d] = -1; loL[d] = -1; }
            int PR = 0;
            if (b != nr - 1) for (uint32_t s = 0; s < b; ++s) if (al[s]) { srcs[PR++] = s; gang[s] = pseudo_angle(x[r], y[r], x[R[s]], y[R[s]], 1.0); }
            if (PR > 1)
            {
                qsort(srcs, PR, sizeof(int), ang_less);
                for (uint32_t i2 = b + 1; i2 < nr; ++i2)
                {   /* stay right */
                    const int ok2 = orient(&P, R[i2], r, S) >= 0;
                    for (int p = 0; p < PR; ++p) m[p] = ok2 && orient(&P, R[i2], r, R[srcs[p]]) >= 0;
                    classify(0, m, PR); loR[i2 - b - 1] = g_lastlo == PR ? -1 : g_lastlo;
                }
                for (uint32_t i2 = a + 1; i2 < nl; ++i2)
                {   /* switch left */
                    const int ok2 = (l == S) || (orient(&P, l, E, L[i2]) >= 0 && orient(&P, L[i2], l, S) >= 0);
                    for (int p = 0; p < PR; ++p) m[p] = ok2 && pseudo_dist(&P, R[srcs[p]], r, L[i2]) <= pseudo_dist(&P, R[srcs[p]], r, l);
                    classify(1, m, PR); loR[nDR + i2 - a - 1] = g_lastlo == PR ? -1 : g_lastlo;
                }
            }
            int PL = 0;
            if (a != nl - 1) for (uint32_t s = 0; s < a; ++s) if (al[nr + s]) { srcs[PL++] = s; gang[s] = pseudo_angle(x[l], y[l], x[L[s]], y[L[s]], -1.0); }
            if (PL > 1)
            {
                qsort(srcs, PL, sizeof(int), ang_less);
                for (uint32_t i2 = a + 1; i2 < nl; ++i2)
                {
                    const int ok2 = orient(&P, L[i2], l, E) >= 0;
                    for (int p = 0; p < PL; ++p) m[p] = ok2 && orient(&P, L[i2], l, L[srcs[p]]) >= 0;
                    classify(2, m, PL); loL[nDR + i2 - a - 1] = g_lastlo == PL ? -1 : g_lastlo;
                }
                for (uint32_t i2 = b + 1; i2 < nr; ++i2)
                {
                    const int ok2 = (r == E) || (orient(&P, r, S, R[i2]) >= 0 && orient(&P, R[i2], r, E) >= 0);
                    for (int p = 0; p < PL; ++p) m[p] = ok2 && pseudo_dist(&P, L[srcs[p]], l, R[i2]) <= pseudo_dist(&P, L[srcs[p]], l, r);
                    classify(3, m, PL); loL[i2 - b - 1] = g_lastlo == PL ? -1 : g_lastlo;
                }
            }
            {   /* distinct (loR, loL) among the destinations that receive anything */
                static unsigned char seen[512 * 512 / 8 + 8];
                int ne = 0, nd = 0;
                for (int d = 0; d < DD; ++d) if (loR[d] >= 0 || loL[d] >= 0)
                {
                    ++ne;
                    const int key = (loR[d] + 1) * 300 + (loL[d] + 1);
                    if (!(seen[key >> 3] & (1 << (key & 7)))) { seen[key >> 3] |= 1 << (key & 7); ++nd; }
                }
                for (int d = 0; d < DD; ++d) if (loR[d] >= 0 || loL[d] >= 0) { const int key = (loR[d] + 1) * 300 + (loL[d] + 1); seen[key >> 3] = 0; }
                g_ne += ne; g_nd += nd; g_dd += DD;
            }
        }
        free(alive);
        layers_clear(Ls, nlayers);
    }
    printf("n %u pairs %lu sampled %lu\n", n, pairs, used);
    const char* kn[4] = { "R stay", "R switch", "L stay", "L switch" };
    for (int k = 0; k < 4; ++k)
    {
        uint64_t tot = 0; for (int c = 0; c < K_N; ++c) tot += cls[k][c];
        printf("%-9s pass %.1f%% of %lu evals; (dest,group) classes:", kn[k], 100.0 * passes[k] / (evals[k] ? evals[k] : 1), evals[k]);
        for (int c = 0; c < K_N; ++c) printf(" %s %.2f%%", kname[c], 100.0 * cls[k][c] / (tot ? tot : 1));
        printf("\n          runs:"); for (int r = 0; r < 8; ++r) printf(" [%d]%lu", r, runs_hist[k][r]); printf("\n");
    }
    printf("destinations %lu, non-empty %lu, distinct (loR,loL) %lu\n", g_dd, g_ne, g_nd);
    return 0;
}
