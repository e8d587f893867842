/* Development tool (not product, not test): per-cell statistics of the AD layered DP, to size the cell-wavefront kernel.
 * Includes the oracle restatement and walks the layer tables of every stride-th candidate pair.
 * Build: gcc -O2 -o /tmp/ad_cell_stats tools/ad_cell_stats.c -lm ; run: /tmp/ad_cell_stats points.txt amax dmax stride */
#define ADK_ORACLE_DEVSTATS 1
#include "../oracle/adk_oracle.c"
#include <stdio.h>

static uint64_t hist_ent[16], hist_edges[16], hist_band[16], hist_len[16], hist_dests[16];
static uint64_t w_ent[16], w_band[16], w_edges[16], w_tot;
static int bucket(uint64_t v) { int b = 0; while (v > 1 && b < 15) { v >>= 1; ++b; } return b; }

int main(int argc, char** argv)
{
    FILE* f = fopen(argv[1], "r");
    const double amax = atof(argv[2]), dmax = atof(argv[3]);
    const int stride = atoi(argv[4]);
    static double x[4096], y[4096];
    uint32_t n = 0;
    while (fscanf(f, "%lf %lf", &x[n], &y[n]) == 2) ++n;
    pts_t P = { x, y, n };
    int32_t* cw = (int32_t*) malloc(sizeof(int32_t) * (size_t) n * (n - 1));
    build_cw(&P, cw, NULL);
    tables_t T = alloc_tables(n);
    sweep_tables(&P, cw, &T);
    const size_t nlayers = n - 1;
    layer_t* Ls = (layer_t*) calloc(nlayers, sizeof(layer_t));
    int32_t* L = malloc(4 * (n + 2)); int32_t* R = malloc(4 * (n + 2));
    uint64_t pairs = 0, used = 0, tot_states = 0, tot_cells_alive = 0, tot_cells = 0, tot_edges = 0, cells = 0;
    uint64_t max_states = 0, sum_band_x_dests = 0;
    for (uint32_t i = 0; i < n; ++i) for (uint32_t j = i + 1; j < n; ++j)
    {
        if (!(dist2(&P, i, j) <= dmax * dmax)) continue;
        if ((pairs++ % stride) != 0) continue;
        uint32_t nl, nr; lune_lists(&P, i, j, L, &nl, R, &nr);
        uint64_t pk = 0; double pa = 0;
        process_segment(&P, &T, L, nl, R, nr, UINT64_MAX, amax, Ls, nlayers, &cells, &pk, &pa);
        ++used;
        /* per (cell, side, edge): count, kmin, kmax */
        const size_t NC = (size_t) nl * nr, NE = NC * (nl + nr);
        uint16_t* ecnt = calloc(NE, 2); uint16_t* kmin = malloc(NC * 2); uint16_t* kmax = calloc(NC, 2); uint32_t* cent = calloc(NC, 4);
        memset(kmin, 0xff, NC * 2);
        uint64_t states = 0;
        for (size_t k = 0; k < nlayers; ++k) for (size_t t = 0; t < Ls[k].cnt; ++t)
        {
            const entry_t* e = &Ls[k].e[t];
            uint32_t a, b, edge;
            if (e->side >= SIDE_RR) { a = e->o; b = e->ss; edge = e->s; } else { a = e->ss; b = e->o; edge = nr + e->s; }
            const size_t c = (size_t) a * nr + b;
            ++ecnt[c * (nl + nr) + edge]; ++cent[c]; ++states;
            if (k < kmin[c]) kmin[c] = k; if (k > kmax[c]) kmax[c] = k;
        }
        tot_states += states; if (states > max_states) max_states = states;
        tot_cells += NC;
        for (size_t c = 0; c < NC; ++c) if (cent[c])
        {
            ++tot_cells_alive;
            uint32_t edges = 0;
            for (uint32_t e = 0; e < nl + nr; ++e) if (ecnt[c * (nl + nr) + e]) { ++edges; ++hist_len[bucket(ecnt[c * (nl + nr) + e])]; }
            tot_edges += edges;
            const uint32_t a = c / nr, b = c % nr, D = (nr - 1 - b) + (nl - 1 - a);
            ++hist_ent[bucket(cent[c])]; ++hist_edges[bucket(edges)]; ++hist_band[bucket(kmax[c] - kmin[c] + 1)]; ++hist_dests[bucket(D ? D : 1)];
            sum_band_x_dests += (uint64_t) (kmax[c] - kmin[c] + 1) * D;
            { const uint64_t w = (uint64_t) cent[c] * D; w_tot += w; w_ent[bucket(cent[c])] += w; w_band[bucket(kmax[c] - kmin[c] + 1)] += w; w_edges[bucket(edges)] += w; }
        }
        free(ecnt); free(kmin); free(kmax); free(cent);
        layers_clear(Ls, nlayers);
    }
    printf("n %u pairs %lu sampled %lu states %lu (avg %.0f max %lu) cells(ref) %lu\n", n, pairs, used, tot_states, (double) tot_states / used, max_states, cells);
    printf("cells alive %lu of %lu (%.1f%%), entries/alive cell %.1f, edges/alive cell %.1f, band*dests per alive cell %.0f\n", tot_cells_alive, tot_cells,
           100.0 * tot_cells_alive / tot_cells, (double) tot_states / tot_cells_alive, (double) tot_edges / tot_cells_alive, (double) sum_band_x_dests / tot_cells_alive);
    printf("transitions passing every test %lu = %.1f%% of the cell updates\n", adk_dev_relax, 100.0 * adk_dev_relax / cells);
    const char* names[] = { "entries/cell", "edges/cell", "band/cell", "len/(cell,edge)", "dests/cell" };
    uint64_t* hs[] = { hist_ent, hist_edges, hist_band, hist_len, hist_dests };
    for (int h = 0; h < 5; ++h) { printf("%-16s", names[h]); for (int b = 0; b < 12; ++b) printf(" [%d]%lu", 1 << b, hs[h][b]); printf("\n"); }
    const char* wn[] = { "W entries/cell", "W band/cell", "W edges/cell" };
    uint64_t* ws[] = { w_ent, w_band, w_edges };
    for (int h = 0; h < 3; ++h) { printf("%-16s", wn[h]); for (int b = 0; b < 12; ++b) printf(" [%d]%.1f%%", 1 << b, 100.0 * ws[h][b] / w_tot); printf("\n"); }
    return 0;
}
