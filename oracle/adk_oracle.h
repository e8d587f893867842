/* TEST INFRASTRUCTURE — CPU restatement (plain C, IEEE binary64) of the reference's two hot paths.
 *
 * This is the parity checker for the CUDA solver.  It is NOT part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may call it.
 * Each function in adk_oracle.c cites the reference lines (relative to /root/reference) it follows.
 * Pinned against: the reference's 100 golden fixtures (data/samples/tests, via tests/golden/) and the
 * live outputs of the unmodified reference build oracle/_ref/libadkref.so (tests/test_oracle_cpu.py).
 */
#ifndef ADK_ORACLE_H
#define ADK_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct adko_result
{
    int32_t found;
    uint32_t count;
    double area;
    uint32_t hull_len;
    uint32_t diam_valid;
    uint32_t diam[2];
    int64_t entries_alloc;     /* BenchmarkInfo::myAllocatedEntriesCount */
    int64_t entries_required;  /* BenchmarkInfo::myRequiredEntriesCount  */
    uint64_t cells;            /* A: iterations of Eppstein.cpp:205-235; AD: iterations of AntipodalOptimized.cpp:198,225,271,298 */
    uint64_t states;           /* AD: hash entries over all candidate pairs */
    uint64_t pairs;            /* AD: candidate pairs within the diameter bound */
    double solve_ms;
} adko_result;

#define ADKO_KMAX_INF UINT64_MAX

/* MT::EppsteinAlgorithmWithBenchmarkInfo (Eppstein.cpp:128-277). hull: caller buffer, capacity n. */
int adko_solve_area(const double* x, const double* y, uint32_t n, uint64_t kmax, double amax,
                    int reconstruct, adko_result* out, uint32_t* hull);

/* The same solve restricted to every nshards-th anchor of the bottom-up anchor order (Eppstein.cpp:177): out is this shard's best
 * (with its hull when reconstruct is set), cells / entries_required this shard's share, *best_anchor_pos the position of the best's
 * anchor in that order.  Merge by (count desc, area asc, best_anchor_pos asc).  tests/golden/make_oracle_sharded.py --area. */
int adko_solve_area_shard(const double* x, const double* y, uint32_t n, uint64_t kmax, double amax, int reconstruct,
                          uint32_t shard, uint32_t nshards, adko_result* out, uint32_t* hull, uint32_t* best_anchor_pos);

/* MT::AntipodalOptimizedAlgorithmWithBenchmarkInfo (AntipodalOptimized.cpp:452-545). */
int adko_solve_area_diameter(const double* x, const double* y, uint32_t n, uint64_t kmax, double amax, double dmax,
                             int reconstruct, adko_result* out, uint32_t* hull);

/* The same solve restricted to every nshards-th candidate pair of the enumeration order (no reconstruction): out->count, area and
 * diam are this shard's best, cells / states this shard's share, pairs the number of ALL candidate pairs.  Merge the shards by
 * (count desc, area asc, (diam[0], diam[1]) asc).  tests/golden/make_oracle_sharded.py. */
int adko_solve_area_diameter_shard(const double* x, const double* y, uint32_t n, uint64_t kmax, double amax, double dmax,
                                   uint32_t shard, uint32_t nshards, adko_result* out);

/* SortPointsClockWiseAroundPoint for every centre (Utils.cpp:295-316): order[i*(n-1)+p]. */
int adko_cw_order(const double* x, const double* y, uint32_t n, int32_t* order);

/* CountPointsBelowAllSegments (Utils.cpp:74-143): the sweep, row-major n*n tables. */
int adko_count_tables(const double* x, const double* y, uint32_t n, int32_t* below, int32_t* coll, int32_t* right_below);

/* The same three tables from the definitions in the reference's disabled self-check (Utils.cpp:145-223). */
int adko_count_tables_bruteforce(const double* x, const double* y, uint32_t n, int32_t* below, int32_t* coll, int32_t* right_below);

/* PointsInTriangle (Utils.cpp:251-293) for m index triples, using the sweep tables. */
int adko_points_in_triangle(const double* x, const double* y, uint32_t n, const uint32_t* tri, uint32_t m, int32_t* out);

/* locGetFirstLeftAndRight (Eppstein.cpp:62-114): slope order around centre j for anchor i.
 * out: point indices, capacity n; returns the list length (or <0 on error). */
int adko_slope_order(const double* x, const double* y, uint32_t n, uint32_t i, uint32_t j, int32_t* out);

/* locPrepareLeftAndRightPoints (AntipodalOptimized.cpp:39-109): L = [s..t], R = [t..s]; capacity n each. */
int adko_lune_lists(const double* x, const double* y, uint32_t n, uint32_t s, uint32_t t,
                    int32_t* left, uint32_t* nleft, int32_t* right, uint32_t* nright);

#ifdef __cplusplus
}
#endif
#endif
