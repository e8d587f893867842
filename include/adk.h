/* adk.h — C ABI of the B200-native (sm_100a) area / area+diameter bounded largest-convex-subset solver.
 *
 * This is the drop-in boundary for the two hot paths of gianmarcopicarella/area-diameter-algorithm.
 * Everything above it (the C++ mirror of namespace MT in area-diameter-algorithm_b200/host/, the reference's own
 * src/bench, src/test, src/main, src/py-binding) stays host code; everything below it is hand-written CUDA.
 * Plain pointers and sizes only: no C++ types, no torch types, no ownership transfer.
 *
 * Reference interfaces replaced (paths relative to the reference repo):
 *   adk_solve_area            <- MT::EppsteinAlgorithmWithBenchmarkInfo           src/common/Eppstein.h:17-23,  Eppstein.cpp:128-277
 *   adk_solve_area_diameter   <- MT::AntipodalOptimizedAlgorithmWithBenchmarkInfo  src/common/AntipodalOptimized.h:18-25, AntipodalOptimized.cpp:452-545
 *                                MT::AntipodalAlgorithmWithBenchmarkInfo           src/common/Antipodal.h:18-25 (served by the same kernels)
 *   adk_*_partial / adk_merge / adk_*_finish : the same two solves split into "my shard of anchors / candidate pairs",
 *                                a host-side lexicographic merge and the witness reconstruction, for one-process-per-GPU sharding.
 *   adk_solve_batch           <- the per-slide loop over thesis.AreaDiameterAlgorithmOptimized (Area-Selector, quality_of_results_comparison.py:48-55)
 *   adk_count_tables          <- MT::CountPointsBelowAllSegments                  src/common/Utils.h:41-43,   Utils.cpp:74-143
 *   adk_points_in_triangle    <- MT::PointsInTriangle                             src/common/Utils.h:45-49,   Utils.cpp:251-293
 *   adk_cw_order              <- MT::SortPointsClockWiseAroundPoint               src/common/Utils.h:53,      Utils.cpp:311-316
 *
 * There is NO CPU fallback behind these entry points: without a CUDA device adk_create fails with ADK_ERR_CUDA.
 */
#ifndef ADK_H
#define ADK_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ADK_VERSION 1

/* status codes */
#define ADK_OK 0
#define ADK_ERR_CUDA 1        /* a CUDA runtime call failed; adk_last_error() has the text */
#define ADK_ERR_ARG 2         /* bad argument (NULL, n too large, shard out of range, ...) */
#define ADK_ERR_NOMEM 3       /* the instance needs more device memory than the budget allows */

/* flags */
#define ADK_FLAG_RECONSTRUCT 1u   /* fill hull[] (and, for the area path, the hull's diameter pair) */
#define ADK_FLAG_COUNTERS 2u      /* fill the BenchmarkInfo-style counters */

#define ADK_KMAX_INF UINT64_MAX   /* aMaxPointsCount = (size_t) -1 */
#define ADK_MAX_POINTS 4096u

typedef struct adk_ctx adk_ctx;

/* One solve's answer: mirrors MT::ConvexArea (Utils.h:20-26) + MT::BenchmarkInfo (Utils.h:35-39). */
typedef struct adk_result
{
    int32_t found;              /* 0 = std::nullopt */
    uint32_t count;             /* myPointsCount */
    double area;                /* myHullArea */
    uint32_t hull_len;          /* number of entries written to hull[] (0 unless ADK_FLAG_RECONSTRUCT) */
    uint32_t diam_valid;        /* myDiameterOpt.has_value() */
    uint32_t diam[2];           /* myFirstIndex, mySecondIndex */
    int64_t entries_alloc;      /* myAllocatedEntriesCount (-1 when not requested) */
    int64_t entries_required;   /* myRequiredEntriesCount  (-1 when not requested) */
    uint64_t cells_nominal;     /* DP cell updates the reference executes for this call */
    uint64_t cells_executed;    /* DP cell updates the kernels executed */
    uint64_t states;            /* area+diameter path: DP states created (reference "entries"; after adk_area_diameter_finish: those of the
                                   winning candidate alone); area path: prefix-minimum change points */
    uint64_t pairs;             /* area+diameter path: candidate pairs within the diameter bound */
    double kernel_ms;           /* device time of the whole call (CUDA events on the solver's stream) */
    double dp_ms;               /* device time of the dominant DP kernel alone */
} adk_result;

/* A rank's local best: what is exchanged between GPUs (32 bytes). */
typedef struct adk_partial
{
    int32_t found;
    uint32_t count;
    double area;
    uint64_t tie_key;           /* iteration-order key: smaller = met earlier by the reference */
    uint32_t w0, w1;            /* witness: area path (anchor, last vertex); area+diameter path (s, t) */
} adk_partial;

/* Run statistics of a partial solve (all device-side counts). */
typedef struct adk_stats
{
    uint64_t cells_nominal, cells_executed, states, pairs;
    double kernel_ms, dp_ms;
    uint64_t launches;          /* kernels launched by the call */
    uint64_t arena_bytes;       /* device workspace in use */
} adk_stats;

/* ---- context ------------------------------------------------------------------------------------------------ */
int adk_create(adk_ctx** out_ctx, int device);        /* device = CUDA ordinal; -1 = current device */
void adk_destroy(adk_ctx* ctx);
const char* adk_last_error(const adk_ctx* ctx);       /* text of the last failure on this context */
int adk_set_memory_budget(adk_ctx* ctx, uint64_t bytes);   /* cap for DP workspaces (default: 75 % of free memory) */
int adk_set_stream(adk_ctx* ctx, void* cuda_stream);       /* run on the caller's cudaStream_t (NULL = legacy default stream) */
uint64_t adk_launch_count(const adk_ctx* ctx);             /* kernels launched by this context since adk_create */
/* By default a context keeps the precompute products (count tables, angular orders) while consecutive solves pass the same
 * coordinates (a diameter sweep).  enabled = 0: every solve recomputes them, as the reference does on every call
 * (Eppstein.cpp:143-155, AntipodalOptimized.cpp:469-481); bench.py times this mode. */
int adk_set_reuse(adk_ctx* ctx, int enabled);

/* ---- whole solves (host buffers in, host buffers out; H2D/D2H inside) ------------------------------------------ */
int adk_solve_area(adk_ctx* ctx, const double* x, const double* y, uint32_t n,
                   uint64_t kmax, double amax, uint32_t flags, adk_result* out, uint32_t* hull /* capacity n */);

int adk_solve_area_diameter(adk_ctx* ctx, const double* x, const double* y, uint32_t n,
                            uint64_t kmax, double amax, double dmax, uint32_t flags, adk_result* out, uint32_t* hull);

/* ---- sharded solves: shard `shard` of `nshards` takes every nshards-th unit of the cost-ordered work list (anchors; candidate
 *      pairs by cost class, then pair key: the same deal on every rank, nothing is exchanged to agree on it).  finish is cheapest
 *      on the rank whose partial won the merge (its records are still resident); elsewhere it re-runs the winning unit. ------- */
int adk_area_partial(adk_ctx* ctx, const double* x, const double* y, uint32_t n, uint64_t kmax, double amax,
                     uint32_t shard, uint32_t nshards, adk_partial* out, adk_stats* stats);
int adk_area_finish(adk_ctx* ctx, const double* x, const double* y, uint32_t n, uint64_t kmax, double amax,
                    const adk_partial* best, uint32_t flags, adk_result* out, uint32_t* hull);

int adk_area_diameter_partial(adk_ctx* ctx, const double* x, const double* y, uint32_t n, uint64_t kmax, double amax,
                              double dmax, uint32_t shard, uint32_t nshards, adk_partial* out, adk_stats* stats);
int adk_area_diameter_finish(adk_ctx* ctx, const double* x, const double* y, uint32_t n, uint64_t kmax, double amax,
                             double dmax, const adk_partial* best, uint32_t flags, adk_result* out, uint32_t* hull);

/* ---- batches of independent instances (BASELINE config 5: the cell samples of one study, one instance per slide;
 *      replaces the Python loop over thesis.AreaDiameterAlgorithmOptimized in
 *      external/Area-Selector/run/quality_of_results_comparison.py:48-55,194-197 and the per-call marshalling of
 *      src/py-binding/main.cpp:10-18) -------------------------------------------------------------------------------- */
#define ADK_ALGO_AREA 0u              /* MT::EppsteinAlgorithm */
#define ADK_ALGO_AREA_DIAMETER 1u     /* MT::AntipodalOptimizedAlgorithm (also serves MT::AntipodalAlgorithm) */

typedef struct adk_instance
{
    const double* x;            /* host SoA coordinates, borrowed for the call */
    const double* y;
    uint32_t n;
    uint32_t algo;              /* ADK_ALGO_* */
    uint64_t kmax;              /* aMaxPointsCount (ADK_KMAX_INF = no bound) */
    double amax, dmax;          /* aMaxArea, aMaxDiameter (dmax ignored by ADK_ALGO_AREA) */
    uint32_t flags;             /* ADK_FLAG_* */
    uint32_t hull_cap;          /* capacity of hull[] (>= n when ADK_FLAG_RECONSTRUCT) */
    uint32_t* hull;             /* caller buffer or NULL */
} adk_instance;

/* Instance-level sharding ("C5 batching"): longest-processing-time-first assignment of `ninst` instances to
 * `nshards` ranks by estimated cost (n^3 for the area path, n^4 for the area+diameter path).  owner[i] = rank of instance i.
 * Deterministic and identical on every rank; host side, no device work. */
int adk_batch_assign(const adk_instance* inst, uint32_t ninst, uint32_t nshards, uint32_t* owner /* ninst */);

/* Solves the instances owned by `shard` (all of them when nshards == 1), heaviest first, on this context's device.
 * out[i].found = -1 marks an instance owned by another shard (its adk_result is otherwise zero).  There is no data-path
 * exchange between ranks: the caller gathers the results it needs (a host gather of adk_result records). */
int adk_solve_batch(adk_ctx* ctx, const adk_instance* inst, uint32_t ninst, uint32_t shard, uint32_t nshards, adk_result* out);

/* ---- the tiled area-only workflow of the reference's Python driver in one call: replaces create_patch_dict
 *      (external/Area-Selector/src/convex_area_selector/convex_polygon_finder.py:139-219) + the window loop of run_area_only_tiled
 *      (run/quality_of_results_comparison.py:77-103).  The windows (patch_size x patch_size, np.arange grid from min - 0.01 with
 *      step_size) are generated, counted and ordered on the device; they are visited by population, largest first, each one solved
 *      with the area path from device-resident coordinates, until a window holds fewer points than the best polygon so far.
 *      hull[] and diam[] index the caller's points (the first of equal points, as the reference maps them back). ----------------- */
typedef struct adk_tiled_stats
{
    uint64_t patches;           /* non-empty windows */
    uint64_t solved;            /* windows visited before the early stop */
} adk_tiled_stats;

int adk_solve_area_tiled(adk_ctx* ctx, const double* x, const double* y, uint32_t n, double patch_size, double step_size,
                         uint64_t kmax, double amax, adk_result* out, uint32_t* hull /* capacity n */, adk_tiled_stats* stats);

/* Lexicographic best of `count` partials: count desc, area asc, tie_key asc (host side, no device work). */
int adk_merge(const adk_partial* parts, uint32_t count, adk_partial* out);

/* ---- several GPUs of one box behind one handle (replaces the same two reference calls, src/bench/main.cpp:89,100 and
 *      src/py-binding/main.cpp:29,59, when more than one device is configured).  One blocking call per solve, as in the reference;
 *      inside it one host thread per device solves that device's shard of the anchors / candidate pairs, ONE ncclAllGather of the
 *      32-byte adk_partial records + the same adk_merge on every rank is the "max-reduce of (count, area, witness)", the rank that
 *      owns the winning unit reconstructs the witness from its resident records and ONE ncclBroadcast hands record + hull to every
 *      rank.  NCCL is bound at run time (libnccl.so.2); without it adk_multi_create fails — there is no fallback. ----------------- */
typedef struct adk_multi adk_multi;

typedef struct adk_multi_stats
{
    uint32_t ndev, pad;
    uint64_t cells_nominal, cells_executed, states, pairs, launches;   /* sums over the devices (pairs: the common candidate count) */
    double max_kernel_ms, max_dp_ms, sum_dp_ms;                        /* device time of the slowest rank; DP kernel time, max and sum */
    double max_exchange_ms;                                            /* all-gather + reconstruction + broadcast, slowest rank */
} adk_multi_stats;

int adk_multi_create(adk_multi** out, const int* devices, int ndev);  /* devices == NULL or ndev <= 0: CUDA devices 0 .. ndev-1 / all */
void adk_multi_destroy(adk_multi* m);
const char* adk_multi_last_error(const adk_multi* m);
int adk_multi_device_count(const adk_multi* m);
int adk_multi_set_reuse(adk_multi* m, int enabled);                    /* adk_set_reuse on every device */
int adk_multi_solve_area(adk_multi* m, const double* x, const double* y, uint32_t n, uint64_t kmax, double amax, uint32_t flags,
                         adk_result* out, uint32_t* hull /* capacity n */);
int adk_multi_solve_area_diameter(adk_multi* m, const double* x, const double* y, uint32_t n, uint64_t kmax, double amax, double dmax,
                                  uint32_t flags, adk_result* out, uint32_t* hull);
int adk_multi_last_stats(const adk_multi* m, adk_multi_stats* out);    /* of the last solve */

/* ---- device-resident variants (pointers are DEVICE pointers; used by bench.py's kernel-only timing) ------------ */
int adk_area_partial_dev(adk_ctx* ctx, const double* d_x, const double* d_y, uint32_t n, uint64_t kmax, double amax,
                         uint32_t shard, uint32_t nshards, adk_partial* out, adk_stats* stats);
int adk_area_diameter_partial_dev(adk_ctx* ctx, const double* d_x, const double* d_y, uint32_t n, uint64_t kmax,
                                  double amax, double dmax, uint32_t shard, uint32_t nshards, adk_partial* out,
                                  adk_stats* stats);

int adk_area_finish_dev(adk_ctx* ctx, const double* d_x, const double* d_y, uint32_t n, uint64_t kmax, double amax,
                        const adk_partial* best, uint32_t flags, adk_result* out, uint32_t* hull);
int adk_area_diameter_finish_dev(adk_ctx* ctx, const double* d_x, const double* d_y, uint32_t n, uint64_t kmax, double amax,
                                 double dmax, const adk_partial* best, uint32_t flags, adk_result* out, uint32_t* hull);

/* ---- precompute stages, exposed for parity tests ---------------------------------------------------------------- */
int adk_count_tables(adk_ctx* ctx, const double* x, const double* y, uint32_t n,
                     int32_t* below /* n*n */, int32_t* coll /* n*n */, int32_t* right_below /* n */);
int adk_points_in_triangle(adk_ctx* ctx, const double* x, const double* y, uint32_t n,
                           const uint32_t* triples /* 3*m */, uint32_t m, int32_t* out /* m */);
int adk_cw_order(adk_ctx* ctx, const double* x, const double* y, uint32_t n, int32_t* order /* n*(n-1) */);
/* The same clockwise lists as adk_cw_order for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  NO COUNTERPART IN THE REFERENCE.
 * d_x, d_y (n doubles each) and d_order (n*(n-1) int32) are DEVICE pointers owned by the caller.  The call allocates nothing
 * (no cudaMalloc / cudaFree / allocating library routine): the n x n angle and distance tables of the stage live in a static
 * device array of 2 * ADK_NOALLOC_MAX_POINTS^2 doubles (4 MiB, held by every GPU that loads the library until exit).
 * n > ADK_NOALLOC_MAX_POINTS is rejected with ADK_ERR_ARG and a text in adk_last_error.  The two kernels are queued on the
 * context's stream (adk_set_stream) and the call returns WITHOUT synchronising; the first call on a device also resolves the
 * array's address (cudaGetSymbolAddress), so make one call before starting a capture.  Every call shares the array: calls that
 * overlap on the GPU (several streams, contexts or threads) race. */
#define ADK_NOALLOC_MAX_POINTS 512u
int adk_cw_order_noalloc(adk_ctx* ctx, const double* d_x, const double* d_y, uint32_t n, int32_t* d_order /* n*(n-1) */);
/* slope order around centre j for anchor i (Eppstein.cpp:62-114); out capacity n; *len receives the length */
int adk_slope_order(adk_ctx* ctx, const double* x, const double* y, uint32_t n, uint32_t i, uint32_t j,
                    int32_t* out, uint32_t* len);
/* lune lists of the candidate pair (s,t) (AntipodalOptimized.cpp:39-109); capacities n+2 */
int adk_lune_lists(adk_ctx* ctx, const double* x, const double* y, uint32_t n, uint32_t s, uint32_t t,
                   int32_t* left, uint32_t* nleft, int32_t* right, uint32_t* nright);

#ifdef __cplusplus
}
#endif
#endif
