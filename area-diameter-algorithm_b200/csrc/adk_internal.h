// adk_internal.h — host-side context and launcher declarations shared by the .cu files.  Not installed.
#pragma once
#include <cstdint>
#include <cstddef>
#include <string>
#include <vector>
#include <cuda_runtime.h>

#include "../../include/adk.h"

namespace adk
{
// A grow-only device buffer.
struct DevBuf
{
    void* p = nullptr;
    size_t cap = 0;
    template <typename T> T* as() const { return static_cast<T*>(p); }
};

struct Ctx
{
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr, ev3 = nullptr;
    std::string err;
    uint64_t budget = 0;           // DP workspace cap in bytes
    uint64_t launches = 0;         // kernels launched by the current call
    uint64_t total_launches = 0;   // ... by all earlier calls
    bool own_stream = true;
    bool debug = false;            // ADK_DEBUG, read once in adk_create
    uint32_t force_cell_warps = 0; // ADK_FORCE_CELL_WARPS (experiments only), read once in adk_create
    uint32_t area_cluster = 0;     // ADK_AREA_CLUSTER=1|2|4|8: CTAs per anchor in the row sweep (0 = chosen per launch), read once in adk_create
    bool area_wavefront = false;   // ADK_AREA_WAVEFRONT=1 (A/B measurements of the old level wavefront), read once in adk_create

    // point set currently resident
    uint32_t n = 0;
    DevBuf x, y;
    // precompute products
    DevBuf ang, d2c;               // n*n double: angle / squared distance of p seen from centre c
    DevBuf cw;                     // n*(n-1) int32: clockwise neighbour lists
    DevBuf modpi;                  // n*(n-1) int32: neighbours by direction mod pi (descending), ties nearer first
    DevBuf rank;                   // n*n uint16: rank[c*n+p] = position of p in modpi[c]
    DevBuf below, coll, rb;        // count tables (int32)
    DevBuf yorder, above;          // n int32: points by ascending y; above[i] = #points with y >= y_i (excluding i)
    bool have_cw = false, have_modpi = false, have_tables = false, have_yorder = false;

    // scratch / arenas
    DevBuf scratch0, scratch1, scratch2, scratch3, scratch4, scratch5, scratch6;
    DevBuf arena;                  // AD per-CTA slices (directories + layer vectors)
    DevBuf ad_order, ad_ovf, ad_cta;   // AD: processing order, overflow lists, per-CTA bests
    DevBuf ad_aux;                 // AD: per-CTA order / suffix-start tables and spill workspaces
    DevBuf area_boards;            // A: per-cluster band progress boards of the row sweep
    DevBuf area_slot_off;          // A: start of every resident cluster's table slot in scratch5 (entries)
    DevBuf area_mcr;               // A: per table entry, the largest c among the records that read it (stop rule of the row sweep)
    DevBuf area_keep_hull;         // A: per-cluster polygon of the cluster's best anchor (walked back inside the row sweep)
    bool area_keep_valid = false;  // ... valid for the resident point set and these bounds (see area_finish)
    uint32_t area_keep_slots = 0, area_keep_stride = 0;
    uint64_t area_keep_K = 0;
    double area_keep_amax = 0.0;
    DevBuf tile_buf;               // tiled workflow: the slide, the window grid, populations / order, one window instance
    uint64_t ad_V = 0, ad_Q = 0;   // per-slice capacities the arena was initialised for (grow-only)
    uint32_t ad_L = 0, ad_slices = 0;
    size_t ad_slice_bytes = 0;
    // the winning candidate of the last area+diameter partial solve, still resident in its arena slice (see ad_finish)
    bool ad_keep_valid = false;
    uint32_t ad_keep_slice = 0, ad_keep_s = 0, ad_keep_t = 0, ad_keep_nl = 0, ad_keep_nr = 0, ad_keep_count = 0;
    uint64_t ad_keep_kmax = 0, ad_keep_states = 0;
    double ad_keep_amax = 0.0;
    unsigned char ad_keep_layout[96] = {0};   // the CellLayout the slice was written with
    std::vector<double> host_xy;   // host copy of the resident point set (x then y): same points => precompute products stay valid
    const double* dev_x = nullptr; // device pointers the resident set was last taken from (a finish after its own partial
    const double* dev_y = nullptr; //   is recognised without reading anything back)
    bool points_valid = false;
    bool reuse = true;             // keep the precompute products across solves of the same point set (adk_set_reuse)
    DevBuf small;                  // small result records
    void* h_pinned = nullptr;      // pinned host staging
    size_t h_pinned_cap = 0;
};

int set_cuda_error(Ctx* c, cudaError_t e, const char* what);
int ensure(Ctx* c, DevBuf& b, size_t bytes);
int ensure_pinned(Ctx* c, size_t bytes);
void release_idle(Ctx* c, const DevBuf* keep);
uint64_t workspace_budget(Ctx* c, const DevBuf& target);
// Makes the context's device current for the duration of an entry point and restores the caller's device afterwards.
struct DeviceGuard
{
    int prev = -1, dev = -1;
    bool ok = false;
    explicit DeviceGuard(int d) : dev(d) { if (cudaGetDevice(&prev) != cudaSuccess) prev = -1; ok = prev == d || cudaSetDevice(d) == cudaSuccess; }
    ~DeviceGuard() { if (prev >= 0 && prev != dev) cudaSetDevice(prev); }
};
#define ADK_CUDA(ctx, call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return adk::set_cuda_error((ctx), e__, #call); } while (0)
#define ADK_TRY(call) do { int r__ = (call); if (r__ != ADK_OK) return r__; } while (0)

// ---- precompute (adk_precompute.cu) ----
int upload_points(Ctx* c, const double* x, const double* y, uint32_t n, bool device_ptrs, bool new_solve);
int run_angles_and_orders(Ctx* c, bool need_cw, bool need_modpi);
int run_count_tables(Ctx* c);
int run_yorder(Ctx* c);
int run_pit_query(Ctx* c, const uint32_t* h_triples, uint32_t m, int32_t* h_out);
int cw_order_noalloc(Ctx* c, const double* d_x, const double* d_y, uint32_t n, int32_t* d_order);

// ---- area path (adk_area.cu) ----
int area_partial(Ctx* c, uint64_t kmax, double amax, uint32_t shard, uint32_t nshards, adk_partial* out, adk_stats* st);
int area_finish(Ctx* c, uint64_t kmax, double amax, const adk_partial* best, uint32_t flags, adk_result* out, uint32_t* hull);
int area_slope_order(Ctx* c, uint32_t i, uint32_t j, int32_t* h_out, uint32_t* h_len);

// ---- tiled area-only workflow (adk_tiled.cu) ----
int area_tiled(Ctx* c, const double* hx, const double* hy, uint32_t n, double patch, double step, uint64_t kmax, double amax,
               adk_result* out, uint32_t* hull, adk_tiled_stats* st);

// ---- area+diameter path (adk_ad.cu) ----
int ad_partial(Ctx* c, uint64_t kmax, double amax, double dmax, uint32_t shard, uint32_t nshards, adk_partial* out, adk_stats* st);
int ad_finish(Ctx* c, uint64_t kmax, double amax, double dmax, const adk_partial* best, uint32_t flags, adk_result* out, uint32_t* hull);
int ad_lune_lists(Ctx* c, uint32_t s, uint32_t t, int32_t* h_left, uint32_t* nl, int32_t* h_right, uint32_t* nr);

}  // namespace adk
