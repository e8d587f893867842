// adk_capi.cu — the C ABI (include/adk.h): context management and the thin entry points over the kernels.
#include "adk_internal.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

namespace adk
{

int set_cuda_error(Ctx* c, cudaError_t e, const char* what)
{
    if (c)
    {
        c->err = std::string("CUDA error: ") + cudaGetErrorString(e) + " at " + what;
    }
    return ADK_ERR_CUDA;
}

// Frees the large workspaces that the buffer being grown does not depend on: the AD arena (area path growing) or the
// area path's tables (AD path growing).  cudaFree synchronises, so nothing in flight still reads them.
void release_idle(Ctx* c, const DevBuf* keep)
{
    auto drop = [&](DevBuf& b) { if (&b != keep && b.p) { cudaFree(b.p); b.p = nullptr; b.cap = 0; } };
    if (keep == &c->arena) { drop(c->scratch5); drop(c->scratch4); drop(c->scratch2); drop(c->scratch1); drop(c->area_mcr); c->area_keep_valid = false; }
    else if (c->arena.p)
    {
        drop(c->arena); drop(c->ad_aux);
        c->ad_keep_valid = false;
        c->ad_V = c->ad_Q = 0; c->ad_L = 0; c->ad_slices = 0; c->ad_slice_bytes = 0;
    }
}

// Bytes a workspace may grow to right now: the configured budget, capped by what the device really has free
// (counting the buffer's own current allocation), after releasing the other path's idle workspaces if needed.
uint64_t workspace_budget(Ctx* c, const DevBuf& target)
{
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) { cudaGetLastError(); return c->budget; }
    uint64_t avail = (uint64_t) ((free_b + target.cap) * 0.92);
    if (avail < c->budget)
    {
        release_idle(c, &target);
        if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) avail = (uint64_t) ((free_b + target.cap) * 0.92);
    }
    return std::min<uint64_t>(c->budget, avail);
}

int ensure(Ctx* c, DevBuf& b, size_t bytes)
{
    if (bytes <= b.cap) return ADK_OK;
    if (b.p) { cudaError_t e = cudaFree(b.p); b.p = nullptr; b.cap = 0; if (e != cudaSuccess) return set_cuda_error(c, e, "cudaFree"); }
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&b.p, want);
    if (e != cudaSuccess)
    {
        cudaGetLastError();
        want = bytes;
        e = cudaMalloc(&b.p, want);
    }
    if (e != cudaSuccess)
    {
        cudaGetLastError();
        release_idle(c, &b);
        e = cudaMalloc(&b.p, want);
        if (e != cudaSuccess) { b.p = nullptr; cudaGetLastError(); c->err = "device allocation failed"; return ADK_ERR_NOMEM; }
    }
    b.cap = want;
    return ADK_OK;
}

int ensure_pinned(Ctx* c, size_t bytes)
{
    if (bytes <= c->h_pinned_cap) return ADK_OK;
    if (c->h_pinned) cudaFreeHost(c->h_pinned);
    c->h_pinned = nullptr; c->h_pinned_cap = 0;
    const size_t want = std::max<size_t>(bytes * 2, 1 << 16);
    ADK_CUDA(c, cudaMallocHost(&c->h_pinned, want));
    c->h_pinned_cap = want;
    return ADK_OK;
}

static void free_buf(DevBuf& b) { if (b.p) cudaFree(b.p); b.p = nullptr; b.cap = 0; }

}  // namespace adk

using adk::Ctx;

struct adk_ctx { Ctx c; };

static thread_local std::string g_create_error;

extern "C" {

int adk_create(adk_ctx** out_ctx, int device)
{
    if (!out_ctx) return ADK_ERR_ARG;
    *out_ctx = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
    {
        g_create_error = std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        cudaGetLastError();
        return ADK_ERR_CUDA;   // there is no CPU fallback
    }
    if (device < 0) { if (cudaGetDevice(&device) != cudaSuccess) return ADK_ERR_CUDA; }
    if (device >= ndev) return ADK_ERR_ARG;
    adk::DeviceGuard guard(device);
    if (!guard.ok) return ADK_ERR_CUDA;
    adk_ctx* h = new (std::nothrow) adk_ctx();
    if (!h) return ADK_ERR_NOMEM;
    Ctx* c = &h->c;
    c->device = device;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) c->sm_count = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { delete h; return ADK_ERR_CUDA; }
    if (cudaEventCreate(&c->ev0) != cudaSuccess || cudaEventCreate(&c->ev1) != cudaSuccess || cudaEventCreate(&c->ev2) != cudaSuccess ||
        cudaEventCreate(&c->ev3) != cudaSuccess)
    {
        g_create_error = "cudaEventCreate failed";
        adk_destroy(h);
        return ADK_ERR_CUDA;
    }
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) c->budget = (uint64_t) (free_b * 0.75);
    else c->budget = 8ull << 30;
    c->debug = getenv("ADK_DEBUG") != nullptr;
    if (const char* f = getenv("ADK_FORCE_CELL_WARPS")) c->force_cell_warps = (uint32_t) atoi(f);
    if (const char* f = getenv("ADK_AREA_WAVEFRONT")) c->area_wavefront = f[0] == '1';
    if (const char* f = getenv("ADK_AREA_CLUSTER")) { const int v = atoi(f); if (v == 1 || v == 2 || v == 4 || v == 8) c->area_cluster = (uint32_t) v; }
    *out_ctx = h;
    return ADK_OK;
}

void adk_destroy(adk_ctx* h)
{
    if (!h) return;
    Ctx* c = &h->c;
    adk::DeviceGuard guard(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    adk::DevBuf* bufs[] = {&c->x, &c->y, &c->ang, &c->d2c, &c->cw, &c->modpi, &c->rank, &c->below, &c->coll, &c->rb, &c->yorder,
                           &c->above, &c->scratch0, &c->scratch1, &c->scratch2, &c->scratch3, &c->scratch4, &c->scratch5,
                           &c->scratch6, &c->arena, &c->small, &c->ad_order, &c->ad_ovf, &c->ad_cta, &c->ad_aux, &c->area_boards, &c->area_keep_hull, &c->area_mcr, &c->area_slot_off, &c->tile_buf};
    for (adk::DevBuf* b : bufs) adk::free_buf(*b);
    if (c->h_pinned) cudaFreeHost(c->h_pinned);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->ev2) cudaEventDestroy(c->ev2);
    if (c->ev3) cudaEventDestroy(c->ev3);
    if (c->stream && c->own_stream) cudaStreamDestroy(c->stream);
    delete h;
}

const char* adk_last_error(const adk_ctx* h) { return h ? h->c.err.c_str() : g_create_error.c_str(); }

int adk_set_memory_budget(adk_ctx* h, uint64_t bytes)
{
    if (!h || bytes < (1ull << 20)) return ADK_ERR_ARG;
    h->c.budget = bytes;
    return ADK_OK;
}

int adk_set_stream(adk_ctx* h, void* cuda_stream)
{
    if (!h) return ADK_ERR_ARG;
    Ctx* c = &h->c;
    adk::DeviceGuard guard(c->device);
    if (!guard.ok) return ADK_ERR_CUDA;
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->stream && c->own_stream) cudaStreamDestroy(c->stream);
    c->stream = static_cast<cudaStream_t>(cuda_stream);
    c->own_stream = false;
    return ADK_OK;
}

int adk_set_reuse(adk_ctx* h, int enabled)
{
    if (!h) return ADK_ERR_ARG;
    h->c.reuse = enabled != 0;
    return ADK_OK;
}

uint64_t adk_launch_count(const adk_ctx* h) { return h ? h->c.total_launches + h->c.launches : 0; }

#define ADK_ENTER(h) if (!(h)) return ADK_ERR_ARG; Ctx* c = &(h)->c; c->err.clear(); c->total_launches += c->launches; c->launches = 0; \
    adk::DeviceGuard guard__(c->device); if (!guard__.ok) return ADK_ERR_CUDA

static void fill_result_from(adk_result* out, const adk_stats& st, uint32_t flags)
{
    out->cells_nominal = st.cells_nominal; out->cells_executed = st.cells_executed;
    out->states = st.states; out->pairs = st.pairs; out->kernel_ms = st.kernel_ms; out->dp_ms = st.dp_ms;
    out->entries_alloc = -1; out->entries_required = -1;
    (void) flags;
}

int adk_area_partial(adk_ctx* h, const double* x, const double* y, uint32_t n, uint64_t kmax, double amax,
                     uint32_t shard, uint32_t nshards, adk_partial* out, adk_stats* stats)
{
    ADK_ENTER(h);
    if (!out) return ADK_ERR_ARG;
    ADK_TRY(adk::upload_points(c, x, y, n, false, true));
    return adk::area_partial(c, kmax, amax, shard, nshards, out, stats);
}

int adk_area_partial_dev(adk_ctx* h, const double* d_x, const double* d_y, uint32_t n, uint64_t kmax, double amax,
                         uint32_t shard, uint32_t nshards, adk_partial* out, adk_stats* stats)
{
    ADK_ENTER(h);
    if (!out) return ADK_ERR_ARG;
    ADK_TRY(adk::upload_points(c, d_x, d_y, n, true, true));
    return adk::area_partial(c, kmax, amax, shard, nshards, out, stats);
}

int adk_area_finish(adk_ctx* h, const double* x, const double* y, uint32_t n, uint64_t kmax, double amax,
                    const adk_partial* best, uint32_t flags, adk_result* out, uint32_t* hull)
{
    ADK_ENTER(h);
    if (!out || !best || ((flags & ADK_FLAG_RECONSTRUCT) && !hull)) return ADK_ERR_ARG;
    std::memset(out, 0, sizeof(*out));
    out->entries_alloc = -1; out->entries_required = -1;
    if (!best->found) return ADK_OK;
    ADK_TRY(adk::upload_points(c, x, y, n, false, false));
    return adk::area_finish(c, kmax, amax, best, flags, out, hull);
}

int adk_area_finish_dev(adk_ctx* h, const double* d_x, const double* d_y, uint32_t n, uint64_t kmax, double amax,
                        const adk_partial* best, uint32_t flags, adk_result* out, uint32_t* hull)
{
    ADK_ENTER(h);
    if (!out || !best || ((flags & ADK_FLAG_RECONSTRUCT) && !hull)) return ADK_ERR_ARG;
    std::memset(out, 0, sizeof(*out));
    out->entries_alloc = -1; out->entries_required = -1;
    if (!best->found) return ADK_OK;
    ADK_TRY(adk::upload_points(c, d_x, d_y, n, true, false));
    return adk::area_finish(c, kmax, amax, best, flags, out, hull);
}

int adk_solve_area(adk_ctx* h, const double* x, const double* y, uint32_t n, uint64_t kmax, double amax, uint32_t flags,
                   adk_result* out, uint32_t* hull)
{
    ADK_ENTER(h);
    if (!out || ((flags & ADK_FLAG_RECONSTRUCT) && !hull)) return ADK_ERR_ARG;
    std::memset(out, 0, sizeof(*out));
    out->entries_alloc = -1; out->entries_required = -1;
    if (n < 3 || kmax < 3) return ADK_OK;
    ADK_TRY(adk::upload_points(c, x, y, n, false, true));
    adk_partial part; adk_stats st;
    ADK_TRY(adk::area_partial(c, kmax, amax, 0, 1, &part, &st));
    const uint64_t launches = c->launches;
    ADK_TRY(adk::area_finish(c, kmax, amax, &part, flags, out, hull));
    (void) launches;
    fill_result_from(out, st, flags);
    if (flags & ADK_FLAG_COUNTERS)
    {   // Eppstein.cpp:171-175: the reference's dense table size; "required" is reported by the DP kernels
        const uint64_t K = std::min<uint64_t>(kmax, n);
        out->entries_alloc = (int64_t) ((K + 1) * 2 * (uint64_t) n * n);
        out->entries_required = (int64_t) st.states;      // area path: prefix-minimum change points (Eppstein.cpp:228-232)
    }
    return ADK_OK;
}

int adk_area_diameter_partial(adk_ctx* h, const double* x, const double* y, uint32_t n, uint64_t kmax, double amax,
                              double dmax, uint32_t shard, uint32_t nshards, adk_partial* out, adk_stats* stats)
{
    ADK_ENTER(h);
    if (!out) return ADK_ERR_ARG;
    ADK_TRY(adk::upload_points(c, x, y, n, false, true));
    return adk::ad_partial(c, kmax, amax, dmax, shard, nshards, out, stats);
}

int adk_area_diameter_partial_dev(adk_ctx* h, const double* d_x, const double* d_y, uint32_t n, uint64_t kmax, double amax,
                                  double dmax, uint32_t shard, uint32_t nshards, adk_partial* out, adk_stats* stats)
{
    ADK_ENTER(h);
    if (!out) return ADK_ERR_ARG;
    ADK_TRY(adk::upload_points(c, d_x, d_y, n, true, true));
    return adk::ad_partial(c, kmax, amax, dmax, shard, nshards, out, stats);
}

int adk_area_diameter_finish(adk_ctx* h, const double* x, const double* y, uint32_t n, uint64_t kmax, double amax,
                             double dmax, const adk_partial* best, uint32_t flags, adk_result* out, uint32_t* hull)
{
    ADK_ENTER(h);
    if (!out || !best || ((flags & ADK_FLAG_RECONSTRUCT) && !hull)) return ADK_ERR_ARG;
    std::memset(out, 0, sizeof(*out));
    out->entries_alloc = -1; out->entries_required = -1;
    if (!best->found) return ADK_OK;
    ADK_TRY(adk::upload_points(c, x, y, n, false, false));
    const int rc = adk::ad_finish(c, kmax, amax, dmax, best, flags, out, hull);
    out->states = out->entries_required > 0 ? (uint64_t) out->entries_required : 0;      // the winning candidate's states (its DP re-run in the reference, :529-535)
    out->entries_required = -1;
    return rc;
}

int adk_area_diameter_finish_dev(adk_ctx* h, const double* d_x, const double* d_y, uint32_t n, uint64_t kmax, double amax,
                                 double dmax, const adk_partial* best, uint32_t flags, adk_result* out, uint32_t* hull)
{
    ADK_ENTER(h);
    if (!out || !best || ((flags & ADK_FLAG_RECONSTRUCT) && !hull)) return ADK_ERR_ARG;
    std::memset(out, 0, sizeof(*out));
    out->entries_alloc = -1; out->entries_required = -1;
    if (!best->found) return ADK_OK;
    ADK_TRY(adk::upload_points(c, d_x, d_y, n, true, false));
    const int rc = adk::ad_finish(c, kmax, amax, dmax, best, flags, out, hull);
    out->states = out->entries_required > 0 ? (uint64_t) out->entries_required : 0;      // the winning candidate's states (its DP re-run in the reference, :529-535)
    out->entries_required = -1;
    return rc;
}

int adk_solve_area_diameter(adk_ctx* h, const double* x, const double* y, uint32_t n, uint64_t kmax, double amax,
                            double dmax, uint32_t flags, adk_result* out, uint32_t* hull)
{
    ADK_ENTER(h);
    if (!out || ((flags & ADK_FLAG_RECONSTRUCT) && !hull)) return ADK_ERR_ARG;
    std::memset(out, 0, sizeof(*out));
    out->entries_alloc = -1; out->entries_required = -1;
    if (std::min<uint64_t>(n, kmax) < 3) return ADK_OK;
    ADK_TRY(adk::upload_points(c, x, y, n, false, true));
    adk_partial part; adk_stats st;
    ADK_TRY(adk::ad_partial(c, kmax, amax, dmax, 0, 1, &part, &st));
    ADK_TRY(adk::ad_finish(c, kmax, amax, dmax, &part, flags, out, hull));
    // AntipodalOptimized.cpp:519-524, 529-535: every pair's hash entries, plus the winner's once more when its DP is re-run
    const int64_t again = (out->found && (flags & ADK_FLAG_RECONSTRUCT) && out->entries_required > 0) ? out->entries_required : 0;
    out->entries_required = -1;
    fill_result_from(out, st, flags);
    if (flags & ADK_FLAG_COUNTERS) { out->entries_alloc = (int64_t) st.states + again; out->entries_required = (int64_t) st.states + again; }
    return ADK_OK;
}

static double batch_cost(const adk_instance& in)
{
    const double n = (double) in.n;
    return in.algo == ADK_ALGO_AREA_DIAMETER ? n * n * n * n : n * n * n;
}

int adk_batch_assign(const adk_instance* inst, uint32_t ninst, uint32_t nshards, uint32_t* owner)
{
    if (((!inst || !owner) && ninst) || nshards == 0) return ADK_ERR_ARG;
    std::vector<uint32_t> order(ninst);
    for (uint32_t i = 0; i < ninst; ++i) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return batch_cost(inst[a]) > batch_cost(inst[b]); });
    std::vector<double> load(nshards, 0.0);
    for (uint32_t i : order)
    {   // heaviest instance first, onto the least loaded rank (lowest rank on ties)
        uint32_t best = 0;
        for (uint32_t r = 1; r < nshards; ++r) if (load[r] < load[best]) best = r;
        owner[i] = best; load[best] += batch_cost(inst[i]);
    }
    return ADK_OK;
}

int adk_solve_batch(adk_ctx* h, const adk_instance* inst, uint32_t ninst, uint32_t shard, uint32_t nshards, adk_result* out)
{
    if (!h) return ADK_ERR_ARG;
    if ((!inst && ninst) || !out || nshards == 0 || shard >= nshards) { h->c.err = "adk_solve_batch: bad argument"; return ADK_ERR_ARG; }
    std::vector<uint32_t> owner(ninst);
    int rc = adk_batch_assign(inst, ninst, nshards, owner.data());
    if (rc != ADK_OK) return rc;
    std::vector<uint32_t> mine;
    for (uint32_t i = 0; i < ninst; ++i)
    {
        std::memset(&out[i], 0, sizeof(adk_result));
        if (owner[i] == shard) mine.push_back(i); else out[i].found = -1;
    }
    std::stable_sort(mine.begin(), mine.end(), [&](uint32_t a, uint32_t b) { return batch_cost(inst[a]) > batch_cost(inst[b]); });
    for (uint32_t i : mine)
    {
        const adk_instance& in = inst[i];
        if ((in.flags & ADK_FLAG_RECONSTRUCT) && (!in.hull || in.hull_cap < in.n)) { h->c.err = "adk_solve_batch: hull buffer too small"; return ADK_ERR_ARG; }
        if (in.algo == ADK_ALGO_AREA) rc = adk_solve_area(h, in.x, in.y, in.n, in.kmax, in.amax, in.flags, &out[i], in.hull);
        else if (in.algo == ADK_ALGO_AREA_DIAMETER) rc = adk_solve_area_diameter(h, in.x, in.y, in.n, in.kmax, in.amax, in.dmax, in.flags, &out[i], in.hull);
        else { h->c.err = "adk_solve_batch: unknown algo"; rc = ADK_ERR_ARG; }
        if (rc != ADK_OK) return rc;
    }
    return ADK_OK;
}

int adk_solve_area_tiled(adk_ctx* h, const double* x, const double* y, uint32_t n, double patch_size, double step_size,
                         uint64_t kmax, double amax, adk_result* out, uint32_t* hull, adk_tiled_stats* stats)
{
    ADK_ENTER(h);
    if (!out || !hull) return ADK_ERR_ARG;
    return adk::area_tiled(c, x, y, n, patch_size, step_size, kmax, amax, out, hull, stats);
}

int adk_merge(const adk_partial* parts, uint32_t count, adk_partial* out)
{
    if (!parts || !out) return ADK_ERR_ARG;
    std::memset(out, 0, sizeof(*out));
    for (uint32_t i = 0; i < count; ++i)
    {
        const adk_partial& p = parts[i];
        if (!p.found) continue;
        const bool better = !out->found || p.count > out->count ||
                            (p.count == out->count && (p.area < out->area || (p.area == out->area && p.tie_key < out->tie_key)));
        if (better) *out = p;
    }
    return ADK_OK;
}

int adk_count_tables(adk_ctx* h, const double* x, const double* y, uint32_t n, int32_t* below, int32_t* coll, int32_t* right_below)
{
    ADK_ENTER(h);
    if (!below || !coll || !right_below) return ADK_ERR_ARG;
    ADK_TRY(adk::upload_points(c, x, y, n, false, false));
    ADK_TRY(adk::run_count_tables(c));
    ADK_CUDA(c, cudaMemcpyAsync(below, c->below.p, sizeof(int32_t) * (size_t) n * n, cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaMemcpyAsync(coll, c->coll.p, sizeof(int32_t) * (size_t) n * n, cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaMemcpyAsync(right_below, c->rb.p, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaStreamSynchronize(c->stream));
    return ADK_OK;
}

int adk_points_in_triangle(adk_ctx* h, const double* x, const double* y, uint32_t n, const uint32_t* triples, uint32_t m, int32_t* out)
{
    ADK_ENTER(h);
    if (!triples || !out) return ADK_ERR_ARG;
    for (uint32_t t = 0; t < 3 * m; ++t) if (triples[t] >= n) return ADK_ERR_ARG;
    ADK_TRY(adk::upload_points(c, x, y, n, false, false));
    return adk::run_pit_query(c, triples, m, out);
}

int adk_cw_order(adk_ctx* h, const double* x, const double* y, uint32_t n, int32_t* order)
{
    ADK_ENTER(h);
    if (!order || n < 2) return ADK_ERR_ARG;
    ADK_TRY(adk::upload_points(c, x, y, n, false, false));
    ADK_TRY(adk::run_angles_and_orders(c, true, false));
    ADK_CUDA(c, cudaMemcpyAsync(order, c->cw.p, sizeof(int32_t) * (size_t) n * (n - 1), cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaStreamSynchronize(c->stream));
    return ADK_OK;
}

int adk_cw_order_noalloc(adk_ctx* h, const double* d_x, const double* d_y, uint32_t n, int32_t* d_order)
{
    ADK_ENTER(h);
    return adk::cw_order_noalloc(c, d_x, d_y, n, d_order);
}

int adk_slope_order(adk_ctx* h, const double* x, const double* y, uint32_t n, uint32_t i, uint32_t j, int32_t* out, uint32_t* len)
{
    ADK_ENTER(h);
    if (!out || !len) return ADK_ERR_ARG;
    ADK_TRY(adk::upload_points(c, x, y, n, false, false));
    return adk::area_slope_order(c, i, j, out, len);
}

int adk_lune_lists(adk_ctx* h, const double* x, const double* y, uint32_t n, uint32_t s, uint32_t t,
                   int32_t* left, uint32_t* nleft, int32_t* right, uint32_t* nright)
{
    ADK_ENTER(h);
    if (!left || !right || !nleft || !nright || s >= n || t >= n || s == t) return ADK_ERR_ARG;
    ADK_TRY(adk::upload_points(c, x, y, n, false, false));
    return adk::ad_lune_lists(c, s, t, left, nleft, right, nright);
}

}  // extern "C"
