// adk_multi.cu — several GPUs of one box behind ONE handle, inside the C ABI (include/adk.h, adk_multi_*).
//
// The reference's callers (src/bench/main.cpp:89,100, src/py-binding/main.cpp:29,59) make one blocking call per solve from one
// host thread.  adk_multi keeps that shape: one call, and inside it one host thread per device runs that device's shard of the
// anchors / candidate pairs (adk_*_partial on its own adk_ctx), the ranks exchange their 32-byte local bests with ONE
// ncclAllGather over NVLink (NCCL has no lexicographic max; an all-gather of 32 B per rank followed by the same adk_merge on
// every rank IS the "max-reduce of (count, area, witness)" of the north_star), the rank that owns the winning unit reconstructs
// the witness from its resident records, and one ncclBroadcast hands the result record + hull to every rank.
//
// NCCL is bound lazily (dlopen of libnccl.so.2: the copy already mapped by the process, e.g. PyTorch's, or the system's), so that
// single-GPU users of libadk.so never load it.  There is no fallback: if NCCL cannot be loaded adk_multi_create fails.
#include "adk_internal.h"

#include <dlfcn.h>

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstring>
#include <mutex>
#include <thread>

namespace
{
// ---- the handful of NCCL entry points used, declared here so that neither nccl.h nor libnccl is needed at build time
typedef struct ncclComm* ncclComm_t;
typedef int ncclResult_t;                       // ncclSuccess = 0
constexpr int kNcclUint8 = 1;                   // ncclDataType_t: ncclUint8
struct NcclApi
{
    void* lib = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
    std::string err;
    bool load()
    {
        if (lib) return true;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) { lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (lib) break; }
        if (!lib) { err = std::string("cannot load NCCL: ") + dlerror(); return false; }
        auto sym = [&](const char* n) { void* p = dlsym(lib, n); if (!p) err = std::string("NCCL symbol missing: ") + n; return p; };
        CommInitAll = reinterpret_cast<decltype(CommInitAll)>(sym("ncclCommInitAll"));
        CommDestroy = reinterpret_cast<decltype(CommDestroy)>(sym("ncclCommDestroy"));
        AllGather = reinterpret_cast<decltype(AllGather)>(sym("ncclAllGather"));
        Broadcast = reinterpret_cast<decltype(Broadcast)>(sym("ncclBroadcast"));
        GetErrorString = reinterpret_cast<decltype(GetErrorString)>(sym("ncclGetErrorString"));
        GetVersion = reinterpret_cast<decltype(GetVersion)>(sym("ncclGetVersion"));
        return CommInitAll && CommDestroy && AllGather && Broadcast && GetErrorString;
    }
};
NcclApi g_nccl;
std::mutex g_nccl_mu;

constexpr uint32_t kHdrWords = 16;      // result record travelling with the hull: found, count, area (2 words), hull_len, diam_valid, diam[2], ...

struct Job
{
    int algo = 0;                       // ADK_ALGO_*
    const double* x = nullptr; const double* y = nullptr;
    uint32_t n = 0; uint64_t kmax = 0; double amax = 0.0, dmax = 0.0; uint32_t flags = 0;
};

struct Rank
{
    int device = 0;
    adk_ctx* ctx = nullptr;
    ncclComm_t comm = nullptr;
    cudaStream_t stream = nullptr;
    unsigned char* d_send = nullptr;    // 32 B
    unsigned char* d_recv = nullptr;    // ndev * 32 B
    uint32_t* d_res = nullptr;          // kHdrWords + ADK_MAX_POINTS words
    unsigned char* h_pin = nullptr;     // pinned: ndev * 32 B + (kHdrWords + ADK_MAX_POINTS) * 4 B
    std::thread worker;
    // per-job output
    int rc = ADK_OK;
    std::string err;
    adk_stats st{};
    adk_result res{};
    std::vector<uint32_t> hull;
    float exchange_ms = 0.f;
};
}  // namespace

struct adk_multi
{
    std::vector<Rank> ranks;
    std::string err;
    // a tiny fork/join: the workers sleep between solves
    std::mutex mu;
    std::condition_variable cv_go, cv_done;
    uint64_t epoch = 0;
    int pending = 0;
    bool quit = false;
    Job job;
    adk_multi_stats last{};
};

namespace
{
int fail(Rank& r, int rc, const std::string& what) { r.rc = rc; r.err = what; return rc; }

#define MR_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return fail(r, ADK_ERR_CUDA, std::string("CUDA error: ") + cudaGetErrorString(e__) + " at " #call); } while (0)
#define MR_NCCL(call) do { ncclResult_t e__ = (call); if (e__ != 0) return fail(r, ADK_ERR_CUDA, std::string("NCCL error: ") + g_nccl.GetErrorString(e__) + " at " #call); } while (0)

// One rank's share of one solve.  Every rank takes exactly the same collective calls in the same order, whatever it finds.
int rank_solve(adk_multi* m, uint32_t me)
{
    Rank& r = m->ranks[me];
    const Job& j = m->job;
    const uint32_t nd = (uint32_t) m->ranks.size();
    r.rc = ADK_OK; r.err.clear();
    std::memset(&r.res, 0, sizeof(r.res)); r.res.entries_alloc = -1; r.res.entries_required = -1;
    std::memset(&r.st, 0, sizeof(r.st));
    MR_CUDA(cudaSetDevice(r.device));
    adk_partial mine; std::memset(&mine, 0, sizeof(mine));
    int rc = j.algo == ADK_ALGO_AREA ? adk_area_partial(r.ctx, j.x, j.y, j.n, j.kmax, j.amax, me, nd, &mine, &r.st)
                                     : adk_area_diameter_partial(r.ctx, j.x, j.y, j.n, j.kmax, j.amax, j.dmax, me, nd, &mine, &r.st);
    if (rc != ADK_OK) { r.err = adk_last_error(r.ctx); std::memset(&mine, 0, sizeof(mine)); }
    const int partial_rc = rc;
    // ---- the exchange: 32 bytes per rank
    cudaEvent_t e0, e1;
    MR_CUDA(cudaEventCreate(&e0)); MR_CUDA(cudaEventCreate(&e1));
    adk_partial* h_parts = reinterpret_cast<adk_partial*>(r.h_pin);
    uint32_t* h_res = reinterpret_cast<uint32_t*>(r.h_pin + sizeof(adk_partial) * nd);
    MR_CUDA(cudaEventRecord(e0, r.stream));
    MR_CUDA(cudaMemcpyAsync(r.d_send, &mine, sizeof(mine), cudaMemcpyHostToDevice, r.stream));
    MR_NCCL(g_nccl.AllGather(r.d_send, r.d_recv, sizeof(adk_partial), kNcclUint8, r.comm, r.stream));
    MR_CUDA(cudaMemcpyAsync(h_parts, r.d_recv, sizeof(adk_partial) * nd, cudaMemcpyDeviceToHost, r.stream));
    MR_CUDA(cudaStreamSynchronize(r.stream));
    adk_partial best;
    adk_merge(h_parts, nd, &best);
    uint32_t owner = 0;
    if (best.found)
        for (uint32_t q = 0; q < nd; ++q)
            if (h_parts[q].found && h_parts[q].count == best.count && h_parts[q].tie_key == best.tie_key && h_parts[q].area == best.area) { owner = q; break; }
    // ---- the owner of the winning unit reconstructs (its records are still resident) and broadcasts record + hull
    const uint32_t words = kHdrWords + j.n;
    int finish_rc = ADK_OK;
    if (me == owner)
    {
        std::memset(h_res, 0, sizeof(uint32_t) * words);
        if (partial_rc == ADK_OK)
        {
            adk_result fr;
            finish_rc = j.algo == ADK_ALGO_AREA ? adk_area_finish(r.ctx, j.x, j.y, j.n, j.kmax, j.amax, &best, j.flags, &fr, h_res + kHdrWords)
                                                : adk_area_diameter_finish(r.ctx, j.x, j.y, j.n, j.kmax, j.amax, j.dmax, &best, j.flags, &fr, h_res + kHdrWords);
            if (finish_rc == ADK_OK)
            {
                h_res[0] = (uint32_t) fr.found; h_res[1] = fr.count; std::memcpy(h_res + 2, &fr.area, sizeof(double));
                h_res[4] = fr.hull_len; h_res[5] = fr.diam_valid; h_res[6] = fr.diam[0]; h_res[7] = fr.diam[1];
                std::memcpy(h_res + 10, &fr.states, sizeof(uint64_t));        // the winner's own states (area+diameter path)
            }
            else r.err = adk_last_error(r.ctx);
        }
        h_res[8] = (uint32_t) finish_rc;
        MR_CUDA(cudaMemcpyAsync(r.d_res, h_res, sizeof(uint32_t) * words, cudaMemcpyHostToDevice, r.stream));
    }
    MR_NCCL(g_nccl.Broadcast(r.d_res, r.d_res, sizeof(uint32_t) * words, kNcclUint8, (int) owner, r.comm, r.stream));
    MR_CUDA(cudaMemcpyAsync(h_res, r.d_res, sizeof(uint32_t) * words, cudaMemcpyDeviceToHost, r.stream));
    MR_CUDA(cudaEventRecord(e1, r.stream));
    MR_CUDA(cudaStreamSynchronize(r.stream));
    cudaEventElapsedTime(&r.exchange_ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (partial_rc != ADK_OK) return fail(r, partial_rc, r.err);
    if (h_res[8] != ADK_OK) return fail(r, (int) h_res[8], me == owner ? r.err : std::string("the owning rank could not reconstruct the witness"));
    r.res.found = (int32_t) h_res[0]; r.res.count = h_res[1]; std::memcpy(&r.res.area, h_res + 2, sizeof(double));
    r.res.hull_len = h_res[4]; r.res.diam_valid = h_res[5]; r.res.diam[0] = h_res[6]; r.res.diam[1] = h_res[7];
    std::memcpy(&r.res.states, h_res + 10, sizeof(uint64_t));
    r.hull.assign(h_res + kHdrWords, h_res + kHdrWords + std::min<uint32_t>(r.res.hull_len, j.n));
    return ADK_OK;
}

void worker_main(adk_multi* m, uint32_t me)
{
    uint64_t seen = 0;
    for (;;)
    {
        {
            std::unique_lock<std::mutex> lk(m->mu);
            m->cv_go.wait(lk, [&] { return m->quit || m->epoch != seen; });
            if (m->quit) return;
            seen = m->epoch;
        }
        rank_solve(m, me);
        {
            std::lock_guard<std::mutex> lk(m->mu);
            if (--m->pending == 0) m->cv_done.notify_all();
        }
    }
}

int run_job(adk_multi* m, const Job& job, adk_result* out, uint32_t* hull)
{
    if (!m || !out || ((job.flags & ADK_FLAG_RECONSTRUCT) && !hull)) return ADK_ERR_ARG;
    std::memset(out, 0, sizeof(*out)); out->entries_alloc = -1; out->entries_required = -1;
    m->err.clear();
    if (!job.x || !job.y || job.n == 0 || job.n > ADK_MAX_POINTS) { m->err = "bad point arrays / n out of range"; return ADK_ERR_ARG; }
    if (std::min<uint64_t>(job.n, job.kmax) < 3) return ADK_OK;             // Eppstein.cpp:137-140, AntipodalOptimized.cpp:461-464
    {
        std::lock_guard<std::mutex> lk(m->mu);
        m->job = job; m->pending = (int) m->ranks.size(); ++m->epoch;
    }
    m->cv_go.notify_all();
    {
        std::unique_lock<std::mutex> lk(m->mu);
        m->cv_done.wait(lk, [&] { return m->pending == 0; });
    }
    adk_multi_stats& s = m->last;
    std::memset(&s, 0, sizeof(s));
    s.ndev = (uint32_t) m->ranks.size();
    for (const Rank& r : m->ranks)
    {
        if (r.rc != ADK_OK) { m->err = "device " + std::to_string(r.device) + ": " + r.err; return r.rc; }
        s.cells_nominal += r.st.cells_nominal; s.cells_executed += r.st.cells_executed; s.states += r.st.states; s.launches += r.st.launches;
        s.pairs = std::max<uint64_t>(s.pairs, r.st.pairs);
        s.max_kernel_ms = std::max(s.max_kernel_ms, r.st.kernel_ms); s.max_dp_ms = std::max(s.max_dp_ms, r.st.dp_ms);
        s.sum_dp_ms += r.st.dp_ms;
        s.max_exchange_ms = std::max<double>(s.max_exchange_ms, r.exchange_ms);
    }
    const Rank& r0 = m->ranks[0];
    *out = r0.res;
    const uint64_t winner_states = r0.res.states;
    out->cells_nominal = s.cells_nominal; out->cells_executed = s.cells_executed; out->states = s.states; out->pairs = s.pairs;
    out->kernel_ms = s.max_kernel_ms; out->dp_ms = s.max_dp_ms;
    out->entries_alloc = -1; out->entries_required = -1;
    if (job.flags & ADK_FLAG_COUNTERS)
    {
        if (job.algo == ADK_ALGO_AREA)
        {   // Eppstein.cpp:171-175: the reference's dense table size; required = prefix-minimum change points over all anchors
            const uint64_t K = std::min<uint64_t>(job.kmax, job.n);
            out->entries_alloc = (int64_t) ((K + 1) * 2 * (uint64_t) job.n * job.n);
            out->entries_required = (int64_t) s.states;
        }
        else
        {   // AntipodalOptimized.cpp:519-524, 529-535: every pair's hash entries, plus the winner's once more (its DP is re-run there)
            const uint64_t again = (out->found && (job.flags & ADK_FLAG_RECONSTRUCT)) ? winner_states : 0;
            out->entries_alloc = (int64_t) (s.states + again); out->entries_required = (int64_t) (s.states + again);
        }
    }
    if (hull) for (uint32_t t = 0; t < out->hull_len && t < job.n; ++t) hull[t] = r0.hull[t];
    return ADK_OK;
}
}  // namespace

extern "C" {

int adk_multi_create(adk_multi** out, const int* devices, int ndev)
{
    if (!out) return ADK_ERR_ARG;
    *out = nullptr;
    int have = 0;
    if (cudaGetDeviceCount(&have) != cudaSuccess || have == 0) { cudaGetLastError(); return ADK_ERR_CUDA; }   // there is no CPU fallback
    if (ndev <= 0) { ndev = have; devices = nullptr; }
    if (ndev > have) return ADK_ERR_ARG;
    std::vector<int> devs(ndev);
    for (int i = 0; i < ndev; ++i)
    {
        devs[i] = devices ? devices[i] : i;
        if (devs[i] < 0 || devs[i] >= have) return ADK_ERR_ARG;
        for (int k = 0; k < i; ++k) if (devs[k] == devs[i]) return ADK_ERR_ARG;
    }
    {
        std::lock_guard<std::mutex> lk(g_nccl_mu);
        if (!g_nccl.load()) return ADK_ERR_CUDA;
    }
    int prev = -1; cudaGetDevice(&prev);
    adk_multi* m = new (std::nothrow) adk_multi();
    if (!m) return ADK_ERR_NOMEM;
    m->ranks.resize(ndev);
    std::vector<ncclComm_t> comms(ndev, nullptr);
    bool ok = g_nccl.CommInitAll(comms.data(), ndev, devs.data()) == 0;
    for (int i = 0; ok && i < ndev; ++i)
    {
        Rank& r = m->ranks[i];
        r.device = devs[i]; r.comm = comms[i];
        ok = adk_create(&r.ctx, devs[i]) == ADK_OK && cudaSetDevice(devs[i]) == cudaSuccess &&
             cudaStreamCreateWithFlags(&r.stream, cudaStreamNonBlocking) == cudaSuccess &&
             cudaMalloc(&r.d_send, 64) == cudaSuccess && cudaMalloc(&r.d_recv, sizeof(adk_partial) * ndev) == cudaSuccess &&
             cudaMalloc(&r.d_res, sizeof(uint32_t) * (kHdrWords + ADK_MAX_POINTS)) == cudaSuccess &&
             cudaMallocHost(&r.h_pin, sizeof(adk_partial) * ndev + sizeof(uint32_t) * (kHdrWords + ADK_MAX_POINTS)) == cudaSuccess;
    }
    if (prev >= 0) cudaSetDevice(prev);
    if (!ok) { for (int i = 0; i < ndev; ++i) if (!m->ranks[i].comm) m->ranks[i].comm = comms[i]; adk_multi_destroy(m); return ADK_ERR_CUDA; }
    for (int i = 0; i < ndev; ++i) m->ranks[i].worker = std::thread(worker_main, m, (uint32_t) i);
    *out = m;
    return ADK_OK;
}

void adk_multi_destroy(adk_multi* m)
{
    if (!m) return;
    {
        std::lock_guard<std::mutex> lk(m->mu);
        m->quit = true;
    }
    m->cv_go.notify_all();
    int prev = -1; cudaGetDevice(&prev);
    for (Rank& r : m->ranks)
    {
        if (r.worker.joinable()) r.worker.join();
        cudaSetDevice(r.device);
        if (r.stream) { cudaStreamSynchronize(r.stream); cudaStreamDestroy(r.stream); }
        if (r.comm && g_nccl.CommDestroy) g_nccl.CommDestroy(r.comm);
        if (r.d_send) cudaFree(r.d_send);
        if (r.d_recv) cudaFree(r.d_recv);
        if (r.d_res) cudaFree(r.d_res);
        if (r.h_pin) cudaFreeHost(r.h_pin);
        if (r.ctx) adk_destroy(r.ctx);
    }
    if (prev >= 0) cudaSetDevice(prev);
    delete m;
}

const char* adk_multi_last_error(const adk_multi* m) { return m ? m->err.c_str() : (g_nccl.err.empty() ? "adk_multi: no handle" : g_nccl.err.c_str()); }

int adk_multi_device_count(const adk_multi* m) { return m ? (int) m->ranks.size() : 0; }

int adk_multi_set_reuse(adk_multi* m, int enabled)
{
    if (!m) return ADK_ERR_ARG;
    for (Rank& r : m->ranks) adk_set_reuse(r.ctx, enabled);
    return ADK_OK;
}

int adk_multi_solve_area(adk_multi* m, const double* x, const double* y, uint32_t n, uint64_t kmax, double amax, uint32_t flags,
                         adk_result* out, uint32_t* hull)
{
    Job j; j.algo = ADK_ALGO_AREA; j.x = x; j.y = y; j.n = n; j.kmax = kmax; j.amax = amax; j.dmax = 0.0; j.flags = flags;
    return run_job(m, j, out, hull);
}

int adk_multi_solve_area_diameter(adk_multi* m, const double* x, const double* y, uint32_t n, uint64_t kmax, double amax, double dmax,
                                  uint32_t flags, adk_result* out, uint32_t* hull)
{
    Job j; j.algo = ADK_ALGO_AREA_DIAMETER; j.x = x; j.y = y; j.n = n; j.kmax = kmax; j.amax = amax; j.dmax = dmax; j.flags = flags;
    return run_job(m, j, out, hull);
}

int adk_multi_last_stats(const adk_multi* m, adk_multi_stats* out)
{
    if (!m || !out) return ADK_ERR_ARG;
    *out = m->last;
    return ADK_OK;
}

}  // extern "C"
