// adk_area.cu — path A: area-bounded largest convex subset (MT::EppsteinAlgorithm*, Eppstein.cpp:128-277) on sm_100a.
//
// B200-first reformulation (DESIGN.md "Path A"):
//   * Everything that does not depend on the level m is hoisted into a per-anchor PREP kernel: slope lists
//     (rotations of the per-centre direction-mod-pi order), triangle counts c(i,j,l), triangle areas, the
//     left-of-line mask.  The reference recomputes all of it K-2 times (Eppstein.cpp:199-201,209-214).
//   * Only "source" cells (left of pi->pj, area <= A) can ever change a row's running minimum, and only
//     positions right of the line are ever read back, so each row stores its prefix minimum at its sources
//     only (compact table T[level][row base + k]); the read T[m-c][l][j] becomes one precomputed gather index.
//   * Levels form a wavefront; one CTA owns an anchor, one warp owns a row, the prefix minimum is a warp
//     shuffle scan with a carried running value.  Levels live in a ring of (c_max + 1) slots; the sweep over m
//     stops as soon as no deeper level can be finite.
#include "adk_internal.h"
#include "adk_device.cuh"
#include "adk_area_plan.h"

#include <cooperative_groups.h>
#include <cstddef>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace adk
{

struct SrcRec { double area; uint32_t gidx; uint16_t c; uint16_t c3; };   // 16 B, one per source cell; c3 bit 15: the source is the
                                                                           // first element of its row's slope list (for the counter)
constexpr uint32_t kReachBuckets = 8;     // area buckets of an entry's readers (stop rule of the row sweep)
constexpr uint32_t kNoPrev = 0xffffffffu;   // nothing before the read position: level 2 reads 0, deeper levels read +inf

struct AnchorInfo
{
    int32_t pt;            // anchor point index
    int32_t a;             // capacity: |Above(pt)|
    uint32_t order;        // rank in the reference's anchor order (y ascending)
    uint32_t rows;         // rows actually present (j loop of Eppstein.cpp:196)
    uint64_t meta_off;     // first SrcRec / srcl entry of this anchor (row stride = a)
    uint64_t row_off;      // first rowpt / nsrc / rowbase entry
    uint64_t t_off;        // first T entry
    uint32_t U;            // total sources
    uint32_t cmax;         // largest c over sources
    uint32_t W;            // ring slots
    uint32_t best_m;       // 0 = nothing found
    double best_area;
    int32_t best_row;
    uint32_t levels_run;
    uint64_t list_cells;   // sum of slope-list lengths = reference cells per level
    uint64_t nonempty_rows;
    uint64_t required;     // prefix-minimum change points beyond the first list element (Eppstein.cpp:228-232)
    unsigned long long lvl[3];   // best area (order-preserving bits) of the level in flight and its two neighbours
    int32_t best_row_any;
    uint32_t defer_overflow;   // a row met more than 32 eps-collinear points beyond its vertex: outside the contract, reported as an error
    uint32_t pad2_[2];
    unsigned long long t0_ns, t1_ns;   // development: when the row pipeline started / finished this anchor (ADK_DEBUG)
};

// --------------------------------------------------------------------------------------------- PREP kernel

constexpr int kPrepThreads = 1024;   // one CTA per SM (64 registers): the heaviest anchor and the winner's single-anchor re-run are one CTA's critical path

__global__ void __launch_bounds__(kPrepThreads)
k_area_prep(const double* __restrict__ x, const double* __restrict__ y, uint32_t n, Tables T,
            const int32_t* __restrict__ cw, const int32_t* __restrict__ modpi, const uint16_t* __restrict__ rank,
            AnchorInfo* __restrict__ info, uint32_t first, uint32_t count,
            SrcRec* __restrict__ recs, uint16_t* __restrict__ srcl, int32_t* __restrict__ rowpt,
            uint32_t* __restrict__ nsrc, uint32_t* __restrict__ rowbase,
            uint16_t* __restrict__ qscratch, uint32_t qstride, uint64_t K, double amax, uint32_t* __restrict__ mcr)
{
    extern __shared__ uint16_t slot_of[];    // n entries: point -> row slot, 0xffff = not above the anchor
    __shared__ uint32_t s_start, s_rows, s_cmax, s_total;
    __shared__ unsigned long long s_cells, s_nonempty;
    __shared__ int32_t s_def[kPrepThreads / 32][32];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const uint32_t len = n - 1;
    uint16_t* Q = qscratch + (size_t) blockIdx.x * qstride;

    // anchors arrive heaviest first: deal them to the CTAs in snake order so that every CTA gets the same mix
    for (uint32_t round = 0; round * gridDim.x < count; ++round)
    {
        const uint32_t off = round * gridDim.x + ((round & 1u) ? gridDim.x - 1u - blockIdx.x : blockIdx.x);
        if (off >= count) continue;                  // uniform per CTA: no barrier is skipped by part of a block
        const uint32_t idx = first + off;
        AnchorInfo& A = info[idx];
        const int i = A.pt;
        const uint32_t acap = (uint32_t) A.a;
        const double ix = x[i], iy = y[i];
        const int32_t* ring_i = cw + (size_t) i * len;
        SrcRec* arec = recs + A.meta_off;
        uint16_t* asrc = srcl + A.meta_off;
        int32_t* arowpt = rowpt + A.row_off;
        uint32_t* ansrc = nsrc + A.row_off;
        uint32_t* abase = rowbase + A.row_off;

        if (threadIdx.x == 0) { s_start = len; s_rows = len; s_cmax = 0; s_total = 0; s_cells = 0; s_nonempty = 0; }
        for (uint32_t p = threadIdx.x; p < n; p += blockDim.x) slot_of[p] = 0xffff;
        __syncthreads();
        // Phase 0a: first clockwise position whose predecessor is below the anchor (Eppstein.cpp:32-42)
        for (uint32_t t = threadIdx.x; t + 1 < len; t += blockDim.x)
            if (y[ring_i[t]] < iy && y[ring_i[t + 1]] >= iy) atomicMin(&s_start, t + 1);
        __syncthreads();
        const uint32_t start = s_start == len ? 0u : s_start;
        // Phase 0b: the j loop runs while the clockwise neighbour stays above the anchor (Eppstein.cpp:196)
        for (uint32_t t = threadIdx.x; t < len; t += blockDim.x)
            if (y[ring_i[(start + t) % len]] < iy) atomicMin(&s_rows, t);
        __syncthreads();
        const uint32_t rows = min(s_rows, acap);
        for (uint32_t t = threadIdx.x; t < rows; t += blockDim.x)
        {
            const int p = ring_i[(start + t) % len];
            arowpt[t] = p;
            slot_of[p] = (uint16_t) t;
        }
        for (size_t t = threadIdx.x; t < (size_t) rows * rows; t += blockDim.x) Q[t] = 0xffff;
        __syncthreads();

        // Phase 1: one warp per row.  Walk the centre's direction-mod-pi ring starting just after the anchor.
        for (uint32_t jj = warp; jj < rows; jj += nwarps)
        {
            const int j = arowpt[jj];
            const double jx = x[j], jy = y[j];
            const int32_t* ring_j = modpi + (size_t) j * len;
            const uint32_t r0 = rank[(size_t) j * n + i];
            const int cij = T.C(i, j);
            uint32_t ns = 0, kept = 0, ndef = 0;
            SrcRec* rrec = arec + (size_t) jj * acap;
            uint16_t* rsrc = asrc + (size_t) jj * acap;
            // two chunks of 32 ring positions per step: the loads of both (ring -> slot -> coordinates -> count tables, four
            // dependent round trips) are in flight together, the lists are then extended in order
            struct Cand { bool keep, src, defer; int p; uint32_t sl; double area; uint32_t cfull, c3; };
            auto eval = [&](uint32_t t) -> Cand
            {
                Cand e{false, false, false, -1, 0xffffu, 0.0, 0u, 0u};
                if (t < len)
                {
                    e.p = ring_j[(r0 + t) % len];
                    e.sl = slot_of[e.p];
                    if (e.sl != 0xffff)
                    {
                        const double px = x[e.p], py = y[e.p];
                        const double s = signed_area(ix, iy, jx, jy, px, py);
                        if (near_zero(s)) e.defer = py > jy;              // Eppstein.cpp:76-82
                        else
                        {
                            e.keep = true;
                            if (s > 0.0)
                            {   // left of pi->pj: a candidate predecessor (Eppstein.cpp:209-214)
                                e.area = fabs(s);
                                const int cnt = cij + T.C(j, e.p) + points_in_triangle(T, i, ix, iy, j, jx, jy, e.p, px, py);
                                e.cfull = 1u + (uint32_t) cnt;
                                e.c3 = e.cfull + (uint32_t) T.C(e.p, i);  // the m == 3 extra term (Eppstein.cpp:25-28)
                                e.src = e.area <= amax && (uint64_t) e.cfull <= K;
                            }
                        }
                    }
                }
                return e;
            };
            auto commit = [&](const Cand& e)
            {
                const uint32_t km = __ballot_sync(0xffffffffu, e.keep), sm = __ballot_sync(0xffffffffu, e.src);
                const uint32_t dm = __ballot_sync(0xffffffffu, e.defer);
                const uint32_t lt = (1u << lane) - 1u;
                if (e.keep) Q[(size_t) jj * rows + e.sl] = (uint16_t) (ns + __popc(sm & lt));
                if (e.src)
                {
                    const uint32_t k = ns + __popc(sm & lt);
                    const uint32_t head = (k == 0 && kept == 0 && (km & lt) == 0u) ? 0x8000u : 0u;
                    rsrc[k] = (uint16_t) e.sl;
                    rrec[k] = SrcRec{e.area, kNoPrev, (uint16_t) e.cfull, (uint16_t) (e.c3 | head)};
                }
                if (e.defer)
                {
                    const uint32_t d = ndef + __popc(dm & lt);
                    if (d < 32) s_def[warp][d] = e.p;
                }
                if (ndef + __popc(dm) > 32u && lane == 0) A.defer_overflow = 1u;     // never truncate silently
                ns += __popc(sm); kept += __popc(km); ndef = min(32u, ndef + __popc(dm));
            };
#pragma unroll 1
            for (uint32_t base = 1; base < len; base += 64)
            {
                const Cand e0 = eval(base + lane), e1 = eval(base + 32u + lane);
                commit(e0);
                commit(e1);
            }
            __syncwarp();
            if (ndef && lane == 0)
            {   // points collinear with pi, pj beyond pj close the list, nearer first (Eppstein.cpp:78-81, Utils.cpp:301-306)
                for (uint32_t u = 1; u < ndef; ++u)
                {
                    const int p = s_def[warp][u]; const double dp = dist2(jx, jy, x[p], y[p]);
                    uint32_t v = u;
                    while (v > 0 && dist2(jx, jy, x[s_def[warp][v - 1]], y[s_def[warp][v - 1]]) > dp) { s_def[warp][v] = s_def[warp][v - 1]; --v; }
                    s_def[warp][v] = p;
                }
                for (uint32_t u = 0; u < ndef; ++u)
                {
                    const int p = s_def[warp][u];
                    const uint32_t sl = slot_of[p];
                    Q[(size_t) jj * rows + sl] = (uint16_t) ns;
                    const double area = fabs(signed_area(ix, iy, jx, jy, x[p], y[p]));
                    const uint32_t cfull = 1u + (uint32_t) (cij + T.C(j, p));   // PointsInTriangle = 0 for a collinear triple
                    if (area <= amax && (uint64_t) cfull <= K)
                    {
                        rsrc[ns] = (uint16_t) sl;
                        rrec[ns] = SrcRec{area, kNoPrev, (uint16_t) cfull, (uint16_t) (cfull | ((ns == 0 && kept == 0) ? 0x8000u : 0u))};
                        ++ns;
                    }
                    ++kept;
                }
            }
            ns = __shfl_sync(0xffffffffu, ns, 0); kept = __shfl_sync(0xffffffffu, kept, 0);
            if (lane == 0)
            {
                ansrc[jj] = ns;
                atomicAdd(&s_cells, (unsigned long long) kept);
                if (kept) atomicAdd(&s_nonempty, 1ull);
            }
        }
        __syncthreads();
        // Phase 2a: exclusive scan of the per-row source counts (one warp, chunked)
        if (warp == 0)
        {
            const uint32_t chunk = (rows + 31) / 32;
            const uint32_t lo = min(rows, lane * chunk), hi = min(rows, lo + chunk);
            uint32_t sum = 0;
            for (uint32_t r = lo; r < hi; ++r) sum += ansrc[r];
            uint32_t incl = sum;
            for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o); if ((int) lane >= o) incl += v; }
            uint32_t run = incl - sum;
            for (uint32_t r = lo; r < hi; ++r) { abase[r] = run; run += ansrc[r]; }
            if (lane == 31) s_total = incl;
        }
        __syncthreads();
        // Phase 2b: the read T[m-c][l][j] as an index into row l's compact prefix-minimum array.
        // Every table entry also learns WHO READS IT: a record r that reads entry e turns a finite value g of e at level m into a
        // finite value at level m + c_r only if g + area_r <= A.  Readers are bucketed by floor(8 area_r / A); amcr[e][b] becomes the
        // largest c among the readers of e in buckets <= b, so that the row sweep can bound the highest level a finite entry can
        // still feed by m + amcr[e][floor(8 (A - g) / A)] and stop long before "top level + largest c of the anchor".
        // Storage: 8 x 32-bit words per entry while the maxima are scattered (native atomicMax, nobody waits for it), packed to
        // 8 x 16 bits at the start of the same 32 bytes by the prefix pass below.
        uint32_t* amcr = mcr + A.meta_off * kReachBuckets;
        {
            uint4* z = reinterpret_cast<uint4*>(amcr);
            for (uint32_t t = threadIdx.x; t < 2u * s_total; t += blockDim.x) z[t] = make_uint4(0u, 0u, 0u, 0u);
        }
        __syncthreads();
        const double inv_bucket = (double) kReachBuckets / amax;
        uint32_t cm = 0;
        for (uint32_t jj = warp; jj < rows; jj += nwarps)
        {
            const uint32_t ns = ansrc[jj];
            SrcRec* rrec = arec + (size_t) jj * acap;
            const uint16_t* rsrc = asrc + (size_t) jj * acap;
            // four records per lane at a time: the chain record -> Q column -> row base is three dependent memory round trips
#pragma unroll 1
            for (uint32_t k0 = lane; k0 < ns; k0 += 128)
            {
                uint32_t lp[4], q[4], ab[4], cc[4]; double ar[4];
#pragma unroll
                for (uint32_t t = 0; t < 4; ++t) { const uint32_t k = k0 + 32u * t; lp[t] = k < ns ? rsrc[k] : 0u; }
#pragma unroll
                for (uint32_t t = 0; t < 4; ++t)
                {
                    const uint32_t k = k0 + 32u * t;
                    q[t] = k < ns ? Q[(size_t) lp[t] * rows + jj] : 0xffffu;
                    ab[t] = abase[lp[t]];
                    cc[t] = k < ns ? rrec[k].c : 0u;
                    ar[t] = k < ns ? rrec[k].area : 0.0;
                }
#pragma unroll
                for (uint32_t t = 0; t < 4; ++t)
                {
                    const uint32_t k = k0 + 32u * t;
                    if (k >= ns) continue;
                    const uint32_t g = (q[t] == 0xffff || q[t] == 0) ? kNoPrev : ab[t] + q[t] - 1;
                    rrec[k].gidx = g;
                    if (g != kNoPrev)
                    {
                        const int br = min((int) kReachBuckets - 1, max(0, (int) (ar[t] * inv_bucket)));
                        atomicMax(amcr + (size_t) g * kReachBuckets + br, cc[t]);
                    }
                    cm = max(cm, cc[t]);
                }
            }
        }
        __syncthreads();
        for (uint32_t t = threadIdx.x; t < s_total; t += blockDim.x)
        {   // prefix maximum over the buckets, packed to 16 bits in place (c <= ADK_MAX_POINTS)
            uint4* e = reinterpret_cast<uint4*>(amcr + (size_t) t * kReachBuckets);
            const uint4 lo = e[0], hi = e[1];
            uint32_t v[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
#pragma unroll
            for (int b = 1; b < 8; ++b) v[b] = max(v[b], v[b - 1]);
            e[0] = make_uint4(v[0] | (v[1] << 16), v[2] | (v[3] << 16), v[4] | (v[5] << 16), v[6] | (v[7] << 16));
        }
        atomicMax(&s_cmax, cm);
        __syncthreads();
        if (threadIdx.x == 0)
        {
            A.rows = rows; A.U = s_total; A.cmax = s_cmax; A.list_cells = s_cells; A.nonempty_rows = s_nonempty;
            A.best_m = 0; A.best_area = 0.0; A.best_row = -1; A.levels_run = 0; A.required = 0;
            A.lvl[0] = A.lvl[1] = A.lvl[2] = kInfBits; A.best_row_any = 0x7fffffff;
        }
        __syncthreads();
    }
}

// --------------------------------------------------------------------------------------------- DP kernel

constexpr int kDpThreads = 512;

__device__ __forceinline__ double warp_prefix_min(double v, uint32_t lane)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
        const double u = __shfl_up_sync(0xffffffffu, v, o);
        if ((int) lane >= o) v = fmin(v, u);
    }
    return v;
}

// One CTA per anchor (dynamic queue, heaviest anchor first), one warp per row, wavefront over levels.
// One anchor per thread-block CLUSTER: the rows of the anchor are dealt over the warps of all CTAs of the cluster and
// the levels are separated by a cluster barrier.  Anchor costs grow like a_i^3, so the heaviest anchor alone outweighs the
// average CTA's share of a whole solve: one CTA per anchor left half of the SM time idle behind the last heavy anchors
// (profiles/r1c_ncu_k_area_dp_n1000.txt); spreading an anchor over 8 SMs shortens that critical path 8 times.
__global__ void __launch_bounds__(kDpThreads)
k_area_dp(AnchorInfo* __restrict__ info, uint32_t first, uint32_t count, unsigned int* __restrict__ queue,
          const SrcRec* __restrict__ recs, const uint32_t* __restrict__ nsrc, const uint32_t* __restrict__ rowbase,
          double* __restrict__ Tbuf, uint64_t K, double amax, unsigned int* __restrict__ cluster_next)
{
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    const uint32_t crank = cluster.block_rank(), csize = cluster.num_blocks();
    const uint32_t cid = blockIdx.x / csize;
    __shared__ unsigned long long s_lvl, s_req;
    __shared__ int s_row;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const uint32_t gwarp = crank * nwarps + warp, gnwarps = csize * nwarps;
    for (;;)
    {
        if (crank == 0 && threadIdx.x == 0) cluster_next[cid] = atomicAdd(queue, 1u);
        __threadfence();
        cluster.sync();
        const uint32_t q = *reinterpret_cast<volatile unsigned int*>(cluster_next + cid);
        if (q >= count) break;
        AnchorInfo& A = info[first + q];
        const uint32_t rows = A.rows, U = A.U, W = A.W, acap = (uint32_t) A.a, cmax = A.cmax;
        const SrcRec* arec = recs + A.meta_off;
        const uint32_t* ansrc = nsrc + A.row_off;
        const uint32_t* abase = rowbase + A.row_off;
        double* T = Tbuf + A.t_off;
        volatile unsigned long long* g_lvl = A.lvl;          // best area of levels m, m+1, m+2 (mod 3), shared by the cluster
        uint32_t top = 2, levels = 0;
        double best_area = 0.0;
        if (threadIdx.x == 0) { s_lvl = kInfBits; s_row = 0x7fffffff; s_req = 0ull; }
        unsigned int req = 0;
        __syncthreads();
        if (U > 0)
        {
            for (uint64_t m = 3; m <= K; ++m)
            {
                if (crank == 0 && threadIdx.x == 0) A.lvl[(m + 1) % 3] = kInfBits;
                double* Tm = T + (size_t) (m % W) * U;
                for (uint32_t jj = gwarp; jj < rows; jj += gnwarps)
                {
                    const uint32_t ns = ansrc[jj];
                    if (ns == 0) continue;
                    const SrcRec* rrec = arec + (size_t) jj * acap;
                    double* out = Tm + abase[jj];
                    double carry = INFINITY;
                    for (uint32_t k0 = 0; k0 < ns; k0 += 32)
                    {
                        const uint32_t k = k0 + lane;
                        double val = INFINITY;
                        bool head = false;
                        if (k < ns)
                        {
                            const SrcRec r = rrec[k];
                            const uint64_t c = m == 3 ? (r.c3 & 0x7fffu) : r.c;
                            head = (r.c3 & 0x8000u) != 0;
                            if (c + 2 <= m)
                            {   // Eppstein.cpp:210-216 with T[0], T[1] = +inf and T[2] = 0
                                const uint64_t lvl = m - c;
                                double prev;
                                if (lvl == 2) prev = 0.0;
                                else prev = r.gidx == kNoPrev ? INFINITY : __ldcg(&T[(size_t) (lvl % W) * U + r.gidx]);
                                const double cur = __dadd_rn(prev, r.area);
                                if (cur <= amax) val = cur;
                            }
                        }
                        const bool head_finite = k == 0 && head && val < INFINITY;       // counted once as "first element", not as a change
                        val = fmin(warp_prefix_min(val, lane), carry);
                        double before = __shfl_up_sync(0xffffffffu, val, 1);
                        if (lane == 0) before = carry;
                        req += (unsigned int) __popc(__ballot_sync(0xffffffffu, k < ns && val < before))
                               - (unsigned int) __popc(__ballot_sync(0xffffffffu, head_finite));      // Eppstein.cpp:228-232
                        if (k < ns) __stcg(&out[k], val);                           // Eppstein.cpp:234
                        carry = __shfl_sync(0xffffffffu, val, 31);
                    }
                    if (lane == 0 && carry < INFINITY) atomicMin(&s_lvl, area_bits(carry));
                }
                __syncthreads();
                if (threadIdx.x == 0)
                {
                    if (s_lvl != kInfBits) atomicMin(const_cast<unsigned long long*>(&g_lvl[m % 3]), s_lvl);
                    s_lvl = kInfBits;
                    __threadfence();
                }
                cluster.sync();                          // level m is complete in every CTA of the cluster
                ++levels;
                const unsigned long long lb = g_lvl[m % 3];
                if (lb != kInfBits) { top = (uint32_t) m; best_area = bits_area(lb); }
                if (m >= (uint64_t) top + cmax) break;      // no deeper level can be finite
            }
            if (top > 2)
            {   // first row (in j order) that attains the best area at the best level (Eppstein.cpp:219-224)
                const double* Tt = T + (size_t) (top % W) * U;
                for (uint32_t jj = crank * blockDim.x + threadIdx.x; jj < rows; jj += csize * blockDim.x)
                {
                    const uint32_t ns = ansrc[jj];
                    if (ns && __ldcg(&Tt[abase[jj] + ns - 1]) == best_area) atomicMin(&s_row, (int) jj);
                }
            }
        }
        if (lane == 0 && req) atomicAdd(&s_req, (unsigned long long) req);
        __syncthreads();
        if (threadIdx.x == 0)
        {
            if (s_row != 0x7fffffff) atomicMin(&A.best_row_any, s_row);
            if (s_req) atomicAdd(reinterpret_cast<unsigned long long*>(&A.required), s_req);
            __threadfence();
        }
        cluster.sync();
        if (crank == 0 && threadIdx.x == 0)
        {
            A.levels_run = levels;
            if (top > 2) { A.best_m = top; A.best_area = best_area; A.best_row = *reinterpret_cast<volatile int*>(&A.best_row_any); }
        }
    }
}

// --------------------------------------------------------------------------------------------- DP kernel, row sweep
//
// The same recurrence, swept the other way round.  A source l of row j lies to the left of anchor->j, i.e. it comes EARLIER
// in the clockwise row order, so rows can be completed one after the other with ALL levels of a row handled together:
// lanes hold 32 consecutive levels (a band); per source a warp reads one contiguous 256-byte level vector of the
// predecessor entry (shifted by c), adds the triangle, folds it into the running minimum and writes one contiguous
// 256-byte vector.  Traffic per cell = 8 B read + 8 B written (the algorithmic 16 B); the level wavefront above re-reads
// the 16-byte source record at every level and pulls a 64-byte DRAM burst for every 8-byte gather (80 B per cell measured).
// Table layout: T[compact source index][ring of Wb bands][32 levels].
// One CTA per anchor, one row at a time: the warps take up to kSeg sources each, request all their predecessor vectors at once
// (one memory round trip per kSeg x warps sources; asynchronous copies into shared memory, see SrcPlan), form local prefix minima
// in place, exchange the segment minima through shared memory and write the final vectors.  Rows are strictly ordered inside the
// CTA, so no flags are needed; bands are swept one after the other, and after each band the stop rule applies.
#ifndef ADK_ROW_THREADS
#define ADK_ROW_THREADS 256
#endif
#ifndef ADK_ROW_SEG
#define ADK_ROW_SEG 32
#endif
#ifndef ADK_ROW_CTAS
#define ADK_ROW_CTAS 2
#endif
constexpr int kRowThreads = ADK_ROW_THREADS;
constexpr uint32_t kRowWarps = kRowThreads / 32;
constexpr uint32_t kBand = 32;
constexpr uint32_t kSeg = ADK_ROW_SEG;                 // sources per warp and round
constexpr uint32_t kRound = kRowWarps * kSeg;          // sources per CTA and round
// Staging of the predecessor vectors (round 2): the 32 levels a warp needs from one predecessor entry are 256 contiguous bytes of
// its table row, starting at any 8-byte boundary and possibly wrapping at the end of the ring.  They are brought into shared
// memory with 16-byte asynchronous copies (cp.async.cg: L2 -> shared memory, no register in between), 17 per source, so that a
// vector in flight costs 272 bytes of shared memory instead of two registers in every lane: the kernel drops from 124 to under
// 64 registers and the resident warps per SM double.
constexpr uint32_t kStageDoubles = 34;                 // per source: 32 levels + one element of slack at either end (16-byte pieces)
// Per source of a round: address of the first 16-byte piece, byte correction for the pieces that lie past the end of the ring,
// number of pieces (0 = nothing to fetch) and the first piece that is past the first band slot.
struct __align__(16) SrcPlan { unsigned long long first; int32_t wrap; uint16_t pieces, turn; };

__device__ __forceinline__ void async_copy16(void* smem_dst, const void* gmem_src)
{
    const uint32_t d = (uint32_t) __cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void async_copy_wait()
{
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

__device__ __forceinline__ uint32_t level_pos(uint32_t lvl, uint32_t Wb)      // lvl >= 3
{
    const uint32_t z = lvl - 3u;
    return ((z >> 5) % Wb) * kBand + (z & 31u);
}

// Segment of round `rd` of a row with ns sources that warp `warp` handles: the row is cut into ceil(ns / kRound) rounds
// of equal size, a round into kRowWarps equal segments of at most kSeg sources.
__device__ __forceinline__ void row_segment(uint32_t ns, uint32_t rd, uint32_t warp, uint32_t& base, uint32_t& cnt)
{
    const uint32_t rounds = (ns + kRound - 1) / kRound;
    const uint32_t per_round = rounds == 1 ? ns : (ns + rounds - 1) / rounds;
    const uint32_t seg = (per_round + kRowWarps - 1) / kRowWarps;
    const uint32_t end = min(ns, (rd + 1) * per_round);
    base = rd * per_round + warp * seg;
    cnt = base < end ? min(seg, end - base) : 0u;
}

// One anchor per thread-block CLUSTER (round 2).  Cell (row j, band b) only reads rows finished before j in bands <= b, so the
// bands of an anchor form a pipeline: CTA r of the cluster sweeps bands r, r + CS, r + 2 CS, ... and may work on row j of band b
// as soon as band b - 1 has finished row j (one row behind; a progress word per band in global memory, published after every row).
// With CS = 1 this is the one-CTA-per-anchor sweep of round 1; with CS > 1 the heaviest anchor's critical path (96-120 ms at
// n = 1000) is cut by the number of bands in flight, which is what anchors-over-8-GPUs needs.  The ring holds
// ceil(cmax / 32) + CS bands: the oldest band in flight still reads ceil(cmax / 32) bands back while the newest one writes.
// Bands past the last finite level may be started before the stop rule can know (at most CS - 1 of them): they produce +inf only.
__device__ void hull_diameter_dev(const double* x, const double* y, const int32_t* hull, uint32_t cnt, uint32_t* out);

// Local prefix minima of one warp's segment (see k_area_rows).  BAND0: the band that holds level 3, whose lane 0 uses the m == 3
// variant of c (Eppstein.cpp:25-28); the other bands do not pay for the choice.
template <bool BAND0>
__device__ __forceinline__ double prefix_pass(const SrcRec* st_rec, double* cell, uint32_t cnt, uint32_t m, bool active, double amax)
{
    const int mk = active ? (int) m : 0;
    const bool is3 = BAND0 && m == 3u;
    const uint4* rec4 = reinterpret_cast<const uint4*>(st_rec);          // area (x, y) | gidx (z) | c, c3 (w)
    double run = INFINITY;
#pragma unroll 1
    for (uint32_t u0 = 0; u0 < cnt; u0 += 4)
    {
        uint4 rw[4]; double v[4];
#pragma unroll
        for (uint32_t t = 0; t < 4; ++t) rw[t] = rec4[u0 + t];
#pragma unroll
        for (uint32_t t = 0; t < 4; ++t) v[t] = cell[t * kStageDoubles + (rw[t].w & 1u)];
        __syncwarp();                       // every lane has its elements before a neighbour overwrites them
#pragma unroll
        for (uint32_t t = 0; t < 4; ++t)
        {
            const uint32_t c = (BAND0 && is3) ? ((rw[t].w >> 16) & 0x7fffu) : (rw[t].w & 0xffffu);
            const int k = mk - (int) c;
            const bool real = k >= 3 && rw[t].z != kNoPrev;
            const int hi = real ? __double2hiint(v[t]) : (k == 2 ? 0 : 0x7ff00000);
            const int lw = real ? __double2loint(v[t]) : 0;
            const double cur = __dadd_rn(__hiloint2double(hi, lw), __hiloint2double((int) rw[t].y, (int) rw[t].x));
            v[t] = cur <= amax ? cur : INFINITY;
        }
#pragma unroll
        for (uint32_t t = 0; t < 4; ++t) { run = v[t] < run ? v[t] : run; cell[t * kStageDoubles + 1u] = run; }
        cell += 4u * kStageDoubles;
    }
    return run;
}

// The best anchor a cluster has met so far, with its polygon already walked back (see the end of k_area_rows).
struct KeepRec { uint32_t found, m; double area; uint32_t order, len, ok, diam[2]; int32_t pt, row, pad; };

constexpr uint32_t kMaxRing = 256;           // bands on the board: covers any cmax < 32 * (kMaxRing - 8) = 7 936 > ADK_MAX_POINTS
struct BandBoard
{
    unsigned long long prog[kMaxRing];      // (band + 1) << 32 | rows finished, per ring slot
    unsigned int top[kMaxRing];             // highest finite level of the band in that slot (0 = none)
    double area[kMaxRing];                  // its best area
    unsigned int q, levels, reach, pad;     // the anchor the cluster works on; levels executed (all bands); highest level a finite entry can feed
    KeepRec keep;                           // this cluster's best anchor so far; its hull is at keep_hull[cluster * hull_stride ..]
};

__global__ void __launch_bounds__(kRowThreads, ADK_ROW_CTAS)
k_area_rows(AnchorInfo* __restrict__ info, uint32_t first, uint32_t count, unsigned int* __restrict__ queue,
            const SrcRec* __restrict__ recs, const uint32_t* __restrict__ nsrc,
            const uint32_t* __restrict__ rowbase, double* __restrict__ Tbuf, const uint64_t* __restrict__ slot_off, uint64_t K, double amax,
            uint32_t cap_rows, BandBoard* __restrict__ boards,
            const double* __restrict__ x, const double* __restrict__ y, const uint16_t* __restrict__ srcl,
            const int32_t* __restrict__ rowpt, int32_t* __restrict__ keep_hull, uint32_t hull_stride,
            const uint32_t* __restrict__ mcr)
{
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    const uint32_t crank = cluster.block_rank(), CS = cluster.num_blocks();
    const uint32_t cid = blockIdx.x / CS;
    __shared__ __align__(16) SrcRec s_rec[kRowWarps][kSeg];
    __shared__ __align__(16) SrcPlan s_plan[kRowWarps][kSeg];
    __shared__ double s_tot[2][kRowWarps][32];
    __shared__ unsigned long long s_best[kBand];
    __shared__ long long s_req;
    __shared__ int s_row, s_go;
    __shared__ __align__(16) uint16_t s_mc[kRowWarps][kReachBuckets];    // per warp and round: largest reader c per area bucket
    __shared__ unsigned int s_reach;
    // [cap_rows] source counts, [cap_rows] row bases (anchors with more rows read them from global memory), then the staging
    // area: kStageDoubles doubles per source of the round, kSeg sources per warp
    extern __shared__ __align__(16) uint32_t s_rowtab[];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    SrcRec* st_rec = s_rec[warp];
    SrcPlan* st_plan = s_plan[warp];
    double* stage = reinterpret_cast<double*>(s_rowtab + 2u * cap_rows) + (size_t) warp * kSeg * kStageDoubles;
    BandBoard& B = boards[cid];
    volatile unsigned long long* vprog = B.prog;
    if (crank == 0 && threadIdx.x == 0) { B.keep.found = 0; B.keep.m = 0; B.keep.area = 0.0; B.keep.ok = 0; B.keep.len = 0; }
    for (bool first_round = true;; first_round = false)
    {
        __syncthreads();
        if (crank == 0 && threadIdx.x == 0)
        {
            // the first anchor of cluster k is anchor k (its slot is sized for it, see area_run); then the queue behind the first ones
            B.q = first_round ? cid : gridDim.x / CS + atomicAdd(queue, 1u); B.levels = 0; B.reach = 0;
            for (uint32_t t = 0; t < kMaxRing; ++t) { B.prog[t] = 0ull; B.top[t] = 0u; }
            __threadfence();
        }
        cluster.sync();
        const uint32_t q = *reinterpret_cast<volatile unsigned int*>(&B.q);
        if (q >= count) break;
        AnchorInfo& A = info[first + q];
        const uint32_t rows = A.rows, U = A.U, W = A.W, acap = (uint32_t) A.a, cmax = A.cmax;
        const uint32_t Wb = W / kBand, W8 = W * 8u;     // W8: bytes per table row
        const double inv_bucket = (double) kReachBuckets / amax;
        // bands this anchor can fill at all: a polygon above it has at most a + 2 points (the host sizes the ring with the same
        // bound, so a band beyond it would wrap onto a band that is still being written)
        const uint64_t top_level = K < (uint64_t) acap + 2 ? K : (uint64_t) acap + 2;
        const uint32_t nb_cap = (uint32_t) (top_level >= 3 ? (top_level - 2 + kBand - 1) / kBand : 1);
        const SrcRec* arec = recs + A.meta_off;
        const uint4* amcr = reinterpret_cast<const uint4*>(mcr + A.meta_off * kReachBuckets);      // 32 bytes per entry, the first 16 hold 8 x uint16
        const uint32_t* ansrc = nsrc + A.row_off;
        const uint32_t* abase = rowbase + A.row_off;
        if (rows <= cap_rows)
        {
            for (uint32_t t = threadIdx.x; t < rows; t += blockDim.x) { s_rowtab[t] = ansrc[t]; s_rowtab[cap_rows + t] = abase[t]; }
            ansrc = s_rowtab; abase = s_rowtab + cap_rows;
        }
        double* T = Tbuf + slot_off[cid];                    // the cluster's own table slot, reused for every anchor it takes
        if (threadIdx.x == 0)
        {
            s_req = 0; s_row = 0x7fffffff; s_reach = 0;
            if (crank == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(A.t0_ns));
        }
        __syncthreads();
        long long req = 0;
        if (U > 0)
        {
            uint32_t par = 0;                   // which s_tot buffer the next chunk uses
            for (uint32_t b = crank;; b += CS)
            {
                const uint32_t lo = 3u + kBand * b;
                if ((uint64_t) lo > K || b >= nb_cap) break;
                if (b >= CS)
                {   // Bands finish in order, so every band up to b - CS (this CTA's previous one) is finished or about to be: had the
                    // one-band-at-a-time sweep stopped by then (no deeper level can be finite, Eppstein.cpp:192 made adaptive), this
                    // band is not needed.
                    if (threadIdx.x == 0)
                    {
                        const uint32_t t0 = b + CS > kMaxRing ? b + CS - kMaxRing : 0u;      // bands whose board slot is still theirs
                        for (uint32_t t = t0; t + CS <= b; ++t)
                        {   // every band up to b - CS has published its share of `reach`
                            const uint32_t sl = t % kMaxRing;
                            while (vprog[sl] != (((unsigned long long) (t + 1) << 32) | 0xffffffffull)) __nanosleep(64);
                            __threadfence();
                        }
                        const uint64_t lo_done = 3u + kBand * (b - CS);
                        const uint64_t hi_done = K < lo_done + kBand - 1 ? K : lo_done + kBand - 1;
                        // Nothing beyond hi_done can be finite once hi_done has passed every level a finite entry can feed (its level
                        // + the largest c among its readers) and every level the constant level 2 can feed (2 + c).  "top finite level + cmax", the
                        // rule of round 1, is the same bound with the anchor-wide cmax in place of the per-entry one.
                        const uint64_t reach = max((uint64_t) *reinterpret_cast<volatile unsigned int*>(&B.reach), (uint64_t) cmax + 2u);
                        s_go = (hi_done >= K || hi_done >= reach) ? 0 : 1;
                    }
                    __syncthreads();
                    const int go = s_go;
                    __syncthreads();
                    if (!go) break;
                }
                const uint32_t m = lo + lane;
                const bool active = (uint64_t) m <= K;
                const uint32_t bm = b % Wb;
                const uint32_t mypos = bm * kBand + lane;
                const uint32_t bslot = b % kMaxRing, pslot = (b + kMaxRing - 1) % kMaxRing;
                double band_best = INFINITY;    // warp 0: best final value of level m over the rows
                // A row's sources are cut into rounds of at most kRound, and a round evenly into one segment per warp.
                // The source records of a round are requested one round ahead (they depend on nothing).
                uint32_t nj = 0, nr = 0;
                while (nj < rows && ansrc[nj] == 0) ++nj;
                // kept as raw words until the round that uses them: unpacking right behind the loads would wait for them
                const uint4* arec4 = reinterpret_cast<const uint4*>(arec);                 // area (x, y) | gidx (z) | c, c3 (w)
                uint4 nxt = make_uint4(0u, 0u, kNoPrev, 0u);
                uint4 nxt_mc = make_uint4(0u, 0u, 0u, 0u);   // the reader summary of the entry this lane's record produces (see k_area_prep)
                uint32_t reach_l = 0;           // highest level a finite entry written by this lane (level m) can feed
                {
                    uint32_t nb = 0, nc = 0;
                    if (nj < rows) row_segment(ansrc[nj], nr, warp, nb, nc);
                    if (lane < nc) { nxt = arec4[(size_t) nj * acap + nb + lane]; nxt_mc = amcr[2u * (abase[nj] + nb + lane)]; }
                }
                for (uint32_t jj = 0; jj < rows; ++jj)
                {
                    const uint32_t ns = ansrc[jj];
                    if (ns == 0) continue;
                    if (b > 0 && CS > 1)
                    {   // One row behind: band b - 1 must have finished row jj itself.  Then, by induction over the bands, every
                        // earlier band has finished row jj as well — a source of this row reads rows before jj up to
                        // ceil(cmax / 32) bands back, and "band b - 1 finished the rows before jj" alone would leave band b - 2 one
                        // row short of that.
                        if (lane == 0)
                        {
                            for (;;)
                            {
                                const unsigned long long pv_ = vprog[pslot];
                                if ((pv_ >> 32) == (unsigned long long) b && (uint32_t) pv_ >= jj + 1u) break;
                                __nanosleep(64);
                            }
                            __threadfence();
                        }
                        __syncwarp();
                    }
                    const uint32_t obase = abase[jj];
                    const uint32_t rounds = (ns + kRound - 1) / kRound;
                    double carry = INFINITY;
                    for (uint32_t rd = 0; rd < rounds; ++rd)
                    {
                        uint32_t base, cnt;
                        row_segment(ns, rd, warp, base, cnt);
                        __syncwarp();
                        {   // the largest reader c per area bucket over the entries this warp produces in this round (see the stop rule below)
                            const bool mine_ = lane < cnt;
                            const uint32_t q[4] = {mine_ ? nxt_mc.x : 0u, mine_ ? nxt_mc.y : 0u, mine_ ? nxt_mc.z : 0u, mine_ ? nxt_mc.w : 0u};
                            uint32_t packed[4];
#pragma unroll
                            for (uint32_t h = 0; h < 4; ++h)
                                packed[h] = __reduce_max_sync(0xffffffffu, q[h] & 0xffffu) | (__reduce_max_sync(0xffffffffu, q[h] >> 16) << 16);
                            if (lane == 0) *reinterpret_cast<uint4*>(s_mc[warp]) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
                        }
                        if (lane < cnt)
                        {
                            *reinterpret_cast<uint4*>(&st_rec[lane]) = nxt;
                            // Where the 32 levels m - c (m = lo .. lo + 31) of this source's predecessor entry lie: level lo - c sits at
                            // position p of band slot s0 = bm + floor(-c / 32), the window runs on into slot s0 + 1 (both modulo the
                            // ring).  Lane L of the band reads window element p + L; the window is fetched as 16-byte pieces
                            // starting at element p & ~1, so lane L finds its element at index (p & 1) + L of the staged copy.  A
                            // source none of whose levels exists yet (c + 3 > lo + 31, or past K) is not fetched at all.
                            const uint32_t c = nxt.w & 0xffffu, p = (0u - c) & 31u, p2 = p & ~1u;
                            const uint64_t mtop = K < (uint64_t) lo + 31u ? K : (uint64_t) lo + 31u;
                            const bool fetch = nxt.z != kNoPrev && (uint64_t) c + 3u <= mtop;
                            int s0 = (int) bm + ((-(int) c) >> 5);
                            if (s0 < 0) s0 += (int) Wb;
                            const int s1 = s0 + 1 == (int) Wb ? 0 : s0 + 1;
                            SrcPlan pl;
                            pl.first = (unsigned long long) __cvta_generic_to_global(T) + (uint64_t) nxt.z * W8 + (uint64_t) s0 * (kBand * 8u) + p2 * 8u;
                            pl.wrap = (s1 - s0 - 1) * (int) (kBand * 8u);
                            pl.pieces = (uint16_t) (fetch ? 16u + (p & 1u) : 0u);
                            pl.turn = (uint16_t) ((kBand - p2) >> 1);
                            st_plan[lane] = pl;
                        }
                        else if (lane < kSeg)
                        {   // the loops below run in groups of four sources: the padding fetches nothing and adds +inf
                            st_rec[lane] = SrcRec{INFINITY, kNoPrev, 0, 0};
                            st_plan[lane] = SrcPlan{0ull, 0, 0, 0};
                        }
                        __syncwarp();
                        if (++nr >= rounds) { nr = 0; ++nj; while (nj < rows && ansrc[nj] == 0) ++nj; }
                        {
                            uint32_t nb = 0, nc = 0;
                            if (nj < rows) row_segment(ansrc[nj], nr, warp, nb, nc);
                            if (lane < nc) { nxt = arec4[(size_t) nj * acap + nb + lane]; nxt_mc = amcr[2u * (abase[nj] + nb + lane)]; }
                        }
                        // every predecessor vector of the segment is on its way before any of them is used
                        {
                            char* dst = reinterpret_cast<char*>(stage) + lane * 16u;
                            const uint4* plan = reinterpret_cast<const uint4*>(st_plan);
#pragma unroll 1
                            for (uint32_t u0 = 0; u0 < cnt; u0 += 4)
                            {
                                uint4 pw[4];
#pragma unroll
                                for (uint32_t t = 0; t < 4; ++t) pw[t] = plan[u0 + t];             // first (x, y) | wrap (z) | pieces, turn (w)
#pragma unroll
                                for (uint32_t t = 0; t < 4; ++t)
                                {
                                    const long long off = (long long) (int) (lane * 16u) + (lane >= (pw[t].w >> 16) ? (int) pw[t].z : 0);
                                    const unsigned long long first = ((unsigned long long) pw[t].y << 32) | pw[t].x;
                                    if (lane < (pw[t].w & 0xffffu)) async_copy16(dst + t * (kStageDoubles * 8u), reinterpret_cast<const char*>(first + off));
                                }
                                dst += 4u * kStageDoubles * 8u;
                            }
                        }
                        async_copy_wait();
                        __syncwarp();
                        // local prefix minima: lane L takes its element (index parity + L of the staged copy) and leaves the running
                        // minimum of the segment up to u at index L + 1, whatever the parity (the loads of a group of four all come
                        // before its stores).  Branch-free: level m - c is a real table entry (>= 3), the constant level 2 (T[2] = 0)
                        // or nothing (+inf), Eppstein.cpp:210-216; lanes past K behave like level 0.
                        const double run = b == 0 ? prefix_pass<true>(st_rec, stage + lane, cnt, m, active, amax)
                                                  : prefix_pass<false>(st_rec, stage + lane, cnt, m, active, amax);
                        // the first element of the whole list is no change point if it is finite (its running minimum is its own value)
                        bool head_finite = false;
                        if (base == 0 && cnt) head_finite = (st_rec[0].c3 & 0x8000u) && stage[lane + 1u] < INFINITY;
                        s_tot[par][warp][lane] = run;
                        __syncthreads();
                        double cin = carry, all = carry;
#pragma unroll
                        for (uint32_t w = 0; w < kRowWarps; ++w)
                        {
                            const double t = s_tot[par][w][lane];
                            if (w < warp) cin = t < cin ? t : cin;
                            all = t < all ? t : all;
                        }
                        par ^= 1u;
                        double before = cin;
                        if (head_finite) --req;                                          // the list's first element is no change point
                        char* outp = reinterpret_cast<char*>(T + mypos) + (uint64_t) (obase + base) * W8;
                        uint32_t changes = 0;
                        {
                            const double* cell = stage + lane + 1u;
#pragma unroll 4
                            for (uint32_t u = 0; u < cnt; ++u)
                            {
                                const double mine = *cell;
                                const double g = mine < cin ? mine : cin;
                                changes += g < before ? 1u : 0u;                             // Eppstein.cpp:228-232
                                before = g;
                                if (active) __stcg(reinterpret_cast<double*>(outp), g);      // Eppstein.cpp:234
                                outp += W8;
                                cell += kStageDoubles;
                            }
                        }
                        if (cnt && before < INFINITY)
                        {   // Stop rule: the highest level a finite entry of level m can feed.  `before` is the smallest value this lane
                            // wrote in the round; readers that still fit under the area bound have area <= A - g (inflated: never stop
                            // early), and s_mc holds, per area bucket, the largest c among the readers of the round's entries.  One bound per
                            // lane and round instead of one per entry: never below the per-entry bound, and a few instructions per round.
                            const int bs = min((int) kReachBuckets - 1, max(0, (int) (__dadd_rn(__dmul_rn(__dsub_rn(amax, before), 1.000001), 1e-12) * inv_bucket)));
                            reach_l = max(reach_l, m + (uint32_t) s_mc[warp][bs]);
                        }
                        req += changes;
                        carry = all;
                    }
                    band_best = fmin(band_best, carry);
                    __syncthreads();            // the row is complete before any warp reads it as a source of the next rows
                    if (CS > 1 && threadIdx.x == 0) { __threadfence(); vprog[bslot] = ((unsigned long long) (b + 1) << 32) | (jj + 1); }
                }
                if (warp == 0) s_best[lane] = area_bits(band_best);
                reach_l = __reduce_max_sync(0xffffffffu, reach_l);
                if (lane == 0 && reach_l) atomicMax(&s_reach, reach_l);
                __syncthreads();
                if (threadIdx.x == 0)
                {
                    if (s_reach) atomicMax(&B.reach, s_reach);
                    unsigned int tp = 0; double ar = 0.0;
                    for (uint32_t l = 0; l < kBand; ++l)
                        if (s_best[l] != kInfBits) { tp = lo + l; ar = bits_area(s_best[l]); }
                    const uint64_t hi = K < (uint64_t) lo + kBand - 1 ? K : (uint64_t) lo + kBand - 1;
                    B.top[bslot] = tp; B.area[bslot] = ar;
                    atomicAdd(&B.levels, (unsigned int) (hi - lo + 1));
                    __threadfence();
                    vprog[bslot] = ((unsigned long long) (b + 1) << 32) | 0xffffffffull;      // the whole band is finished
                }
                __syncthreads();
            }
        }
        for (int o = 16; o; o >>= 1) req += __shfl_xor_sync(0xffffffffu, req, o);
        if (lane == 0 && req) atomicAdd(reinterpret_cast<unsigned long long*>(&s_req), (unsigned long long) req);
        __syncthreads();
        if (threadIdx.x == 0 && s_req) { atomicAdd(reinterpret_cast<unsigned long long*>(&A.required), (unsigned long long) s_req); __threadfence(); }
        cluster.sync();                         // every band of the anchor is finished
        // The answer of the anchor: the highest finite level over the bands that the one-band-at-a-time sweep would have run (bands
        // started beyond its stopping point hold +inf only), and the first row (in j order) that attains the best area there
        // (Eppstein.cpp:219-224).  The board keeps the last kMaxRing bands; a finite level is never more than cmax levels from the end.
        uint32_t top = 2; double best_area = 0.0;
        uint32_t bands_ran = 0;                 // bands started for this anchor: no more than the ring holds = nothing was overwritten
        if (U > 0)
        {
            for (uint32_t b = 0; (uint64_t) (3u + kBand * b) <= K && b < nb_cap; ++b)
            {
                const unsigned long long tag = vprog[b % kMaxRing] >> 32;
                if (tag < (unsigned long long) (b + 1)) break;                 // this band was never started: the sweep had stopped
                bands_ran = b + 1;
                if (tag > (unsigned long long) (b + 1)) continue;              // its slot belongs to a later band by now
                const unsigned int tp = *reinterpret_cast<volatile unsigned int*>(&B.top[b % kMaxRing]);
                if (tp) { top = tp; best_area = *reinterpret_cast<volatile double*>(&B.area[b % kMaxRing]); }
            }
            if (top > 2)
            {
                const uint32_t tp = level_pos(top, Wb);
                for (uint32_t jj = crank * blockDim.x + threadIdx.x; jj < rows; jj += CS * blockDim.x)
                {
                    const uint32_t ns = ansrc[jj];
                    if (ns && __ldcg(&T[(size_t) (abase[jj] + ns - 1) * W + tp]) == best_area) atomicMin(&s_row, (int) jj);
                }
            }
        }
        __syncthreads();
        if (threadIdx.x == 0 && s_row != 0x7fffffff) { atomicMin(&A.best_row_any, s_row); __threadfence(); }
        cluster.sync();
        if (crank == 0 && threadIdx.x == 0)
        {
            A.levels_run = *reinterpret_cast<volatile unsigned int*>(&B.levels);
            if (top > 2) { A.best_m = top; A.best_area = best_area; A.best_row = *reinterpret_cast<volatile int*>(&A.best_row_any); }
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(A.t1_ns));
        }
        // The anchor's table is still complete in the cluster's slot (when the ring did not wrap): if this anchor is the cluster's
        // best so far, walk its polygon back NOW (Eppstein.cpp:240-270) instead of re-running the winner after the solve.  A cluster
        // takes its anchors in the reference's order, so "strictly better" keeps the first of equals, as the reference does; the
        // solve's winner beats every anchor its cluster met before it and is beaten by none after it, so its polygon is the one
        // its cluster holds at the end.  One warp; the other CTAs wait at the next cluster barrier, nothing overwrites the table.
        if (crank == 0 && warp == 0 && top > 2 && bands_ran <= Wb &&
            (!B.keep.found || top > B.keep.m || (top == B.keep.m && best_area < B.keep.area)))
        {
            const uint16_t* asrc = srcl + A.meta_off;
            const int32_t* arowpt = rowpt + A.row_off;
            int32_t* hull = keep_hull + (size_t) cid * hull_stride;
            uint32_t m = top, h = 0, ok = 1;
            int row = *reinterpret_cast<volatile int*>(&A.best_row_any);
            const int row0 = row;
            if (lane == 0) hull[0] = A.pt;
            h = 1;
            int last = A.pt;
            const uint32_t hcap = rows + 2;          // a polygon has at most rows + 1 vertices (exactly degenerate inputs must not run away)
            while (m > 2)
            {
                if (h >= hcap || row < 0 || (uint32_t) row >= rows) { ok = 0; break; }
                const int j = arowpt[row];
                const uint32_t ns = ansrc[row], ob = abase[row];
                const uint32_t tp = level_pos(m, Wb);
                const double lx = x[last], ly = y[last], jx = x[j], jy = y[j];
                int pick = -1;
                for (int hi = (int) ns - 1; hi >= 0 && pick < 0; hi -= 32)
                {   // the last position where the running minimum changes and the turn at j stays convex
                    const int k = hi - (int) lane;
                    bool cond = false;
                    if (k >= 0)
                    {
                        const double v = __ldcg(&T[(size_t) (ob + k) * W + tp]);
                        const double vp = k ? __ldcg(&T[(size_t) (ob + k - 1) * W + tp]) : INFINITY;
                        if (v != vp)
                        {
                            const int l = arowpt[asrc[(size_t) row * acap + k]];
                            cond = orientation(lx, ly, jx, jy, x[l], y[l]) >= 0;
                        }
                    }
                    const uint32_t bm = __ballot_sync(0xffffffffu, cond);
                    if (bm) pick = hi - (__ffs((int) bm) - 1);
                }
                if (pick < 0) { ok = 0; break; }
                const SrcRec r = arec[(size_t) row * acap + pick];
                const uint32_t c = m == 3 ? (uint32_t) (r.c3 & 0x7fffu) : (uint32_t) r.c;
                if (c > m) { ok = 0; break; }
                m -= c;
                if (lane == 0) hull[h] = j;
                ++h;
                last = j;
                row = (int) asrc[(size_t) row * acap + pick];
            }
            if (ok && (row < 0 || (uint32_t) row >= rows)) ok = 0;
            if (ok) { if (lane == 0) hull[h] = arowpt[row]; ++h; }
            __syncwarp();
            if (lane == 0)
            {
                KeepRec kr;
                kr.found = 1; kr.m = top; kr.area = best_area; kr.order = A.order; kr.len = h; kr.ok = ok; kr.pt = A.pt; kr.row = row0; kr.pad = 0;
                kr.diam[0] = kr.diam[1] = 0;
                if (ok) { __threadfence(); hull_diameter_dev(x, y, hull, h, kr.diam); }
                B.keep = kr;
            }
            __syncwarp();
        }
    }
}

struct AreaBest { int32_t found; uint32_t m; double area; uint32_t order; int32_t row; int32_t pt; int32_t jpt;
                  uint64_t cells_nominal, cells_executed, required; };

// Lexicographic best over anchors: m desc, area asc, reference anchor order asc (Eppstein.cpp:219-224).
__global__ void k_area_reduce(const AnchorInfo* __restrict__ info, uint32_t count, const int32_t* __restrict__ rowpt,
                              uint64_t K, AreaBest* __restrict__ out)
{
    __shared__ unsigned long long s_cn, s_ce, s_rq;
    __shared__ uint32_t s_m; __shared__ unsigned long long s_area; __shared__ uint32_t s_order;
    if (threadIdx.x == 0) { s_cn = 0; s_ce = 0; s_rq = 0; s_m = 0; s_area = kInfBits; s_order = 0xffffffffu; }
    __syncthreads();
    unsigned long long cn = 0, ce = 0, rq = 0; uint32_t bm = 0;
    for (uint32_t t = threadIdx.x; t < count; t += blockDim.x)
    {
        cn += info[t].list_cells * (K - 2);
        ce += (unsigned long long) info[t].levels_run * info[t].U;
        rq += info[t].required + info[t].nonempty_rows * (K - 2);      // every non-empty list counts once per level (:228-232)
        bm = max(bm, info[t].best_m);
    }
    atomicAdd(&s_cn, cn); atomicAdd(&s_ce, ce); atomicAdd(&s_rq, rq); atomicMax(&s_m, bm);
    __syncthreads();
    const uint32_t m = s_m;
    if (m > 0)
    {
        for (uint32_t t = threadIdx.x; t < count; t += blockDim.x)
            if (info[t].best_m == m) atomicMin(&s_area, area_bits(info[t].best_area));
        __syncthreads();
        for (uint32_t t = threadIdx.x; t < count; t += blockDim.x)
            if (info[t].best_m == m && area_bits(info[t].best_area) == s_area) atomicMin(&s_order, info[t].order);
        __syncthreads();
    }
    for (uint32_t t = threadIdx.x; t < count; t += blockDim.x)
        if (m > 0 && info[t].best_m == m && area_bits(info[t].best_area) == s_area && info[t].order == s_order)
        {
            out->found = 1; out->m = m; out->area = info[t].best_area; out->order = info[t].order;
            out->row = info[t].best_row; out->pt = info[t].pt; out->jpt = rowpt[info[t].row_off + info[t].best_row];
        }
    if (threadIdx.x == 0)
    {
        if (m == 0) { out->found = 0; out->m = 0; out->area = 0.0; out->order = 0; out->row = -1; out->pt = -1; out->jpt = -1; }
        out->cells_nominal = s_cn; out->cells_executed = s_ce; out->required = s_rq;
    }
}

// --------------------------------------------------------------------------------------------- reconstruction

// Rotating calipers over the hull cycle (Utils.h:57-109) feeding the farthest-pair search of Utils.cpp:318-338.
__device__ void hull_diameter_dev(const double* x, const double* y, const int32_t* hull, uint32_t cnt, uint32_t* out)
{
    double best = 0.0; uint32_t ba = 0, bb = 0;
#define ADK_HA(p, q, r) signed_area(x[hull[p]], y[hull[p]], x[hull[q]], y[hull[q]], x[hull[r]], y[hull[r]])
#define ADK_NX(v) (((v) + 1) % cnt)
#define ADK_VISIT(p, q) { const double d = dist2(x[hull[p]], y[hull[p]], x[hull[q]], y[hull[q]]); if (d > best) { best = d; ba = (p); bb = (q); } }
    // The walk below terminates on a strictly convex cycle; on a degenerate one (collinear vertices: exactly equal areas in
    // binary64, where the reference's 80-bit arithmetic happens to break the ties) it can circle forever, so it is
    // given a step budget and falls back to the plain farthest pair of the cycle.
    uint32_t budget = 8 * cnt + 32;
    bool gave_up = false;
    if (cnt > 2)
    {
        uint32_t pi = cnt - 1, qi = 0;
        while (ADK_HA(pi, ADK_NX(pi), ADK_NX(qi)) > ADK_HA(pi, ADK_NX(pi), qi)) { qi = ADK_NX(qi); if (--budget == 0) { gave_up = true; break; } }
        const uint32_t q0 = qi;
        while (qi != 0 && !gave_up)
        {
            if (--budget == 0) { gave_up = true; break; }
            pi = ADK_NX(pi);
            ADK_VISIT(pi, qi)
            while (ADK_HA(pi, ADK_NX(pi), ADK_NX(qi)) > ADK_HA(pi, ADK_NX(pi), qi))
            {
                if (--budget == 0) { gave_up = true; break; }
                qi = ADK_NX(qi);
                if (pi != q0 || qi != 0) ADK_VISIT(pi, qi)
                else break;
            }
            if (gave_up) break;
            const double diff = __dsub_rn(ADK_HA(pi, ADK_NX(pi), ADK_NX(qi)), ADK_HA(pi, ADK_NX(pi), qi));
            if (near_zero(diff))
            {
                if (pi != q0 || qi != cnt - 1) { const uint32_t qn = ADK_NX(qi); ADK_VISIT(pi, qn) }
                else break;
            }
        }
    }
    if (gave_up)
    {
        best = 0.0; ba = 0; bb = 0;
        for (uint32_t a = 0; a < cnt; ++a) for (uint32_t b = a + 1; b < cnt; ++b) ADK_VISIT(a, b)
    }
    out[0] = (uint32_t) hull[ba]; out[1] = (uint32_t) hull[bb];
#undef ADK_HA
#undef ADK_NX
#undef ADK_VISIT
}

struct AreaHull { uint32_t len; uint32_t diam[2]; uint32_t ok; };

// Walk back from (m*, anchor, row*) over the anchor's full table (Eppstein.cpp:240-270).  One thread: <= k steps.
__global__ void k_area_backtrack(const double* __restrict__ x, const double* __restrict__ y,
                                 const AnchorInfo* __restrict__ info, const SrcRec* __restrict__ recs,
                                 const uint16_t* __restrict__ srcl, const int32_t* __restrict__ rowpt,
                                 const uint32_t* __restrict__ nsrc, const uint32_t* __restrict__ rowbase,
                                 const double* __restrict__ Tbuf, uint32_t best_m, int32_t best_row,
                                 int32_t* __restrict__ hull, AreaHull* __restrict__ out)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const AnchorInfo& A = info[0];
    const uint32_t U = A.U, W = A.W, acap = (uint32_t) A.a;
    const SrcRec* arec = recs + A.meta_off;
    const uint16_t* asrc = srcl + A.meta_off;
    const int32_t* arowpt = rowpt + A.row_off;
    const uint32_t* ansrc = nsrc + A.row_off;
    const uint32_t* abase = rowbase + A.row_off;
    const double* T = Tbuf + A.t_off;
    uint32_t m = best_m, h = 0, ok = 1;
    int row = best_row;
    hull[h++] = A.pt;
    const uint32_t hcap = A.rows + 2;        // a polygon has at most rows + 1 vertices; on exactly degenerate inputs a step can
    while (m > 2)                            // consume no point (c = 0) and the walk must not run away
    {
        if (h >= hcap) { ok = 0; break; }
        const int last = hull[h - 1], j = arowpt[row];
        const uint32_t ns = ansrc[row];
        const double* Tm = T + (size_t) (m % W) * U + abase[row];
        int pick = -1;
        for (uint32_t k = ns; k-- > 0;)
        {
            const double v = Tm[k], vp = k ? Tm[k - 1] : INFINITY;
            if (v == vp) continue;                              // the running minimum did not change here
            const int l = arowpt[asrc[(size_t) row * acap + k]];
            if (orientation(x[last], y[last], x[j], y[j], x[l], y[l]) >= 0) { pick = (int) k; break; }
        }
        if (pick < 0) { ok = 0; break; }
        const SrcRec r = arec[(size_t) row * acap + pick];
        m -= m == 3 ? (r.c3 & 0x7fffu) : r.c;
        hull[h++] = j;
        row = (int) asrc[(size_t) row * acap + pick];
    }
    hull[h++] = arowpt[row];
    out->len = h; out->ok = ok;
    hull_diameter_dev(x, y, hull, h, out->diam);
}

// Test hook: the slope list of (anchor i, centre j) exactly as k_area_prep walks it.
__global__ void k_slope_list(const double* __restrict__ x, const double* __restrict__ y, uint32_t n,
                             const int32_t* __restrict__ modpi, const uint16_t* __restrict__ rank,
                             int i, int j, int32_t* __restrict__ out, uint32_t* __restrict__ out_len)
{
    __shared__ int32_t s_def[32];
    const uint32_t lane = threadIdx.x, len = n - 1;
    const double ix = x[i], iy = y[i], jx = x[j], jy = y[j];
    const int32_t* ring_j = modpi + (size_t) j * len;
    const uint32_t r0 = rank[(size_t) j * n + i];
    uint32_t kept = 0, ndef = 0;
    for (uint32_t base = 1; base < len; base += 32)
    {
        const uint32_t t = base + lane;
        bool keep = false, defer = false; int p = -1;
        if (t < len)
        {
            p = ring_j[(r0 + t) % len];
            if (y[p] >= iy)
            {
                const double s = signed_area(ix, iy, jx, jy, x[p], y[p]);
                if (near_zero(s)) defer = y[p] > jy; else keep = true;
            }
        }
        const uint32_t km = __ballot_sync(0xffffffffu, keep), dm = __ballot_sync(0xffffffffu, defer);
        const uint32_t lt = (1u << lane) - 1u;
        if (keep) out[kept + __popc(km & lt)] = p;
        if (defer) { const uint32_t d = ndef + __popc(dm & lt); if (d < 32) s_def[d] = p; }
        kept += __popc(km); ndef = min(32u, ndef + __popc(dm));
    }
    __syncwarp();
    if (lane == 0)
    {
        for (uint32_t u = 1; u < ndef; ++u)
        {
            const int p = s_def[u]; const double dp = dist2(jx, jy, x[p], y[p]);
            uint32_t v = u;
            while (v > 0 && dist2(jx, jy, x[s_def[v - 1]], y[s_def[v - 1]]) > dp) { s_def[v] = s_def[v - 1]; --v; }
            s_def[v] = p;
        }
        for (uint32_t u = 0; u < ndef; ++u) out[kept++] = s_def[u];
        *out_len = kept;
    }
}

// --------------------------------------------------------------------------------------------- host side

namespace
{
struct AreaPlan
{
    std::vector<AnchorInfo> info;
    uint64_t meta_total = 0, row_total = 0;
    uint32_t amax_rows = 0;
};

struct SmallOut { unsigned int queue; unsigned int pad; AreaBest best; AreaHull hull; unsigned int cluster_next[1024]; };

int area_fetch_yorder(Ctx* c, std::vector<int32_t>& yorder, std::vector<int32_t>& above)
{
    const uint32_t n = c->n;
    ADK_TRY(run_yorder(c));
    ADK_TRY(ensure_pinned(c, sizeof(int32_t) * 2 * (size_t) n));
    int32_t* h = static_cast<int32_t*>(c->h_pinned);
    ADK_CUDA(c, cudaMemcpyAsync(h, c->yorder.p, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaMemcpyAsync(h + n, c->above.p, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaStreamSynchronize(c->stream));
    yorder.assign(h, h + n);
    above.assign(h + n, h + 2 * n);
    return ADK_OK;
}

void area_plan_add(AreaPlan& P, int32_t pt, int32_t a, uint32_t order)
{
    AnchorInfo A;
    std::memset(&A, 0, sizeof(A));
    A.pt = pt; A.a = a; A.order = order;
    A.meta_off = P.meta_total; A.row_off = P.row_total;
    P.meta_total += (uint64_t) a * (uint64_t) a;
    P.row_total += (uint64_t) a;
    P.amax_rows = std::max<uint32_t>(P.amax_rows, (uint32_t) a);
    P.info.push_back(A);
}

// Runs PREP for every anchor in the plan, then the DP in memory-bounded batches.  full_m != 0: keep every
// level up to full_m (reconstruction); otherwise ring buffers of (c_max + 1) levels.
int area_run(Ctx* c, AreaPlan& P, uint64_t K, double amax, uint32_t full_m, float* dp_ms)
{
    const uint32_t n = c->n;
    const uint32_t cnt = (uint32_t) P.info.size();
    if (dp_ms) *dp_ms = 0.f;
    if (cnt == 0) return ADK_OK;
    ADK_TRY(ensure(c, c->scratch0, sizeof(AnchorInfo) * cnt));
    ADK_TRY(ensure(c, c->scratch1, sizeof(SrcRec) * std::max<uint64_t>(P.meta_total, 1)));
    ADK_TRY(ensure(c, c->scratch2, sizeof(uint16_t) * std::max<uint64_t>(P.meta_total, 1)));
    ADK_TRY(ensure(c, c->area_mcr, sizeof(uint32_t) * kReachBuckets * std::max<uint64_t>(P.meta_total, 1)));
    ADK_TRY(ensure(c, c->scratch3, sizeof(int32_t) * 3 * std::max<uint64_t>(P.row_total, 1)));
    const uint32_t prep_grid = std::min<uint32_t>(cnt, (uint32_t) c->sm_count);   // 1024-thread CTAs, one per SM
    const uint32_t qstride = P.amax_rows * P.amax_rows;
    ADK_TRY(ensure(c, c->scratch4, sizeof(uint16_t) * std::max<uint64_t>((uint64_t) qstride * prep_grid, 1)));
    ADK_TRY(ensure(c, c->small, sizeof(SmallOut)));
    AnchorInfo* d_info = c->scratch0.as<AnchorInfo>();
    SrcRec* d_recs = c->scratch1.as<SrcRec>();
    uint16_t* d_srcl = c->scratch2.as<uint16_t>();
    int32_t* d_rowpt = c->scratch3.as<int32_t>();
    uint32_t* d_nsrc = reinterpret_cast<uint32_t*>(d_rowpt + P.row_total);
    uint32_t* d_rowbase = d_nsrc + P.row_total;
    Tables T{c->below.as<int32_t>(), c->coll.as<int32_t>(), c->rb.as<int32_t>(), n};

    ADK_CUDA(c, cudaMemcpyAsync(d_info, P.info.data(), sizeof(AnchorInfo) * cnt, cudaMemcpyHostToDevice, c->stream));
    if (P.amax_rows > 0)
    {
        k_area_prep<<<prep_grid, kPrepThreads, sizeof(uint16_t) * n, c->stream>>>(
            c->x.as<double>(), c->y.as<double>(), n, T, c->cw.as<int32_t>(), c->modpi.as<int32_t>(), c->rank.as<uint16_t>(),
            d_info, 0, cnt, d_recs, d_srcl, d_rowpt, d_nsrc, d_rowbase, c->scratch4.as<uint16_t>(), qstride, K, amax, c->area_mcr.as<uint32_t>());
        ++c->launches;
        ADK_CUDA(c, cudaGetLastError());
    }
    ADK_CUDA(c, cudaMemcpyAsync(P.info.data(), d_info, sizeof(AnchorInfo) * cnt, cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaStreamSynchronize(c->stream));
    for (const AnchorInfo& A : P.info)
        if (A.defer_overflow)
        {
            c->err = "area path: more than 32 collinear points beyond one polygon vertex (exactly degenerate input, outside the contract)";
            return ADK_ERR_ARG;
        }

    // table plan: ring of (cmax + 1) levels, or every level up to full_m
    // Ring mode runs the row pipeline (k_area_rows, bands of 32 levels); keeping every level for the walk back (one anchor)
    // stays on the level wavefront.  ADK_AREA_WAVEFRONT=1 forces the wavefront everywhere (A/B measurements).
    const bool by_rows = full_m == 0 && !c->area_wavefront;
    const uint64_t budget_entries = std::max<uint64_t>(workspace_budget(c, c->scratch5) / sizeof(double), 1u << 20);
    if (by_rows)
    {
        // One launch for all anchors: every resident CLUSTER owns one table slot sized for the largest anchor and reuses it for the
        // anchors it pulls from the queue, so the table memory is (resident clusters) x (largest ring), whatever the number of anchors.
        // Cluster size: 1 while there are anchors enough to keep every SM busy (nothing is speculated then); with few anchors
        // per GPU (a shard of a multi-GPU solve) the bands of an anchor are pipelined over 2, 4 or 8 CTAs, which shortens the
        // heaviest anchor's critical path.
        uint32_t max_rows = 1;
        for (uint32_t t = 0; t < cnt; ++t) max_rows = std::max<uint32_t>(max_rows, (uint32_t) P.info[t].a);
        const uint32_t cap_rows = (std::min<uint32_t>(max_rows, 4096) + 1u) & ~1u;      // even: the staging area behind the row table stays 16-byte aligned
        const size_t smem = sizeof(uint32_t) * 2 * cap_rows + sizeof(double) * kRowWarps * kSeg * kStageDoubles;
        ADK_CUDA(c, cudaFuncSetAttribute(k_area_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        int per_sm = 0;
        ADK_CUDA(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_area_rows, kRowThreads, smem));
        const uint64_t resident = (uint64_t) std::max(per_sm, 1) * (uint64_t) c->sm_count;
        uint32_t CS = 1;
        if (c->area_cluster) CS = c->area_cluster;
        else
        {   // Anchor times go like their source counts U.  One launch lasts max(all the work over the resident CTAs, the heaviest
            // anchor alone); a cluster of CS CTAs divides the second term by `gain` (bands in flight) and multiplies the first by
            // `waste` (bands started past the stopping point, the one-row lag) — both measured at n = 1000 (profiles/r2_area_*).
            static const double waste[4] = {1.0, 1.25, 1.6, 2.2}, gain[4] = {1.0, 1.9, 3.4, 5.5};
            double sumU = 0.0, maxU = 0.0;
            for (uint32_t t = 0; t < cnt; ++t) { sumU += (double) P.info[t].U; maxU = std::max(maxU, (double) P.info[t].U); }
            double best = 0.0;
            for (int v = 0; v < 4; ++v)
            {
                const double t = std::max(sumU * waste[v] / (double) resident, maxU / gain[v]);
                if (v == 0 || t < 0.95 * best) { best = v == 0 ? t : std::min(best, t); CS = 1u << v; }
            }
        }
        // Table slots: one per resident cluster, sized for the anchors that cluster can meet; every anchor's ring fills its slot
        // (adk_area_plan.h, checked on the CPU by tests/dropin/slot_plan_check.cpp).
        std::vector<AnchorShape> shapes(cnt);
        for (uint32_t t = 0; t < cnt; ++t) shapes[t] = AnchorShape{(uint32_t) P.info[t].a, P.info[t].U, P.info[t].cmax};
        SlotPlan plan;
        if (!plan_area_slots(shapes, K, CS, resident, budget_entries, kBand, plan))
        {
            c->err = "area path: one anchor's level tables exceed the memory budget";
            return ADK_ERR_NOMEM;
        }
        const uint64_t slots = plan.slots;
        const std::vector<uint64_t>& slot_off = plan.off;
        for (uint32_t t = 0; t < cnt; ++t) { P.info[t].W = plan.W[t]; P.info[t].t_off = 0; }
        ADK_TRY(ensure(c, c->scratch5, sizeof(double) * std::max<uint64_t>(slot_off[slots], 32)));
        ADK_TRY(ensure(c, c->area_slot_off, sizeof(uint64_t) * (slots + 1)));
        ADK_CUDA(c, cudaMemcpyAsync(c->area_slot_off.p, slot_off.data(), sizeof(uint64_t) * (slots + 1), cudaMemcpyHostToDevice, c->stream));
        const uint64_t slot_entries = slot_off[slots] / slots;        // mean, for the debug line
        ADK_TRY(ensure(c, c->area_boards, sizeof(BandBoard) * slots));
        const uint32_t hull_stride = (n + 8) & ~3u;
        ADK_TRY(ensure(c, c->area_keep_hull, sizeof(int32_t) * (size_t) hull_stride * slots));
        c->area_keep_valid = false;
        ADK_CUDA(c, cudaMemcpyAsync(d_info, P.info.data(), sizeof(AnchorInfo) * cnt, cudaMemcpyHostToDevice, c->stream));
        ADK_CUDA(c, cudaMemsetAsync(c->small.p, 0, sizeof(unsigned int) * 2, c->stream));
        ADK_CUDA(c, cudaEventRecord(c->ev2, c->stream));
        SmallOut* so_ = c->small.as<SmallOut>();
        {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3((uint32_t) slots * CS); cfg.blockDim = dim3(kRowThreads); cfg.dynamicSmemBytes = smem; cfg.stream = c->stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = CS; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr; cfg.numAttrs = 1;
            ADK_CUDA(c, cudaLaunchKernelEx(&cfg, k_area_rows, d_info, 0u, cnt, &so_->queue, (const SrcRec*) d_recs, (const uint32_t*) d_nsrc,
                                           (const uint32_t*) d_rowbase, c->scratch5.as<double>(), (const uint64_t*) c->area_slot_off.as<uint64_t>(), K, amax, cap_rows,
                                           c->area_boards.as<BandBoard>(), (const double*) c->x.as<double>(), (const double*) c->y.as<double>(),
                                           (const uint16_t*) d_srcl, (const int32_t*) d_rowpt, c->area_keep_hull.as<int32_t>(), hull_stride,
                                           (const uint32_t*) c->area_mcr.as<uint32_t>()));
        }
        c->area_keep_valid = true; c->area_keep_slots = (uint32_t) slots; c->area_keep_stride = hull_stride;
        c->area_keep_K = K; c->area_keep_amax = amax;
        ++c->launches;
        ADK_CUDA(c, cudaGetLastError());
        ADK_CUDA(c, cudaEventRecord(c->ev3, c->stream));
        if (dp_ms)
        {
            ADK_CUDA(c, cudaEventSynchronize(c->ev3));
            ADK_CUDA(c, cudaEventElapsedTime(dp_ms, c->ev2, c->ev3));
        }
        if (c->debug)
        {
            std::vector<AnchorInfo> h(cnt);
            ADK_CUDA(c, cudaMemcpyAsync(h.data(), d_info, sizeof(AnchorInfo) * cnt, cudaMemcpyDeviceToHost, c->stream));
            ADK_CUDA(c, cudaStreamSynchronize(c->stream));
            unsigned long long tmin = ~0ull, tmax = 0; double sum = 0.0, longest = 0.0;
            for (auto& a : h) { tmin = std::min(tmin, a.t0_ns); tmax = std::max(tmax, a.t1_ns); sum += (double) (a.t1_ns - a.t0_ns); longest = std::max(longest, (double) (a.t1_ns - a.t0_ns)); }
            { const AnchorInfo* w = nullptr; for (auto& a : h) if (!w || a.best_m > w->best_m) w = &a;
              if (w) std::fprintf(stderr, "[adk] best anchor: pt %d order %u m %u U %u cmax %u W %u (bands %u) levels_run %u rows %u a %d; slot_entries %llu CS %u\n", w->pt, w->order,
                                  w->best_m, w->U, w->cmax, w->W, w->W / 32, w->levels_run, w->rows, w->a, (unsigned long long) slot_entries, CS); }
            { double run = 0.0, useful = 0.0;
              for (auto& a : h) { run += (double) a.levels_run * a.U; const uint32_t tb = a.best_m > 2 ? (a.best_m - 2 + 31) / 32 : 1; useful += (double) tb * 32 * a.U; }
              std::fprintf(stderr, "[adk] executed level x source cells %.4g, of which up to each anchor's top band %.4g (%.1f %%)\n", run, useful, 100.0 * useful / run); }
            std::fprintf(stderr, "[adk] rows kernel: %u anchors on %llu CTAs, span %.2f ms, sum of anchor times %.2f ms, longest anchor %.2f ms\n",
                         cnt, (unsigned long long) slots * CS, (tmax - tmin) * 1e-6, sum * 1e-6, longest * 1e-6);
        }
        return ADK_OK;
    }
    uint32_t b0 = 0;
    while (b0 < cnt)
    {
        uint64_t total = 0;
        uint32_t b1 = b0;
        while (b1 < cnt)
        {
            AnchorInfo& A = P.info[b1];
            const uint64_t levels = K >= 3 ? K - 2 : 0;
            uint64_t W = full_m ? (uint64_t) full_m + 1 : std::min<uint64_t>((uint64_t) A.cmax + 1, std::max<uint64_t>(levels, 1));
            if (W < 1) W = 1;
            const uint64_t need = W * (uint64_t) A.U;
            if (b1 > b0 && total + need > budget_entries) break;
            if (need > budget_entries) { c->err = "area path: one anchor's level tables exceed the memory budget"; return ADK_ERR_NOMEM; }
            A.W = (uint32_t) W; A.t_off = total;
            total += need;
            ++b1;
        }
        ADK_TRY(ensure(c, c->scratch5, sizeof(double) * std::max<uint64_t>(total, 1)));
        ADK_CUDA(c, cudaMemcpyAsync(d_info + b0, P.info.data() + b0, sizeof(AnchorInfo) * (b1 - b0), cudaMemcpyHostToDevice, c->stream));
        ADK_CUDA(c, cudaMemsetAsync(c->small.p, 0, sizeof(unsigned int) * 2, c->stream));
        // cluster size by the depth of the heaviest anchor of the batch (rows / 64 warps' worth), 1 .. 8 CTAs per anchor
        uint32_t max_rows = 1;
        for (uint32_t t = b0; t < b1; ++t) max_rows = std::max<uint32_t>(max_rows, (uint32_t) P.info[t].a);
        uint32_t cs = 1;
        while (cs < 8 && max_rows >= 128u * cs) cs *= 2;
        const uint32_t nclusters = std::max<uint32_t>(1, std::min<uint32_t>(b1 - b0, (uint32_t) c->sm_count * 2 / cs));
        ADK_CUDA(c, cudaEventRecord(c->ev2, c->stream));
        {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(nclusters * cs); cfg.blockDim = dim3(kDpThreads); cfg.dynamicSmemBytes = 0; cfg.stream = c->stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr; cfg.numAttrs = 1;
            SmallOut* so_ = c->small.as<SmallOut>();
            ADK_CUDA(c, cudaLaunchKernelEx(&cfg, k_area_dp, d_info, b0, b1 - b0, &so_->queue, (const SrcRec*) d_recs, (const uint32_t*) d_nsrc,
                                           (const uint32_t*) d_rowbase, c->scratch5.as<double>(), K, amax, so_->cluster_next));
        }
        ++c->launches;
        ADK_CUDA(c, cudaGetLastError());
        ADK_CUDA(c, cudaEventRecord(c->ev3, c->stream));
        if (dp_ms || b1 < cnt)
        {
            ADK_CUDA(c, cudaEventSynchronize(c->ev3));
            float ms = 0.f;
            ADK_CUDA(c, cudaEventElapsedTime(&ms, c->ev2, c->ev3));
            if (dp_ms) *dp_ms += ms;
        }
        b0 = b1;
    }
    return ADK_OK;
}
}  // namespace

int area_partial(Ctx* c, uint64_t kmax, double amax, uint32_t shard, uint32_t nshards, adk_partial* out, adk_stats* st)
{
    std::memset(out, 0, sizeof(*out));
    if (st) std::memset(st, 0, sizeof(*st));
    const uint32_t n = c->n;
    if (nshards == 0 || shard >= nshards) { c->err = "shard out of range"; return ADK_ERR_ARG; }
    if (n < 3 || kmax < 3) return ADK_OK;            // Eppstein.cpp:137-140
    const uint64_t K = std::min<uint64_t>(kmax, n);  // Eppstein.cpp:162
    ADK_CUDA(c, cudaEventRecord(c->ev0, c->stream));
    ADK_TRY(run_angles_and_orders(c, true, true));
    ADK_TRY(run_count_tables(c));
    std::vector<int32_t> yorder, above;
    ADK_TRY(area_fetch_yorder(c, yorder, above));
    // anchors in the reference's order are also in decreasing-cost order: deal them round-robin to the shards
    AreaPlan P;
    for (uint32_t r = shard; r < n; r += nshards)
    {
        const int32_t pt = yorder[r];
        if (above[pt] >= 2) area_plan_add(P, pt, above[pt], r);
    }
    float dp_ms = 0.f;
    ADK_TRY(area_run(c, P, K, amax, 0, &dp_ms));
    const uint32_t cnt = (uint32_t) P.info.size();
    AreaBest hb;
    std::memset(&hb, 0, sizeof(hb));
    if (cnt)
    {
        SmallOut* so = c->small.as<SmallOut>();
        k_area_reduce<<<1, 256, 0, c->stream>>>(c->scratch0.as<AnchorInfo>(), cnt, c->scratch3.as<int32_t>(), K, &so->best);
        ++c->launches;
        ADK_CUDA(c, cudaGetLastError());
        ADK_CUDA(c, cudaMemcpyAsync(&hb, &so->best, sizeof(hb), cudaMemcpyDeviceToHost, c->stream));
    }
    ADK_CUDA(c, cudaEventRecord(c->ev1, c->stream));
    ADK_CUDA(c, cudaEventSynchronize(c->ev1));
    float ms = 0.f;
    ADK_CUDA(c, cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    if (hb.found)
    {
        out->found = 1; out->count = hb.m; out->area = hb.area;
        out->tie_key = ((uint64_t) hb.order << 32) | (uint32_t) hb.row;
        out->w0 = (uint32_t) hb.pt; out->w1 = (uint32_t) hb.jpt;
    }
    if (st)
    {
        st->cells_nominal = hb.cells_nominal; st->cells_executed = hb.cells_executed; st->states = hb.required;
        st->kernel_ms = ms; st->dp_ms = dp_ms; st->launches = c->launches;
        st->arena_bytes = c->scratch1.cap + c->scratch2.cap + c->scratch3.cap + c->scratch4.cap + c->scratch5.cap;
    }
    return ADK_OK;
}

int area_finish(Ctx* c, uint64_t kmax, double amax, const adk_partial* best, uint32_t flags, adk_result* out, uint32_t* hull)
{
    const uint32_t n = c->n;
    out->found = best->found;
    if (!best->found) return ADK_OK;
    out->count = best->count; out->area = best->area;
    if (!(flags & ADK_FLAG_RECONSTRUCT)) return ADK_OK;
    const uint64_t K = std::min<uint64_t>(kmax, n);
    if (best->w0 >= n) { c->err = "area_finish: bad witness"; return ADK_ERR_ARG; }
    if (c->debug) std::fprintf(stderr, "[adk] area_finish: keep valid %d K %llu/%llu amax %g/%g slots %u\n", (int) c->area_keep_valid,
                               (unsigned long long) K, (unsigned long long) c->area_keep_K, amax, c->area_keep_amax, c->area_keep_slots);
    if (c->area_keep_valid && c->area_keep_K == K && c->area_keep_amax == amax && c->area_keep_slots > 0)
    {   // the partial solve on this context walked the polygon of every cluster's best anchor back while its table was still
        // there: if the winner is one of them (it is, when it was found on this GPU) there is nothing to re-run
        const uint32_t slots = c->area_keep_slots;
        std::vector<KeepRec> recs(slots);
        ADK_CUDA(c, cudaMemcpy2DAsync(recs.data(), sizeof(KeepRec), reinterpret_cast<const char*>(c->area_boards.p) + offsetof(BandBoard, keep),
                                      sizeof(BandBoard), sizeof(KeepRec), slots, cudaMemcpyDeviceToHost, c->stream));
        ADK_CUDA(c, cudaStreamSynchronize(c->stream));
        if (c->debug)
        {
            uint32_t nf = 0; for (uint32_t t = 0; t < slots; ++t) nf += recs[t].found ? 1u : 0u;
            std::fprintf(stderr, "[adk] area_finish: %u kept records over %u clusters\n", nf, slots);
        }
        if (c->debug)
            for (uint32_t t = 0; t < slots; ++t)
                if (recs[t].found && (recs[t].m + 2 >= best->count || t < 3))
                    std::fprintf(stderr, "[adk] keep[%u]: m %u area %.17g order %u pt %d len %u ok %u   (winner: m %u area %.17g order %u pt %u)\n", t, recs[t].m,
                                 recs[t].area, recs[t].order, recs[t].pt, recs[t].len, recs[t].ok, best->count, best->area, (uint32_t) (best->tie_key >> 32), best->w0);
        for (uint32_t t = 0; t < slots; ++t)
        {
            const KeepRec& kr = recs[t];
            if (!kr.found || !kr.ok || kr.m != best->count || kr.order != (uint32_t) (best->tie_key >> 32) || (uint32_t) kr.pt != best->w0 ||
                kr.area != best->area || kr.len == 0 || kr.len > n) continue;
            ADK_TRY(ensure_pinned(c, sizeof(int32_t) * ((size_t) n + 8)));
            ADK_CUDA(c, cudaMemcpyAsync(c->h_pinned, c->area_keep_hull.as<int32_t>() + (size_t) t * c->area_keep_stride, sizeof(int32_t) * kr.len,
                                        cudaMemcpyDeviceToHost, c->stream));
            ADK_CUDA(c, cudaStreamSynchronize(c->stream));
            const int32_t* hp = static_cast<const int32_t*>(c->h_pinned);
            out->hull_len = kr.len;
            for (uint32_t q = 0; q < kr.len; ++q) hull[q] = (uint32_t) hp[q];
            out->diam_valid = 1; out->diam[0] = kr.diam[0]; out->diam[1] = kr.diam[1];      // Eppstein.cpp:271-273
            return ADK_OK;
        }
    }
    ADK_TRY(run_angles_and_orders(c, true, true));
    ADK_TRY(run_count_tables(c));
    std::vector<int32_t> yorder, above;
    ADK_TRY(area_fetch_yorder(c, yorder, above));
    AreaPlan P;
    area_plan_add(P, (int32_t) best->w0, above[best->w0], (uint32_t) (best->tie_key >> 32));
    // re-run the winning anchor keeping every level up to m*, then walk back (Eppstein.cpp:240-274)
    ADK_TRY(area_run(c, P, std::min<uint64_t>(K, best->count), amax, best->count, nullptr));
    ADK_TRY(ensure(c, c->scratch6, sizeof(int32_t) * ((size_t) n + 4)));
    SmallOut* so = c->small.as<SmallOut>();
    int32_t* d_rowpt = c->scratch3.as<int32_t>();
    uint32_t* d_nsrc = reinterpret_cast<uint32_t*>(d_rowpt + P.row_total);
    uint32_t* d_rowbase = d_nsrc + P.row_total;
    k_area_backtrack<<<1, 32, 0, c->stream>>>(c->x.as<double>(), c->y.as<double>(), c->scratch0.as<AnchorInfo>(),
                                             c->scratch1.as<SrcRec>(), c->scratch2.as<uint16_t>(), d_rowpt, d_nsrc, d_rowbase,
                                             c->scratch5.as<double>(), best->count, (int32_t) (best->tie_key & 0xffffffffu),
                                             c->scratch6.as<int32_t>(), &so->hull);
    ++c->launches;
    ADK_CUDA(c, cudaGetLastError());
    AreaHull hh;
    ADK_TRY(ensure_pinned(c, sizeof(int32_t) * ((size_t) n + 4)));
    ADK_CUDA(c, cudaMemcpyAsync(&hh, &so->hull, sizeof(hh), cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaMemcpyAsync(c->h_pinned, c->scratch6.p, sizeof(int32_t) * ((size_t) n + 4), cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaStreamSynchronize(c->stream));
    if (!hh.ok) { c->err = "area_finish: reconstruction walked off the table"; return ADK_ERR_ARG; }
    out->hull_len = hh.len;
    const int32_t* hp = static_cast<const int32_t*>(c->h_pinned);
    for (uint32_t t = 0; t < hh.len && t < n; ++t) hull[t] = (uint32_t) hp[t];
    out->diam_valid = 1; out->diam[0] = hh.diam[0]; out->diam[1] = hh.diam[1];   // Eppstein.cpp:271-273
    return ADK_OK;
}

int area_slope_order(Ctx* c, uint32_t i, uint32_t j, int32_t* h_out, uint32_t* h_len)
{
    const uint32_t n = c->n;
    if (n < 3 || i >= n || j >= n || i == j) { c->err = "slope_order: bad indices"; return ADK_ERR_ARG; }
    ADK_TRY(run_angles_and_orders(c, false, true));
    ADK_TRY(ensure(c, c->scratch6, sizeof(int32_t) * ((size_t) n + 4)));
    ADK_TRY(ensure(c, c->small, sizeof(SmallOut)));
    k_slope_list<<<1, 32, 0, c->stream>>>(c->x.as<double>(), c->y.as<double>(), n, c->modpi.as<int32_t>(), c->rank.as<uint16_t>(),
                                          (int) i, (int) j, c->scratch6.as<int32_t>(), &c->small.as<SmallOut>()->queue);
    ++c->launches;
    ADK_CUDA(c, cudaGetLastError());
    unsigned int len = 0;
    ADK_CUDA(c, cudaMemcpyAsync(&len, &c->small.as<SmallOut>()->queue, sizeof(len), cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaStreamSynchronize(c->stream));
    ADK_CUDA(c, cudaMemcpyAsync(h_out, c->scratch6.p, sizeof(int32_t) * len, cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaStreamSynchronize(c->stream));
    *h_len = len;
    return ADK_OK;
}

}  // namespace adk
