// adk_ad.cu — path AD: area- and diameter-bounded largest convex subset on sm_100a
// (MT::AntipodalOptimizedAlgorithm*, AntipodalOptimized.cpp:452-545; also serves MT::AntipodalAlgorithm*).
//
// B200-first reformulation (DESIGN.md "Path AD"):
//   * Candidate diameter pairs (s,t) are independent (AntipodalOptimized.cpp:495-527): a persistent grid pulls
//     them from a device queue, ONE CANDIDATE PER CTA.
//   * The reference's per-layer std::unordered_map<key, Entry> becomes a dense per-layer table indexed by the
//     state (side, o, s, ss) holding the area as an order-preserving 64-bit pattern; "insert or keep the smaller"
//     (AntipodalOptimized.cpp:207-217) is one atomicMin, and a per-layer bitmap is the frontier.  Tables live in a
//     per-CTA slice of an L2/HBM arena that every CTA leaves clean (each consumed state is reset), so the arena is
//     initialised once per context, not once per candidate.
//   * Layers are a wavefront (every transition adds c >= 1 points): bitmap scan -> CTA-local queue -> one warp per
//     state, lanes over the candidate next vertices (the loops at AntipodalOptimized.cpp:198,225,271,298).
//   * No back-pointers are stored: the winning candidate is re-run keeping its tables and the witness is recovered
//     by re-deriving, for each state, the predecessor whose relaxed value equals the stored one.
#include "adk_internal.h"
#include "adk_device.cuh"

#include <algorithm>
#include <cmath>
#include <cstring>

namespace adk
{

struct PairRec { uint32_t s, t; uint32_t nl, nr; };   // nl, nr include s and t themselves

struct PairStats
{
    unsigned int npairs_all;        // candidates within the diameter bound (all shards)
    unsigned int nqueued;           // candidates of this shard that can hold at least one interior point
    unsigned int max_nl, max_nr, max_lim;
    unsigned int pad;
    unsigned long long max_cells;   // max over queued candidates of (lim + 1) * S
    unsigned long long max_S;       // max states per layer
};

struct CtaBest { uint32_t found; uint32_t count; double area; unsigned long long key; uint32_t s, t; };

struct AdCounters { unsigned long long cells, states; unsigned int queue; unsigned int pad; };

// Lune membership of p for the candidate (s,t) (AntipodalOptimized.cpp:45-58): 0 = outside, 1 = left list, 2 = right list
__device__ __forceinline__ int lune_side(double sx, double sy, double tx, double ty, double d2, double px, double py)
{
    if (!(dist2(sx, sy, px, py) <= d2 && dist2(tx, ty, px, py) <= d2)) return 0;
    const double stx = __dsub_rn(tx, sx), sty = __dsub_rn(ty, sy);
    const double tsx = __dsub_rn(sx, tx), tsy = __dsub_rn(sy, ty);
    const double d0 = __dadd_rn(__dmul_rn(stx, __dsub_rn(px, sx)), __dmul_rn(sty, __dsub_rn(py, sy)));
    const double d1 = __dadd_rn(__dmul_rn(tsx, __dsub_rn(px, tx)), __dmul_rn(tsy, __dsub_rn(py, ty)));
    if (!(d0 >= 0.0 && d1 >= 0.0)) return 0;
    return orientation(sx, sy, tx, ty, px, py) <= 0 ? 2 : 1;   // collinear goes right (:60)
}

// Squared projection on s->t, clamped (CustomMath.cpp:67-87) for p other than s, t.
__device__ __forceinline__ double proj_dist2(double sx, double sy, double tx, double ty, double d2, double px, double py)
{
    const double dot = __dadd_rn(__dmul_rn(__dsub_rn(tx, sx), __dsub_rn(px, sx)), __dmul_rn(__dsub_rn(ty, sy), __dsub_rn(py, sy)));
    double r = __ddiv_rn(__dmul_rn(dot, dot), d2);
    r = r < 0.0 ? 0.0 : r;
    r = r > d2 ? d2 : r;
    return r;
}

// --------------------------------------------------------------------------------------------- candidate enumeration

// One warp per pair (i < j) within the diameter bound: count its lune, size its tables, queue it.
__global__ void k_ad_pairs(const double* __restrict__ x, const double* __restrict__ y, uint32_t n, double dmax2,
                           uint64_t kmax, uint32_t shard, uint32_t nshards,
                           PairRec* __restrict__ pairs, PairStats* __restrict__ stats)
{
    const uint32_t i = blockIdx.x;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const double sx = x[i], sy = y[i];
    for (uint32_t j = i + 1 + warp; j < n; j += nwarps)
    {
        const double tx = x[j], ty = y[j];
        const double d2 = dist2(sx, sy, tx, ty);
        if (!(d2 <= dmax2)) continue;                        // AntipodalOptimized.cpp:500
        uint32_t cl = 0, cr = 0;
        for (uint32_t p = lane; p < n; p += 32)
        {
            if (p == i || p == j) continue;
            const int side = lune_side(sx, sy, tx, ty, d2, x[p], y[p]);
            cl += side == 1; cr += side == 2;
        }
        for (int o = 16; o; o >>= 1) { cl += __shfl_xor_sync(0xffffffffu, cl, o); cr += __shfl_xor_sync(0xffffffffu, cr, o); }
        if (lane == 0)
        {
            atomicAdd(&stats->npairs_all, 1u);
            const uint32_t nl = cl + 2, nr = cr + 2;
            const uint64_t inner = (uint64_t) cl + cr;
            const uint64_t lim = kmax - 2 < inner ? kmax - 2 : inner;      // :137
            const uint64_t key = (uint64_t) i * n + j;
            if (lim > 0 && key % nshards == shard)
            {
                const uint32_t slot = atomicAdd(&stats->nqueued, 1u);
                pairs[slot] = PairRec{i, j, nl, nr};
                const unsigned long long S = (unsigned long long) nl * nr * nr + (unsigned long long) nr * nl * nl;
                atomicMax(&stats->max_nl, nl); atomicMax(&stats->max_nr, nr); atomicMax(&stats->max_lim, (unsigned int) lim);
                atomicMax(&stats->max_S, S); atomicMax(&stats->max_cells, S * (lim + 1));
            }
        }
    }
}

// --------------------------------------------------------------------------------------------- the layered DP

constexpr int kAdThreads = 256;

struct LuneSmem
{
    int32_t* Lidx; int32_t* Ridx; double* Lx; double* Ly; double* Rx; double* Ry;
    int32_t* tidx; double* tproj; unsigned char* lflag;
};

__device__ __forceinline__ LuneSmem carve_smem(unsigned char* base, uint32_t capL, uint32_t capR)
{
    LuneSmem S;
    double* d = reinterpret_cast<double*>(base);
    S.Lx = d; d += capL; S.Ly = d; d += capL; S.Rx = d; d += capR; S.Ry = d; d += capR;
    S.tproj = d; d += capL + capR;
    int32_t* q = reinterpret_cast<int32_t*>(d);
    S.Lidx = q; q += capL; S.Ridx = q; q += capR; S.tidx = q; q += capL + capR;
    S.lflag = reinterpret_cast<unsigned char*>(q);
    return S;
}

__host__ __device__ inline size_t lune_smem_bytes(uint32_t capL, uint32_t capR, uint32_t layers)
{
    return (size_t) (capL + capR) * (2 * sizeof(double) + sizeof(double) + 2 * sizeof(int32_t)) + layers + 16;
}

// Builds L = [s, left points by ascending projection, t] and R = [t, right points by descending projection, s]
// (AntipodalOptimized.cpp:93-109) in shared memory.  All threads of the CTA participate.
__device__ void build_lune(const double* __restrict__ x, const double* __restrict__ y, uint32_t n, const PairRec& pr,
                           LuneSmem& S, uint32_t* s_cnt /* 2 shared counters */)
{
    const double sx = x[pr.s], sy = y[pr.s], tx = x[pr.t], ty = y[pr.t];
    const double d2 = dist2(sx, sy, tx, ty);
    const uint32_t capL = pr.nl;   // unsorted staging: left at tidx[0..], right at tidx[capL..]
    if (threadIdx.x == 0) { s_cnt[0] = 0; s_cnt[1] = 0; }
    __syncthreads();
    for (uint32_t p = threadIdx.x; p < n; p += blockDim.x)
    {
        if (p == pr.s || p == pr.t) continue;
        const double px = x[p], py = y[p];
        const int side = lune_side(sx, sy, tx, ty, d2, px, py);
        if (side == 0) continue;
        const uint32_t k = atomicAdd(&s_cnt[side - 1], 1u);
        const uint32_t at = side == 1 ? k : capL + k;
        S.tidx[at] = (int32_t) p;
        S.tproj[at] = proj_dist2(sx, sy, tx, ty, d2, px, py);
    }
    __syncthreads();
    const uint32_t cl = s_cnt[0], cr = s_cnt[1];
    // rank by counting; equal projections are ordered by point index
    for (uint32_t a = threadIdx.x; a < cl; a += blockDim.x)
    {
        const double pa = S.tproj[a]; const int32_t ia = S.tidx[a];
        uint32_t r = 0;
        for (uint32_t b = 0; b < cl; ++b) { const double pb = S.tproj[b]; r += (pb < pa || (pb == pa && S.tidx[b] < ia)) ? 1u : 0u; }
        S.Lidx[1 + r] = ia; S.Lx[1 + r] = x[ia]; S.Ly[1 + r] = y[ia];
    }
    for (uint32_t a = threadIdx.x; a < cr; a += blockDim.x)
    {
        const double pa = S.tproj[capL + a]; const int32_t ia = S.tidx[capL + a];
        uint32_t r = 0;
        for (uint32_t b = 0; b < cr; ++b) { const double pb = S.tproj[capL + b]; r += (pb > pa || (pb == pa && S.tidx[capL + b] < ia)) ? 1u : 0u; }
        S.Ridx[1 + r] = ia; S.Rx[1 + r] = x[ia]; S.Ry[1 + r] = y[ia];
    }
    if (threadIdx.x == 0)
    {
        S.Lidx[0] = (int32_t) pr.s; S.Lx[0] = sx; S.Ly[0] = sy;
        S.Lidx[cl + 1] = (int32_t) pr.t; S.Lx[cl + 1] = tx; S.Ly[cl + 1] = ty;
        S.Ridx[0] = (int32_t) pr.t; S.Rx[0] = tx; S.Ry[0] = ty;
        S.Ridx[cr + 1] = (int32_t) pr.s; S.Rx[cr + 1] = sx; S.Ry[cr + 1] = sy;
    }
    __syncthreads();
}

// locPointsWithinTriangleCount (AntipodalOptimized.cpp:21-37): p, pp = chain edge, q = fan apex
__device__ __forceinline__ int tri_count_ad(const Tables& T, bool layer0, int p, double px, double py, int pp, double ppx, double ppy,
                                            int q, double qx, double qy)
{
    int c = T.C(p, pp) + T.C(pp, q) + points_in_triangle(T, p, px, py, pp, ppx, ppy, q, qx, qy);
    if (layer0) c += T.C(p, q);
    return c;
}

// Arena slice of one CTA: [vals: cells u64][bitmap: cells/32 + layers words][queue: S u32]
struct Slice { unsigned long long* vals; unsigned int* bits; unsigned int* queue; };

__device__ __forceinline__ Slice carve_slice(unsigned char* arena, size_t slice_bytes, uint32_t cta, unsigned long long max_cells,
                                              unsigned long long max_S, uint32_t max_layers)
{
    unsigned char* base = arena + (size_t) cta * slice_bytes;
    Slice s;
    s.vals = reinterpret_cast<unsigned long long*>(base);
    s.bits = reinterpret_cast<unsigned int*>(base + sizeof(unsigned long long) * max_cells);
    const size_t words = (size_t) (max_cells / 32) + max_layers + 1;
    s.queue = s.bits + words;
    (void) max_S;
    return s;
}

__host__ inline size_t slice_bytes_for(unsigned long long max_cells, unsigned long long max_S, uint32_t max_layers)
{
    const size_t words = (size_t) (max_cells / 32) + max_layers + 1;
    size_t b = sizeof(unsigned long long) * max_cells + sizeof(unsigned int) * (words + max_S);
    return (b + 255) & ~(size_t) 255;
}

// Relax one destination state: "insert if absent, else keep the smaller area" (AntipodalOptimized.cpp:207-217).
__device__ __forceinline__ void relax(unsigned long long* lvals, unsigned int* lbits, uint32_t idx, double area, unsigned int& new_states)
{
    const unsigned long long old = atomicMin(lvals + idx, area_bits(area));
    if (old == kInfBits) { atomicOr(lbits + (idx >> 5), 1u << (idx & 31)); ++new_states; }
}

template <bool KEEP>
__global__ void __launch_bounds__(kAdThreads)
k_ad_dp(const double* __restrict__ x, const double* __restrict__ y, uint32_t n, Tables T,
        const PairRec* __restrict__ pairs, uint32_t npairs, AdCounters* __restrict__ counters,
        unsigned char* __restrict__ arena, size_t slice_bytes, unsigned long long max_cells, unsigned long long max_S,
        uint32_t max_layers, uint32_t capL, uint32_t capR,
        uint64_t kmax, uint32_t nlayers_ref, double amax, CtaBest* __restrict__ cta_best)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t s_cnt[2];
    __shared__ unsigned int s_pair, s_qn;
    __shared__ unsigned long long s_term;
    __shared__ CtaBest s_best;
    LuneSmem S = carve_smem(smem_raw, capL, capR);
    const Slice sl = carve_slice(arena, slice_bytes, blockIdx.x, max_cells, max_S, max_layers);
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    unsigned long long my_cells = 0; unsigned int my_states = 0;
    if (threadIdx.x == 0) { s_best.found = 0; s_best.count = 0; s_best.area = 0.0; s_best.key = ~0ull; s_best.s = s_best.t = 0; }

    for (;;)
    {
        __syncthreads();
        if (threadIdx.x == 0) s_pair = atomicAdd(&counters->queue, 1u);
        __syncthreads();
        const uint32_t pi = s_pair;
        if (pi >= npairs) break;
        const PairRec pr = pairs[pi];
        build_lune(x, y, n, pr, S, s_cnt);
        const uint32_t nl = pr.nl, nr = pr.nr;
        const uint64_t inner = (uint64_t) nl + nr - 4;
        const uint32_t lim = (uint32_t) (kmax - 2 < inner ? kmax - 2 : inner);
        const uint32_t SR = nl * nr * nr, SL = nr * nl * nl, SS = SR + SL;
        const uint32_t WD = (SS + 31) / 32;
        const double Sx = S.Lx[0], Sy = S.Ly[0], Ex = S.Lx[nl - 1], Ey = S.Ly[nl - 1];
        const int Sp = S.Lidx[0], Ep = S.Lidx[nl - 1];
        const double seg2 = dist2(Sx, Sy, Ex, Ey);
        for (uint32_t k = threadIdx.x; k <= lim; k += blockDim.x) S.lflag[k] = 0;
        __syncthreads();
        // layer-0 seeds (AntipodalOptimized.cpp:149-160): LL (o=0, s=0, ss=i) and RR (o=0, s=0, ss=i), area 0
        for (uint32_t i = 1 + threadIdx.x; i < nl; i += blockDim.x)
        {
            const uint32_t idx = SR + i;
            sl.vals[idx] = 0ull; atomicOr(sl.bits + (idx >> 5), 1u << (idx & 31)); ++my_states;
        }
        for (uint32_t i = 1 + threadIdx.x; i < nr; i += blockDim.x)
        {
            const uint32_t idx = i;
            sl.vals[idx] = 0ull; atomicOr(sl.bits + (idx >> 5), 1u << (idx & 31)); ++my_states;
        }
        if (threadIdx.x == 0) S.lflag[0] = 1;
        uint32_t best_k = 0; double best_area = 0.0;
        __syncthreads();

        for (uint32_t k = 0; k <= lim; ++k)
        {
            if (!S.lflag[k]) continue;        // uniform: written before the last barrier
            unsigned long long* lv = sl.vals + (size_t) k * SS;
            unsigned int* lb = sl.bits + (size_t) k * WD;
            if (threadIdx.x == 0) { s_qn = 0; s_term = kInfBits; }
            __syncthreads();
            // frontier: bitmap -> queue
            for (uint32_t w = threadIdx.x; w < WD; w += blockDim.x)
            {
                unsigned int bits = __ldcg(lb + w);
                if (!bits) continue;
                if (!KEEP) lb[w] = 0u;
                const uint32_t cnt = __popc(bits);
                uint32_t at = atomicAdd(&s_qn, cnt);
                while (bits) { const int b = __ffs(bits) - 1; bits &= bits - 1; sl.queue[at++] = w * 32 + b; }
            }
            __syncthreads();
            const uint32_t qn = s_qn;
            // expand: one warp per state
            for (uint32_t q = warp; q < qn; q += nwarps)
            {
                const uint32_t idx = sl.queue[q];
                const double area0 = bits_area(__ldcg(lv + idx));
                if (!KEEP) { if (lane == 0) lv[idx] = kInfBits; }
                const bool right = idx < SR;
                uint32_t o, s, ss;
                if (right) { ss = idx % nr; const uint32_t t = idx / nr; s = t % nr; o = t / nr; }
                else { const uint32_t v = idx - SR; ss = v % nl; const uint32_t t = v / nl; s = t % nl; o = t / nl; }
                // terminal states (AntipodalOptimized.cpp:167-175)
                if ((right && ss == nr - 1) || (!right && ss == nl - 1))
                {
                    if (k > 0 && lane == 0) atomicMin(&s_term, area_bits(area0));
                    continue;
                }
                // "own" chain = the side of the pending edge; "other" chain = the opposite one
                const double* ownx = right ? S.Rx : S.Lx; const double* owny = right ? S.Ry : S.Ly;
                const double* othx = right ? S.Lx : S.Rx; const double* othy = right ? S.Ly : S.Ry;
                const int32_t* owni = right ? S.Ridx : S.Lidx;
                const uint32_t nown = right ? nr : nl, noth = right ? nl : nr;
                const double ax = right ? Sx : Ex, ay = right ? Sy : Ey;      // fan apex of the own chain
                const int ap = right ? Sp : Ep;
                const double bx = right ? Ex : Sx, by = right ? Ey : Sy;      // fan apex of the other chain
                const double ox = othx[o], oy = othy[o];                       // current vertex of the other chain
                const double rx = ownx[s], ry = owny[s];                       // tail of the pending edge
                const double px = ownx[ss], py = owny[ss];                     // head of the pending edge
                if (dist2(ox, oy, px, py) > seg2) continue;                    // :183-186 / :256-259
                const double tri = fabs(signed_area(ax, ay, rx, ry, px, py));  // :188 / :261
                const uint32_t c = 1u + (uint32_t) tri_count_ad(T, k == 0, owni[s], rx, ry, owni[ss], px, py, ap, ax, ay);
                const double area = __dadd_rn(area0, tri);
                if ((uint64_t) c + k >= nlayers_ref) continue;                 // :192-195
                my_cells += (lane == 0) ? (unsigned long long) (nown - 1 - ss) + (noth - 1 - o) : 0ull;
                const uint32_t dst = k + c;
                if (dst > lim || !(area <= amax)) continue;                    // deeper layers are never read; :203
                unsigned long long* dv = sl.vals + (size_t) dst * SS;
                unsigned int* db = sl.bits + (size_t) dst * WD;
                bool any = false;
                // stay on the own chain (:198-220 / :271-293): new pending edge ss -> i
                const uint32_t base_stay = right ? (o * nr + ss) * nr : SR + (o * nl + ss) * nl;
                for (uint32_t i = ss + 1 + lane; i < nown; i += 32)
                {
                    const double qx = ownx[i], qy = owny[i];
                    if (orientation(qx, qy, px, py, rx, ry) >= 0 && orientation(qx, qy, px, py, ax, ay) >= 0)
                    { relax(dv, db, base_stay + i, area, my_states); any = true; }
                }
                // switch to the other chain (:222-248 / :295-324): new pending edge o -> i there
                const double refd = pseudo_dist(rx, ry, px, py, ox, oy);
                const uint32_t base_sw = right ? SR + (ss * nl + o) * nl : (ss * nr + o) * nr;
                for (uint32_t i = o + 1 + lane; i < noth; i += 32)
                {
                    const double qx = othx[i], qy = othy[i];
                    const double d = pseudo_dist(rx, ry, px, py, qx, qy);
                    if (d <= refd && (o == 0 || (orientation(ox, oy, bx, by, qx, qy) >= 0 && orientation(qx, qy, ox, oy, ax, ay) >= 0)))
                    { relax(dv, db, base_sw + i, area, my_states); any = true; }
                }
                if (any) S.lflag[dst] = 1;
            }
            __syncthreads();
            if (s_term != kInfBits) { best_k = k; best_area = bits_area(s_term); }
            __syncthreads();
        }
        // candidate result (AntipodalOptimized.cpp:328-331, 512-517): count desc, area asc, first pair in (i<j) order
        if (threadIdx.x == 0 && best_k > 0)
        {
            const uint32_t cnt = best_k + 2;
            const unsigned long long key = (unsigned long long) pr.s * n + pr.t;
            const bool better = !s_best.found || cnt > s_best.count ||
                                (cnt == s_best.count && (best_area < s_best.area || (best_area == s_best.area && key < s_best.key)));
            if (better) { s_best.found = 1; s_best.count = cnt; s_best.area = best_area; s_best.key = key; s_best.s = pr.s; s_best.t = pr.t; }
        }
    }
    // per-CTA totals
    for (int o = 16; o; o >>= 1) { my_cells += __shfl_xor_sync(0xffffffffu, my_cells, o); my_states += __shfl_xor_sync(0xffffffffu, my_states, o); }
    if (lane == 0) { atomicAdd(&counters->cells, my_cells); atomicAdd(&counters->states, (unsigned long long) my_states); }
    __syncthreads();
    if (threadIdx.x == 0) cta_best[blockIdx.x] = s_best;
}

__global__ void k_ad_reduce(const CtaBest* __restrict__ in, uint32_t count, CtaBest* __restrict__ out)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    CtaBest b; b.found = 0; b.count = 0; b.area = 0.0; b.key = ~0ull; b.s = b.t = 0;
    for (uint32_t i = 0; i < count; ++i)
    {
        const CtaBest c = in[i];
        if (!c.found) continue;
        const bool better = !b.found || c.count > b.count || (c.count == b.count && (c.area < b.area || (c.area == b.area && c.key < b.key)));
        if (better) b = c;
    }
    *out = b;
}

// Resets every table entry a KEEP run left behind (layers 0..lim of slice 0).
__global__ void k_ad_clean(unsigned char* __restrict__ arena, size_t slice_bytes, unsigned long long max_cells, unsigned long long max_S,
                           uint32_t max_layers, uint32_t SS, uint32_t lim)
{
    const Slice sl = carve_slice(arena, slice_bytes, 0, max_cells, max_S, max_layers);
    const uint32_t WD = (SS + 31) / 32;
    const size_t words = (size_t) WD * (lim + 1);
    for (size_t w = (size_t) blockIdx.x * blockDim.x + threadIdx.x; w < words; w += (size_t) gridDim.x * blockDim.x)
    {
        unsigned int bits = sl.bits[w];
        if (!bits) continue;
        sl.bits[w] = 0u;
        const size_t layer = w / WD, off = (w % WD) * 32;
        while (bits) { const int b = __ffs(bits) - 1; bits &= bits - 1; sl.vals[layer * SS + off + b] = kInfBits; }
    }
}

__global__ void k_fill_u64(unsigned long long* __restrict__ p, size_t count, unsigned long long v)
{
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (size_t) gridDim.x * blockDim.x) p[i] = v;
}

__global__ void k_ad_init_arena(unsigned char* __restrict__ arena, size_t slice_bytes, uint32_t nslices,
                                unsigned long long V, size_t words)
{
    // vals <- +inf, bitmaps <- 0 for every slice
    for (uint32_t s = blockIdx.y; s < nslices; s += gridDim.y)
    {
        unsigned char* base = arena + (size_t) s * slice_bytes;
        unsigned long long* vals = reinterpret_cast<unsigned long long*>(base);
        unsigned int* bits = reinterpret_cast<unsigned int*>(base + sizeof(unsigned long long) * V);
        for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < V; i += (size_t) gridDim.x * blockDim.x) vals[i] = kInfBits;
        for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (size_t) gridDim.x * blockDim.x) bits[i] = 0u;
    }
}

// --------------------------------------------------------------------------------------------- witness recovery

struct AdHull { uint32_t len; uint32_t ok; uint32_t nl, nr; };

// Re-derives the chain of states behind the best terminal state of layer kopt and emits the hull in the
// reference's order (locReconstructHull, AntipodalOptimized.cpp:335-436): [t, left chain backwards.., s, right chain backwards..].
__global__ void __launch_bounds__(kAdThreads)
k_ad_backtrack(const double* __restrict__ x, const double* __restrict__ y, uint32_t n, Tables T, PairRec pr,
               unsigned char* __restrict__ arena, size_t slice_bytes, unsigned long long max_cells, unsigned long long max_S,
               uint32_t max_layers, uint32_t kopt, uint32_t nlayers_ref,
               int32_t* __restrict__ left, int32_t* __restrict__ right, int32_t* __restrict__ hull, AdHull* __restrict__ out)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t s_cnt[2];
    LuneSmem S = carve_smem(smem_raw, pr.nl, pr.nr);
    build_lune(x, y, n, pr, S, s_cnt);
    if (threadIdx.x != 0) return;
    const Slice sl = carve_slice(arena, slice_bytes, 0, max_cells, max_S, max_layers);
    const uint32_t nl = pr.nl, nr = pr.nr;
    const uint32_t SR = nl * nr * nr, SL = nr * nl * nl, SS = SR + SL;
    const double Sx = S.Lx[0], Sy = S.Ly[0], Ex = S.Lx[nl - 1], Ey = S.Ly[nl - 1];
    const int Sp = S.Lidx[0], Ep = S.Lidx[nl - 1];
    const double seg2 = dist2(Sx, Sy, Ex, Ey);
    (void) nlayers_ref;
    // best terminal state of layer kopt (:348-355)
    bool e_right = true; uint32_t eo = 0, es = 0, ess = 0; unsigned long long ev = kInfBits;
    {
        const unsigned long long* lv = sl.vals + (size_t) kopt * SS;
        for (uint32_t o = 0; o < nl; ++o) for (uint32_t s = 0; s + 1 < nr; ++s)
        {
            const unsigned long long v = lv[(o * nr + s) * nr + (nr - 1)];
            if (v < ev) { ev = v; e_right = true; eo = o; es = s; ess = nr - 1; }
        }
        for (uint32_t o = 0; o < nr; ++o) for (uint32_t s = 0; s + 1 < nl; ++s)
        {
            const unsigned long long v = lv[SR + (o * nl + s) * nl + (nl - 1)];
            if (v < ev) { ev = v; e_right = false; eo = o; es = s; ess = nl - 1; }
        }
    }
    uint32_t a = 0, b = 0, ok = ev != kInfBits ? 1u : 0u;
    if (e_right) left[a++] = Ep; else right[b++] = Sp;     // :359-366
    uint32_t k = kopt;
    while (ok && k > 0)
    {
        const bool rt = e_right;
        const double* ownx = rt ? S.Rx : S.Lx; const double* owny = rt ? S.Ry : S.Ly;
        const double* othx = rt ? S.Lx : S.Rx; const double* othy = rt ? S.Ly : S.Ry;
        const int32_t* owni = rt ? S.Ridx : S.Lidx; const int32_t* othi = rt ? S.Lidx : S.Ridx;
        const uint32_t nown = rt ? nr : nl, noth = rt ? nl : nr;
        const double ax = rt ? Sx : Ex, ay = rt ? Sy : Ey; const int ap = rt ? Sp : Ep;        // own fan apex
        const double bx = rt ? Ex : Sx, by = rt ? Ey : Sy; const int bp = rt ? Ep : Sp;        // other fan apex
        const uint32_t own_base = rt ? 0u : SR, oth_base = rt ? SR : 0u;
        int32_t* own_out = rt ? right : left; uint32_t& own_cnt = rt ? b : a;
        int32_t* oth_out = rt ? left : right; uint32_t& oth_cnt = rt ? a : b;
        const uint32_t o = eo, s = es, ss = ess;
        bool found = false, via_switch = false; uint32_t prev = 0, cused = 0; unsigned long long pv = kInfBits;
        // (a) the state was produced by staying on its own chain: source (o, p, s) on the same side
        if (dist2(othx[o], othy[o], ownx[s], owny[s]) <= seg2)
            for (uint32_t p = 0; p < s && !found; ++p)
            {
                if (!(orientation(ownx[ss], owny[ss], ownx[s], owny[s], ownx[p], owny[p]) >= 0 &&
                      orientation(ownx[ss], owny[ss], ownx[s], owny[s], ax, ay) >= 0)) continue;
                const double tri = fabs(signed_area(ax, ay, ownx[p], owny[p], ownx[s], owny[s]));
                for (int z = 0; z < 2 && !found; ++z)
                {
                    const uint32_t c = 1u + (uint32_t) tri_count_ad(T, z == 1, owni[p], ownx[p], owny[p], owni[s], ownx[s], owny[s], ap, ax, ay);
                    if (c > k) continue;
                    const uint32_t ks = k - c;
                    if ((z == 1) != (ks == 0)) continue;
                    const unsigned long long sv = sl.vals[(size_t) ks * SS + own_base + (o * nown + p) * nown + s];
                    if (sv == kInfBits) continue;
                    if (area_bits(__dadd_rn(bits_area(sv), tri)) != ev) continue;
                    found = true; via_switch = false; prev = p; cused = c; pv = sv;
                }
            }
        // (b) it was produced by a switch from the other chain: source (o' = s, s' = p, ss' = o) on the other side
        if (!found && dist2(ownx[s], owny[s], othx[o], othy[o]) <= seg2)
            for (uint32_t p = 0; p < o && !found; ++p)
            {
                const double rx = othx[p], ry = othy[p], px = othx[o], py = othy[o];
                const double refd = pseudo_dist(rx, ry, px, py, ownx[s], owny[s]);
                const double d = pseudo_dist(rx, ry, px, py, ownx[ss], owny[ss]);
                if (!(d <= refd)) continue;
                if (!(s == 0 || (orientation(ownx[s], owny[s], ax, ay, ownx[ss], owny[ss]) >= 0 &&
                                 orientation(ownx[ss], owny[ss], ownx[s], owny[s], bx, by) >= 0))) continue;
                const double tri = fabs(signed_area(bx, by, rx, ry, px, py));
                for (int z = 0; z < 2 && !found; ++z)
                {
                    const uint32_t c = 1u + (uint32_t) tri_count_ad(T, z == 1, othi[p], rx, ry, othi[o], px, py, bp, bx, by);
                    if (c > k) continue;
                    const uint32_t ks = k - c;
                    if ((z == 1) != (ks == 0)) continue;
                    const unsigned long long sv = sl.vals[(size_t) ks * SS + oth_base + (s * noth + p) * noth + o];
                    if (sv == kInfBits) continue;
                    if (area_bits(__dadd_rn(bits_area(sv), tri)) != ev) continue;
                    found = true; via_switch = true; prev = p; cused = c; pv = sv;
                }
            }
        if (!found) { ok = 0; break; }
        own_out[own_cnt++] = owni[ss];                       // :374,388,403,417
        k -= cused;
        if (k == 0)
        {
            if (!via_switch) own_out[own_cnt++] = owni[s];   // :378,407
            else oth_out[oth_cnt++] = othi[o];               // :393,422
            break;
        }
        if (!via_switch) { eo = o; es = prev; ess = s; }
        else { e_right = !rt; eo = s; es = prev; ess = o; }
        ev = pv;
    }
    for (uint32_t t = 0; t < a; ++t) hull[t] = left[t];
    for (uint32_t t = 0; t < b; ++t) hull[a + t] = right[t];
    out->len = a + b; out->ok = ok; out->nl = nl; out->nr = nr;
}

// Test hook: the two lune lists of a candidate pair.
__global__ void __launch_bounds__(kAdThreads)
k_ad_lune_lists(const double* __restrict__ x, const double* __restrict__ y, uint32_t n, uint32_t s, uint32_t t,
                int32_t* __restrict__ out_l, int32_t* __restrict__ out_r, uint32_t* __restrict__ out_cnt)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t s_cnt[2];
    __shared__ uint32_t s_n[2];
    // count first so that the staging layout (left block of nl entries, then the right block) is known
    if (threadIdx.x == 0) { s_n[0] = 0; s_n[1] = 0; }
    __syncthreads();
    const double sx = x[s], sy = y[s], tx = x[t], ty = y[t];
    const double d2 = dist2(sx, sy, tx, ty);
    for (uint32_t p = threadIdx.x; p < n; p += blockDim.x)
    {
        if (p == s || p == t) continue;
        const int side = lune_side(sx, sy, tx, ty, d2, x[p], y[p]);
        if (side) atomicAdd(&s_n[side - 1], 1u);
    }
    __syncthreads();
    PairRec pr{s, t, s_n[0] + 2, s_n[1] + 2};
    LuneSmem S = carve_smem(smem_raw, pr.nl, pr.nr);
    build_lune(x, y, n, pr, S, s_cnt);
    for (uint32_t i = threadIdx.x; i < pr.nl; i += blockDim.x) out_l[i] = S.Lidx[i];
    for (uint32_t i = threadIdx.x; i < pr.nr; i += blockDim.x) out_r[i] = S.Ridx[i];
    if (threadIdx.x == 0) { out_cnt[0] = pr.nl; out_cnt[1] = pr.nr; }
}

// --------------------------------------------------------------------------------------------- host side

namespace
{
struct AdSmall { PairStats stats; AdCounters counters; CtaBest best; AdHull hull; uint32_t lune_cnt[2]; };

// Makes sure the arena holds `want_slices` clean slices able to take (V cells, L layers, Q queue entries).
int ad_prepare_arena(Ctx* c, unsigned long long V, uint32_t L, unsigned long long Q, uint32_t want_slices, uint32_t* got_slices)
{
    const unsigned long long nV = std::max<unsigned long long>(V, c->ad_V), nQ = std::max<unsigned long long>(Q, c->ad_Q);
    const uint32_t nL = std::max<uint32_t>(L, c->ad_L);
    const size_t sb = slice_bytes_for(nV, nQ, nL);
    uint64_t max_slices = workspace_budget(c, c->arena) / sb;
    if (max_slices == 0) { c->err = "area+diameter path: one candidate's layer tables exceed the memory budget"; return ADK_ERR_NOMEM; }
    uint32_t slices = (uint32_t) std::min<uint64_t>(want_slices, max_slices);
    const bool same_layout = nV == c->ad_V && nQ == c->ad_Q && nL == c->ad_L && c->ad_slice_bytes == sb;
    if (same_layout && slices <= c->ad_slices) { *got_slices = slices; return ADK_OK; }
    if (same_layout) slices = std::max<uint32_t>(slices, c->ad_slices);
    ADK_TRY(ensure(c, c->arena, sb * (size_t) slices));
    const size_t words = (size_t) (nV / 32) + nL + 1;
    dim3 grid((unsigned) std::min<size_t>((size_t) ((nV + 255) / 256), 1024), std::min<uint32_t>(slices, 65535));
    k_ad_init_arena<<<grid, 256, 0, c->stream>>>(c->arena.as<unsigned char>(), sb, slices, nV, words);
    ++c->launches;
    ADK_CUDA(c, cudaGetLastError());
    c->ad_V = nV; c->ad_Q = nQ; c->ad_L = nL; c->ad_slice_bytes = sb; c->ad_slices = slices;
    *got_slices = std::min<uint32_t>(want_slices, slices);
    return ADK_OK;
}

template <bool KEEP>
int ad_launch_dp(Ctx* c, const PairRec* d_pairs, uint32_t npairs, uint32_t grid, uint32_t capL, uint32_t capR, uint32_t layers,
                 uint64_t kmax, uint32_t nlayers_ref, double amax, AdCounters* d_counters, CtaBest* d_cta_best)
{
    const size_t smem = lune_smem_bytes(capL, capR, layers);
    if (smem > 200 * 1024) { c->err = "area+diameter path: lune lists do not fit in shared memory"; return ADK_ERR_NOMEM; }
    ADK_CUDA(c, cudaFuncSetAttribute(k_ad_dp<KEEP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    Tables T{c->below.as<int32_t>(), c->coll.as<int32_t>(), c->rb.as<int32_t>(), c->n};
    k_ad_dp<KEEP><<<grid, kAdThreads, smem, c->stream>>>(c->x.as<double>(), c->y.as<double>(), c->n, T, d_pairs, npairs, d_counters,
                                                         c->arena.as<unsigned char>(), c->ad_slice_bytes, c->ad_V, c->ad_Q, c->ad_L,
                                                         capL, capR, kmax, nlayers_ref, amax, d_cta_best);
    ++c->launches;
    ADK_CUDA(c, cudaGetLastError());
    return ADK_OK;
}
}  // namespace

int ad_partial(Ctx* c, uint64_t kmax, double amax, double dmax, uint32_t shard, uint32_t nshards, adk_partial* out, adk_stats* st)
{
    std::memset(out, 0, sizeof(*out));
    if (st) std::memset(st, 0, sizeof(*st));
    const uint32_t n = c->n;
    if (nshards == 0 || shard >= nshards) { c->err = "shard out of range"; return ADK_ERR_ARG; }
    if (std::min<uint64_t>(n, kmax) < 3) return ADK_OK;                 // AntipodalOptimized.cpp:461-464
    const uint32_t nlayers_ref = (uint32_t) (std::min<uint64_t>(n, kmax) - 1);   // :483
    const double dmax2 = dmax * dmax;                                  // :486
    ADK_CUDA(c, cudaEventRecord(c->ev0, c->stream));
    ADK_TRY(run_count_tables(c));
    ADK_TRY(ensure(c, c->scratch0, sizeof(PairRec) * (size_t) n * (n - 1) / 2));
    ADK_TRY(ensure(c, c->small, sizeof(AdSmall)));
    AdSmall* sm = c->small.as<AdSmall>();
    ADK_CUDA(c, cudaMemsetAsync(sm, 0, sizeof(AdSmall), c->stream));
    k_ad_pairs<<<n, 256, 0, c->stream>>>(c->x.as<double>(), c->y.as<double>(), n, dmax2, kmax, shard, nshards,
                                         c->scratch0.as<PairRec>(), &sm->stats);
    ++c->launches;
    ADK_CUDA(c, cudaGetLastError());
    PairStats ps;
    ADK_CUDA(c, cudaMemcpyAsync(&ps, &sm->stats, sizeof(ps), cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaStreamSynchronize(c->stream));
    float dp_ms = 0.f;
    CtaBest hb; std::memset(&hb, 0, sizeof(hb));
    AdCounters hc; std::memset(&hc, 0, sizeof(hc));
    if (ps.nqueued > 0)
    {
        uint32_t grid = 0;
        ADK_TRY(ad_prepare_arena(c, ps.max_cells, ps.max_lim + 1, ps.max_S, std::min<uint32_t>(ps.nqueued, (uint32_t) c->sm_count * 4), &grid));
        ADK_TRY(ensure(c, c->scratch1, sizeof(CtaBest) * grid));
        ADK_CUDA(c, cudaEventRecord(c->ev2, c->stream));
        ADK_TRY(ad_launch_dp<false>(c, c->scratch0.as<PairRec>(), ps.nqueued, grid, ps.max_nl, ps.max_nr, ps.max_lim + 1, kmax,
                                    nlayers_ref, amax, &sm->counters, c->scratch1.as<CtaBest>()));
        ADK_CUDA(c, cudaEventRecord(c->ev3, c->stream));
        k_ad_reduce<<<1, 32, 0, c->stream>>>(c->scratch1.as<CtaBest>(), grid, &sm->best);
        ++c->launches;
        ADK_CUDA(c, cudaGetLastError());
        ADK_CUDA(c, cudaMemcpyAsync(&hb, &sm->best, sizeof(hb), cudaMemcpyDeviceToHost, c->stream));
        ADK_CUDA(c, cudaMemcpyAsync(&hc, &sm->counters, sizeof(hc), cudaMemcpyDeviceToHost, c->stream));
    }
    ADK_CUDA(c, cudaEventRecord(c->ev1, c->stream));
    ADK_CUDA(c, cudaEventSynchronize(c->ev1));
    float ms = 0.f;
    ADK_CUDA(c, cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    if (ps.nqueued > 0) ADK_CUDA(c, cudaEventElapsedTime(&dp_ms, c->ev2, c->ev3));
    if (hb.found)
    {
        out->found = 1; out->count = hb.count; out->area = hb.area; out->tie_key = hb.key; out->w0 = hb.s; out->w1 = hb.t;
    }
    if (st)
    {
        st->cells_nominal = hc.cells; st->cells_executed = hc.cells; st->states = hc.states; st->pairs = ps.npairs_all;
        st->kernel_ms = ms; st->dp_ms = dp_ms; st->launches = c->launches; st->arena_bytes = c->arena.cap;
    }
    return ADK_OK;
}

int ad_finish(Ctx* c, uint64_t kmax, double amax, double dmax, const adk_partial* best, uint32_t flags, adk_result* out, uint32_t* hull)
{
    (void) dmax;
    const uint32_t n = c->n;
    out->found = best->found;
    if (!best->found) return ADK_OK;
    out->count = best->count; out->area = best->area;
    out->diam_valid = 1; out->diam[0] = best->w0; out->diam[1] = best->w1;      // AntipodalOptimized.cpp:516
    if (!(flags & ADK_FLAG_RECONSTRUCT)) return ADK_OK;
    if (best->w0 >= n || best->w1 >= n || best->w0 == best->w1 || best->count < 3) { c->err = "ad_finish: bad witness"; return ADK_ERR_ARG; }
    const uint32_t nlayers_ref = (uint32_t) (std::min<uint64_t>(n, kmax) - 1);
    ADK_TRY(run_count_tables(c));
    ADK_TRY(ensure(c, c->small, sizeof(AdSmall)));
    AdSmall* sm = c->small.as<AdSmall>();
    // the winning candidate alone: size it, re-run it keeping the tables (:529-535), walk back, clean up
    ADK_TRY(ensure(c, c->scratch6, sizeof(int32_t) * 3 * ((size_t) n + 4) + sizeof(PairRec) + 64));
    int32_t* d_left = c->scratch6.as<int32_t>();
    int32_t* d_right = d_left + (n + 4);
    int32_t* d_hull = d_right + (n + 4);
    PairRec* d_pair = reinterpret_cast<PairRec*>(d_hull + (n + 4));
    const size_t smem_ll = lune_smem_bytes(n + 2, n + 2, 1);
    if (smem_ll > 200 * 1024) { c->err = "ad_finish: point set too large for the lune kernel"; return ADK_ERR_NOMEM; }
    ADK_CUDA(c, cudaFuncSetAttribute(k_ad_lune_lists, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem_ll));
    k_ad_lune_lists<<<1, kAdThreads, smem_ll, c->stream>>>(c->x.as<double>(), c->y.as<double>(), n, best->w0, best->w1, d_left, d_right, sm->lune_cnt);
    ++c->launches;
    ADK_CUDA(c, cudaGetLastError());
    uint32_t cnt[2];
    ADK_CUDA(c, cudaMemcpyAsync(cnt, sm->lune_cnt, sizeof(cnt), cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaStreamSynchronize(c->stream));
    PairRec pr{best->w0, best->w1, cnt[0], cnt[1]};
    const uint64_t inner = (uint64_t) pr.nl + pr.nr - 4;
    const uint32_t lim = (uint32_t) std::min<uint64_t>(kmax - 2, inner);
    const uint32_t kopt = best->count - 2;
    if (kopt > lim) { c->err = "ad_finish: witness count exceeds the candidate's capacity"; return ADK_ERR_ARG; }
    const unsigned long long S = (unsigned long long) pr.nl * pr.nr * pr.nr + (unsigned long long) pr.nr * pr.nl * pr.nl;
    uint32_t grid = 0;
    ADK_TRY(ad_prepare_arena(c, S * (lim + 1), lim + 1, S, 1, &grid));
    ADK_CUDA(c, cudaMemcpyAsync(d_pair, &pr, sizeof(pr), cudaMemcpyHostToDevice, c->stream));
    ADK_CUDA(c, cudaMemsetAsync(&sm->counters, 0, sizeof(AdCounters), c->stream));
    ADK_TRY(ensure(c, c->scratch1, sizeof(CtaBest) * 4));
    ADK_TRY(ad_launch_dp<true>(c, d_pair, 1, 1, pr.nl, pr.nr, lim + 1, kmax, nlayers_ref, amax, &sm->counters, c->scratch1.as<CtaBest>()));
    Tables T{c->below.as<int32_t>(), c->coll.as<int32_t>(), c->rb.as<int32_t>(), n};
    const size_t smem = lune_smem_bytes(pr.nl, pr.nr, 1);
    ADK_CUDA(c, cudaFuncSetAttribute(k_ad_backtrack, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    k_ad_backtrack<<<1, kAdThreads, smem, c->stream>>>(c->x.as<double>(), c->y.as<double>(), n, T, pr, c->arena.as<unsigned char>(),
                                                       c->ad_slice_bytes, c->ad_V, c->ad_Q, c->ad_L, kopt, nlayers_ref,
                                                       d_left, d_right, d_hull, &sm->hull);
    ++c->launches;
    ADK_CUDA(c, cudaGetLastError());
    k_ad_clean<<<(unsigned) std::min<size_t>(((size_t) ((S + 31) / 32) * (lim + 1) + 255) / 256, 2048), 256, 0, c->stream>>>(
        c->arena.as<unsigned char>(), c->ad_slice_bytes, c->ad_V, c->ad_Q, c->ad_L, (uint32_t) S, lim);
    ++c->launches;
    ADK_CUDA(c, cudaGetLastError());
    AdHull hh;
    ADK_TRY(ensure_pinned(c, sizeof(int32_t) * ((size_t) n + 4)));
    ADK_CUDA(c, cudaMemcpyAsync(&hh, &sm->hull, sizeof(hh), cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaMemcpyAsync(c->h_pinned, d_hull, sizeof(int32_t) * ((size_t) n + 4), cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaStreamSynchronize(c->stream));
    if (!hh.ok) { c->err = "ad_finish: reconstruction could not re-derive a predecessor"; return ADK_ERR_ARG; }
    out->hull_len = hh.len;
    const int32_t* hp = static_cast<const int32_t*>(c->h_pinned);
    for (uint32_t t = 0; t < hh.len && t < n; ++t) hull[t] = (uint32_t) hp[t];
    return ADK_OK;
}

int ad_lune_lists(Ctx* c, uint32_t s, uint32_t t, int32_t* h_left, uint32_t* nl, int32_t* h_right, uint32_t* nr)
{
    const uint32_t n = c->n;
    ADK_TRY(ensure(c, c->scratch6, sizeof(int32_t) * 3 * ((size_t) n + 4) + 128));
    ADK_TRY(ensure(c, c->small, sizeof(AdSmall)));
    AdSmall* sm = c->small.as<AdSmall>();
    int32_t* d_left = c->scratch6.as<int32_t>();
    int32_t* d_right = d_left + (n + 4);
    const size_t smem = lune_smem_bytes(n + 2, n + 2, 1);
    if (smem > 200 * 1024) { c->err = "lune_lists: point set too large"; return ADK_ERR_NOMEM; }
    ADK_CUDA(c, cudaFuncSetAttribute(k_ad_lune_lists, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    k_ad_lune_lists<<<1, kAdThreads, smem, c->stream>>>(c->x.as<double>(), c->y.as<double>(), n, s, t, d_left, d_right, sm->lune_cnt);
    ++c->launches;
    ADK_CUDA(c, cudaGetLastError());
    uint32_t cnt[2];
    ADK_CUDA(c, cudaMemcpyAsync(cnt, sm->lune_cnt, sizeof(cnt), cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaStreamSynchronize(c->stream));
    ADK_CUDA(c, cudaMemcpyAsync(h_left, d_left, sizeof(int32_t) * cnt[0], cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaMemcpyAsync(h_right, d_right, sizeof(int32_t) * cnt[1], cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaStreamSynchronize(c->stream));
    *nl = cnt[0]; *nr = cnt[1];
    return ADK_OK;
}

}  // namespace adk
