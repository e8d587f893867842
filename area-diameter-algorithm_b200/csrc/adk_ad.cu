// adk_ad.cu — path AD: area- and diameter-bounded largest convex subset on sm_100a
// (MT::AntipodalOptimizedAlgorithm*, AntipodalOptimized.cpp:452-545; also serves MT::AntipodalAlgorithm*).
//
// B200-first reformulation (DESIGN.md "Path AD"):
//   * Candidate diameter pairs (s,t) are independent (AntipodalOptimized.cpp:495-527): a persistent grid pulls
//     them from a device queue, ONE CANDIDATE PER CTA.
//   * A state of the reference (side, o, s, ss) is "pending edge s->ss on one chain, head o on the other chain".
//     Grouping states by their two chain heads (a = head on L, b = head on R) gives a grid of CELLS (a,b):
//     every transition out of cell (a,b) advances exactly one head, so cells on one anti-diagonal a+b are independent
//     and every state vector "pending edge b->i with the other head at a" is produced ENTIRELY inside cell (a,b).
//     The reference's per-layer unordered_map (insert-or-keep-smaller, :207-217) therefore needs no atomics and no
//     global table: one warp owns a cell, lanes own destination vertices i, the minimum over predecessors is taken
//     in a shared-memory accumulator acc[layer][lane] and compacted into an append-only record pool.
//   * Geometry (the Orientation / PointLinePseudoDistance tests of :198-247, :271-324) depends on (source edge,
//     destination vertex) only, not on the layer k: it is evaluated once per pair and reused for every layer.
//   * The records stay in the pool, so the witness is recovered by re-running the winning candidate alone and
//     re-deriving, for each state, the predecessor whose relaxed value equals the stored one (no back-pointers).
#include "adk_internal.h"
#include "adk_device.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace adk
{

struct PairRec { uint32_t s, t; uint32_t nl, nr; };   // nl, nr include s and t themselves
constexpr int kCostBuckets = 96;

struct PairStats
{
    unsigned int npairs_all;        // candidates within the diameter bound (all shards)
    unsigned int nqueued;           // candidates of this shard that can hold at least one interior point
    unsigned int max_nl, max_nr, max_lim;
    unsigned int pad;
    unsigned long long max_cells;   // max over queued candidates of nl * nr                      (cells of the head grid)
    unsigned long long max_dirs;    // max over queued candidates of nl * nr * (nl + nr - 2)       (directory words)
    unsigned long long cost_sum, cost_max;   // sum / max over queued candidates of the cost proxy nl * nr * (nl + nr)
    unsigned int hist[kCostBuckets];    // queued candidates per cost bucket (heaviest candidates are started first)
    unsigned int cursor[kCostBuckets];
};

// Cost class of a candidate: 2 * log2(nl * nr * (nl + nr)), in half-octave steps.
__device__ __forceinline__ uint32_t cost_bucket(uint32_t nl, uint32_t nr)
{
    const unsigned long long cost = (unsigned long long) nl * nr * (nl + nr);
    const int l = 63 - __clzll((long long) cost);
    const uint32_t b = 2u * (uint32_t) l + (uint32_t) ((cost >> (l - 1)) & 1ull);     // cost >= 16: l >= 4
    return b < (uint32_t) kCostBuckets ? b : (uint32_t) kCostBuckets - 1;
}

struct CtaBest { uint32_t found; uint32_t count; double area; unsigned long long key; uint32_t s, t;
                 uint32_t nl, nr, slice, round, states, pad; };   // slice/round: where the winner's records still live (see ad_finish)

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

// Queues a candidate of this shard and folds it into the launch statistics.
__device__ __forceinline__ void queue_pair(PairRec* __restrict__ pairs, PairStats* __restrict__ stats, uint32_t i, uint32_t j,
                                           uint32_t nl, uint32_t nr, unsigned int lim)
{
    const uint32_t slot = atomicAdd(&stats->nqueued, 1u);
    pairs[slot] = PairRec{i, j, nl, nr};
    const unsigned long long cells = (unsigned long long) nl * nr;
    atomicMax(&stats->max_nl, nl); atomicMax(&stats->max_nr, nr); atomicMax(&stats->max_lim, lim);
    atomicMax(&stats->max_cells, cells); atomicMax(&stats->max_dirs, cells * (nl + nr - 2));
    atomicAdd(&stats->hist[cost_bucket(nl, nr)], 1u);
    const unsigned long long cost = cells * (nl + nr);
    atomicAdd(&stats->cost_sum, cost); atomicMax(&stats->cost_max, cost);
}

// One warp per pair (i < j) within the diameter bound: count its lune, size its tables.  With one shard the pair is queued
// right away; with several (dense != nullptr) every rank writes the sizes of ALL candidates into a dense n x n table plus a
// per-row histogram of cost classes, and k_ad_deal deals them out (below).
__global__ void k_ad_pairs(const double* __restrict__ x, const double* __restrict__ y, uint32_t n, double dmax2,
                           uint64_t kmax, PairRec* __restrict__ pairs, PairStats* __restrict__ stats,
                           uint32_t* __restrict__ dense, uint32_t* __restrict__ rowhist)
{
    __shared__ uint32_t s_hist[kCostBuckets];
    const uint32_t i = blockIdx.x;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    if (dense) { for (uint32_t b = threadIdx.x; b < (uint32_t) kCostBuckets; b += blockDim.x) s_hist[b] = 0; __syncthreads(); }
    const double sx = x[i], sy = y[i];
    for (uint32_t j = i + 1 + warp; j < n; j += nwarps)
    {
        const double tx = x[j], ty = y[j];
        const double d2 = dist2(sx, sy, tx, ty);
        if (!(d2 <= dmax2)) { if (dense && lane == 0) dense[(size_t) i * n + j] = 0; continue; }      // AntipodalOptimized.cpp:500
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
            if (dense)
            {
                dense[(size_t) i * n + j] = lim > 0 ? (nl | (nr << 16)) : 0u;
                if (lim > 0) atomicAdd(&s_hist[cost_bucket(nl, nr)], 1u);
            }
            else if (lim > 0) queue_pair(pairs, stats, i, j, nl, nr, (unsigned int) lim);
        }
    }
    if (dense)
    {
        __syncthreads();
        for (uint32_t b = threadIdx.x; b < (uint32_t) kCostBuckets; b += blockDim.x) rowhist[(size_t) i * kCostBuckets + b] = s_hist[b];
    }
}

// Deals the candidates over the shards of a multi-GPU solve.  Every rank sees the same dense table, so every rank can
// compute each candidate's position in the common order (cost class descending, then pair key ascending) with integer
// arithmetic only, and keeps the candidates whose position is congruent to its shard: a round-robin over the sorted
// list.  The loads of the shards then differ by well under 1 % (by key modulo they differed by 5-22 % at 8 shards on
// the bench sweep), and no rank needs to talk to another one.  One CTA per row i; warp 0 walks the row in key order.
__global__ void k_ad_deal(uint32_t n, uint64_t kmax, uint32_t shard, uint32_t nshards, const uint32_t* __restrict__ dense,
                          const uint32_t* __restrict__ rowhist, PairRec* __restrict__ pairs, PairStats* __restrict__ stats)
{
    __shared__ uint32_t s_run[kCostBuckets];     // candidates of the class in earlier rows (then: earlier in key order)
    __shared__ uint32_t s_off[kCostBuckets];     // candidates of heavier classes
    const uint32_t i = blockIdx.x, lane = threadIdx.x & 31;
    if (threadIdx.x < (uint32_t) kCostBuckets)
    {
        uint32_t before = 0, total = 0;
        for (uint32_t r = 0; r < n; ++r)
        {
            const uint32_t h = rowhist[(size_t) r * kCostBuckets + threadIdx.x];
            before += r < i ? h : 0u; total += h;
        }
        s_run[threadIdx.x] = before; s_off[threadIdx.x] = total;
    }
    __syncthreads();
    if (threadIdx.x == 0)
    {
        uint32_t run = 0;
        for (int b = kCostBuckets - 1; b >= 0; --b) { const uint32_t t = s_off[b]; s_off[b] = run; run += t; }
    }
    __syncthreads();
    if (threadIdx.x >= 32) return;
    for (uint32_t j0 = i + 1; j0 < n; j0 += 32)
    {
        const uint32_t j = j0 + lane;
        const uint32_t v = j < n ? dense[(size_t) i * n + j] : 0u;
        const uint32_t nl = v & 0xffffu, nr = v >> 16;
        const uint32_t cls = v ? cost_bucket(nl, nr) : 0xffffffffu;
        const uint32_t peers = __match_any_sync(0xffffffffu, cls);
        const uint32_t r = __popc(peers & ((1u << lane) - 1u));
        uint32_t pos = 0;
        if (v) pos = s_off[cls] + s_run[cls] + r;
        __syncwarp();
        if (v && r + 1 == (uint32_t) __popc(peers)) s_run[cls] += r + 1;
        __syncwarp();
        if (v && pos % nshards == shard)
        {
            const uint64_t inner = (uint64_t) nl + nr - 4;
            queue_pair(pairs, stats, i, j, nl, nr, (unsigned int) (kmax - 2 < inner ? kmax - 2 : inner));
        }
    }
}

// Processing order of the queue: heaviest cost class first (a counting sort over kCostBuckets classes), so that the
// long candidates do not end up alone at the tail of the persistent grid.
__global__ void k_ad_order(const PairRec* __restrict__ pairs, uint32_t npairs, PairStats* __restrict__ stats, uint32_t* __restrict__ order)
{
    __shared__ uint32_t s_base[kCostBuckets];
    if (threadIdx.x == 0)
    {
        uint32_t run = 0;
        for (int b = kCostBuckets - 1; b >= 0; --b) { s_base[b] = run; run += stats->hist[b]; }
    }
    __syncthreads();
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < npairs; t += gridDim.x * blockDim.x)
    {
        const uint32_t b = cost_bucket(pairs[t].nl, pairs[t].nr);
        order[s_base[b] + atomicAdd(&stats->cursor[b], 1u)] = t;
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


// --------------------------------------------------------------------------------------------- the cell wavefront

// Cell (a,b): a = index of the current vertex ("head") of the left chain in L, b = head of the right chain in R.
//   RIGHT(a; s->b) : pending edge R[s]->R[b], left head L[a]   == reference Entry{o=a, s, ss=b, side RR/RL}
//   LEFT (b; s->a) : pending edge L[s]->L[a], right head R[b]  == reference Entry{o=b, s, ss=a, side LL/LR}
// Both kinds live in cell (a,b).  Expanding them (AntipodalOptimized.cpp:177-325) produces RIGHT(a; b->i), i > b
// [cell (a,i)] and LEFT(b; a->i), i > a [cell (i,b)]: destination d of cell (a,b) is
//   d <  nr-1-b : RIGHT(a; b -> b+1+d)          d >= nr-1-b : LEFT(b; a -> a+1+(d-(nr-1-b)))
// and the whole layer vector of a destination is produced inside this one cell.
//
// The suffix sweep (round 2).  The sources of a cell fall into two groups, the RIGHT edges R[s]->R[b] and the LEFT edges
// L[s]->L[a].  Sort each group by the direction of its edge around the head (a pseudo-angle in the frame of the candidate
// segment, sources in the half plane in front of the head).  Every transition test of the reference is a half-plane test on
// that direction: "stay" is Orientation(q, head, p_s) >= 0 (:201, :274) and "switch" compares the pseudo-distances of q and of
// the other head from the line p_s -> head (:226, :299), so for ONE destination the sources that pass form a SUFFIX of the
// sorted group (measured: every one of 8.5e7 (destination, group) sets on the bench instance; tools/ad_interval_stats.c).
// The minimum over the passing sources is then a suffix minimum: per cell ONE backward sweep over each group with the
// running minimum in registers (lanes = layers) serves every destination, instead of one relaxation per passing pair:
//   dest(d)[k] = min( SufR[loR(d)][k], SufL[loL(d)][k] ),   Suf_g[r][k] = min over sorted positions p >= r of S_p[k],
//   S_p[k] = area_p[k - c_p] + triangle_p (<= A).
// The tests themselves are still the reference's, evaluated pair by pair in FP64 — a destination whose passing set is NOT
// a suffix (possible only for eps-degenerate inputs) is relaxed pair by pair ("explicit" path), so the result never
// depends on the sort.  Two more consequences: the "stay" tests depend on (s, b, i) only, not on the other head, so their
// suffix starts are tabulated once per candidate; and destinations with equal (loR, loL) have equal layer vectors, which
// are stored once and shared (3.6x fewer records on the bench instance).

#ifndef ADK_CELL_WARPS
#define ADK_CELL_WARPS 4
#endif
constexpr int kCellWarps = ADK_CELL_WARPS;       // warps per CTA (one candidate per CTA) when candidates are plentiful
constexpr int kMidCellWarps = 10, kMaxCellWarps = 20;   // ... and when few, heavy candidates set the critical path (chosen per launch):
                                                 // 2 x 10 or 1 x 20 warps per SM, the same 20 warps and 96 registers as 5 x 4
constexpr int kCellThreads = kCellWarps * 32;
#ifndef ADK_CELL_CTAS
#define ADK_CELL_CTAS 5
#endif
constexpr int kCellCtas = ADK_CELL_CTAS;         // resident CTAs per SM the 4-warp kernel is compiled for (registers per thread follow)
constexpr int kBT = 32;                          // layers per tile = lanes
constexpr int kGrp = 32;                         // sources per group in shared memory
#ifndef ADK_TAB_ROWS
#define ADK_TAB_ROWS 16
#endif
constexpr int kTabRows = ADK_TAB_ROWS;           // rows of the suffix table in shared memory: one per DISTINCT suffix start of the LEFT group that a leader needs
constexpr int kDstCap = 128;                     // destinations per pass with their state in shared memory
constexpr uint32_t kFB = 0xffffu;                // suffix start "none": the destination takes the explicit path for that group
constexpr uint32_t kNil = 0xffffu;

// Layer vector of one (cell, destination): areas of layers kmin .. kmin+len-1 at vpool[off ..], +inf where unreached; nfin of
// them are finite (the states the reference holds for this pending edge: the cell-update and entry counters).  One 64-bit
// directory word: off (28 bits) | kmin (12) | len (12) | nfin (12); layers < 4096 = ADK_MAX_POINTS, pools <= 2^28 records.
struct DirEnt { uint32_t off; uint32_t kmin, len, nfin; };
__device__ __forceinline__ unsigned long long dir_pack(const DirEnt& d)
{
    return (unsigned long long) d.off | ((unsigned long long) d.kmin << 28) | ((unsigned long long) d.len << 40) | ((unsigned long long) d.nfin << 52);
}
__device__ __forceinline__ DirEnt dir_unpack(unsigned long long w)
{
    DirEnt d; d.off = (uint32_t) (w & 0xfffffffull); d.kmin = (uint32_t) (w >> 28) & 0xfffu; d.len = (uint32_t) (w >> 40) & 0xfffu; d.nfin = (uint32_t) (w >> 52);
    return d;
}
constexpr uint32_t kMaxPoolRecords = (1u << 28) - 32u;
// One live source edge of a cell, in two records: what the transition tests read (first vertex of the edge, pseudo-distance
// of the other chain's head from the edge's line, :222 / :295) and what the sweeps read.  ceff = points added by the edge's
// fan triangle (:188-190, with the extra collinear term of a layer-0 seed); pos = position in the sorted order of ALL s < head.
struct SrcGeo { double refd, psx, psy; };
struct SrcVal { double tri; uint32_t off; uint16_t len, kmin, ceff, pos; };
// Per chain edge (s -> h): area and enclosed points of its fan triangle (:188-190 / :261-263), once per candidate.
struct EdgeRec { double tri; uint16_t c, cx; uint32_t pad; };
static_assert(sizeof(SrcGeo) == 24 && sizeof(SrcVal) == 24 && sizeof(EdgeRec) == 16, "record layouts");
template <bool SPILL> struct LoType { typedef uint8_t type; };      // suffix starts / list links: bytes while a group has <= kGrp sources
template <> struct LoType<true> { typedef uint16_t type; };

// Per-CTA slice of the arena (two per CTA: the one being written and the one holding the CTA's best candidate so far).
//   dir   PRODUCER-major: cell (a,b) owns the W = nl+nr-2 words at (a*nr+b)*W; word d belongs to its destination d.  A cell
//         writes all its words with coalesced stores; a consumer reads only the words of its LIVE sources.
//   live  CONSUMER-major bitmaps: cell (a,b) owns WB = ceil(nr/32)+ceil(nl/32) words; bit u of the first part = the RIGHT
//         edge at sorted position u around R[b] carries states, second part = the same for the LEFT edges around L[a].
//         Producers set the bits (atomicOr) of their non-empty destinations; empty ones cost the consumer nothing.
struct CellSlice
{
    unsigned char* alive;         // nl*nr flags: cell wrote its directory words (read by the witness walk)
    unsigned long long* dir;
    uint32_t* live;
    double* vpool;                // layer vectors
};

struct CellLayout { size_t off_dir, off_live, off_v, bytes; uint32_t cap, pad; };

__host__ __device__ inline uint32_t live_words(uint32_t nl, uint32_t nr) { return (nr + 31) / 32 + (nl + 31) / 32; }

__host__ __device__ inline CellLayout cell_layout(unsigned long long max_cells, unsigned long long max_dirs, uint32_t max_nl, uint32_t max_nr, size_t slice_bytes)
{
    CellLayout Lo;
    auto up = [](size_t v) { return (v + 255) & ~(size_t) 255; };
    size_t o = up((size_t) max_cells);
    Lo.off_dir = o; o = up(o + sizeof(unsigned long long) * (size_t) max_dirs);
    Lo.off_live = o; o = up(o + sizeof(uint32_t) * (size_t) max_cells * live_words(max_nl, max_nr));
    Lo.off_v = o;
    const size_t rest = slice_bytes > o + 4096 ? slice_bytes - o - 4096 : 0;
    const size_t cap = rest / sizeof(double);
    Lo.cap = (uint32_t) (cap > kMaxPoolRecords ? kMaxPoolRecords : cap) & ~31u;
    Lo.bytes = slice_bytes; Lo.pad = 0;
    return Lo;
}

__device__ __forceinline__ CellSlice carve_cells(unsigned char* arena, const CellLayout& Lo, uint32_t cta)
{
    unsigned char* base = arena + (size_t) cta * Lo.bytes;
    CellSlice s;
    s.alive = base;
    s.dir = reinterpret_cast<unsigned long long*>(base + Lo.off_dir);
    s.live = reinterpret_cast<uint32_t*>(base + Lo.off_live);
    s.vpool = reinterpret_cast<double*>(base + Lo.off_v);
    return s;
}

// Per-CTA auxiliary region (not kept between candidates): the per-candidate chain tables and one spill workspace per warp
// for the cells that outgrow the shared-memory workspace.  With tri(h) = h(h-1)/2:
//   sigR[tri(b) + r]  = index s < b at sorted position r around R[b]; invR[tri(b) + s] = r          (sigL, invL: around L[a])
//   edgeR[tri(b) + s] = fan triangle of the edge R[s]->R[b]
//   stayR[b*nr + i]   = first sorted position whose edge R[s]->R[b] may continue to R[i] (b = none, kFB = not a suffix)
struct AuxLayout { size_t off_sig[2], off_inv[2], off_edge[2], off_stay[2], off_spill, spill_bytes, bytes; uint32_t grp_cap, dst_cap; };

// workspace of one warp: [suffix table: grp rows x 32 layers][sources: 2 groups x grp x 2 records][loR, loL, next: dst entries][heads: grp+1 x u32]
__host__ __device__ inline size_t warp_ws_bytes(uint32_t grp, uint32_t dst, uint32_t lo_bytes)
{
    return ((size_t) grp * 256 + (size_t) 2 * grp * (sizeof(SrcGeo) + sizeof(SrcVal)) + (size_t) dst * 3 * lo_bytes + (size_t) (grp + 1) * 4 + 15) & ~(size_t) 15;
}
constexpr size_t kWarpWsBytes = ((size_t) kTabRows * 256 + (size_t) 2 * kGrp * 48 + (size_t) kDstCap * 3 + (size_t) (kGrp + 1) * 4 + 15) & ~(size_t) 15;

__host__ __device__ inline AuxLayout aux_layout(uint32_t max_nl, uint32_t max_nr)
{
    AuxLayout A;
    auto up = [](size_t v) { return (v + 255) & ~(size_t) 255; };
    size_t o = 0;
    for (int ch = 0; ch < 2; ++ch)
    {
        const size_t nc = ch ? max_nl : max_nr, tri = nc * nc / 2 + 1;
        A.off_sig[ch] = o; o = up(o + sizeof(uint16_t) * tri);
        A.off_inv[ch] = o; o = up(o + sizeof(uint16_t) * tri);
        A.off_edge[ch] = o; o = up(o + sizeof(EdgeRec) * tri);
        A.off_stay[ch] = o; o = up(o + sizeof(uint16_t) * nc * nc);
    }
    A.off_spill = o;
    A.grp_cap = max_nl > max_nr ? max_nl : max_nr;
    A.dst_cap = (max_nl + max_nr + 3) & ~3u;
    A.spill_bytes = up(warp_ws_bytes(A.grp_cap, A.dst_cap, 2));
    A.bytes = up(o + A.spill_bytes * kMaxCellWarps);
    return A;
}

// The context of the candidate a CTA is working on.  ONE copy per CTA, in shared memory: every warp reads it from there, and the
// out-of-line paths take its address without forcing a per-thread copy into local memory.
struct CellCtx
{
    // candidate
    uint32_t nl, nr, lim, nlayers;
    double Sx, Sy, Ex, Ey, seg2, amax;
    int Sp, Ep;
    LuneSmem S;
    CellSlice sl;
    uint32_t cap;
    // per-candidate tables (global, per CTA)
    uint16_t* sig[2]; uint16_t* inv[2]; uint16_t* stay[2]; EdgeRec* edge[2];      // [0] = right chain, [1] = left chain
    uint32_t W, WR, WB;           // directory words per cell; bitmap words of the RIGHT part / of a cell
    // warp-private workspaces: shared memory (kGrp / kDstCap; warp w at ws_base + w * kWarpWsBytes) and the global spill
    // (spill_grp / spill_dst; warp w at spill_base + w * spill_bytes)
    unsigned char* ws_base;
    unsigned char* spill_base;
    size_t spill_bytes;
    uint32_t spill_grp, spill_dst;
    // CTA-shared
    volatile uint32_t* rowdone;   // rowdone[a] = cells of row a finished (they finish in b order)
    volatile uint32_t* coldone;   // coldone[b] = cells of column b finished (in a order)
    unsigned int* next;     // next cell of the diagonal-major order to hand out
    unsigned int* bump;     // pool entries allocated so far
    unsigned int* overflow;
};

constexpr unsigned kFull = 0xffffffffu;

// Direction of p seen from the head h as a monotone pseudo-angle in the frame of the candidate segment: x axis = sgn * (t - s)
// (towards the side of the chain the sources lie on), sources in (-1, 1), increasing counter-clockwise.  Only the ORDER it
// induces is used, and only for speed (see the header comment), so plain arithmetic is fine here.
__device__ __forceinline__ double pseudo_angle(double fx, double fy, double sgn, double hx, double hy, double px, double py)
{
    const double dx = px - hx, dy = py - hy;
    const double u = sgn * (dx * fx + dy * fy), v = sgn * (fx * dy - fy * dx);
    const double den = fabs(u) + fabs(v);
    if (den == 0.0) return 0.0;
    const double r = v / den;
    return u >= 0.0 ? r : (v >= 0.0 ? 2.0 - r : -2.0 - r);
}

// Builds, for one chain of the candidate, the sorted orders, the fan triangles of its edges and the "stay" suffix starts.
// All warps of the CTA take part (heads dealt round robin); `keys` is the calling warp's scratch (>= n doubles).
__device__ __noinline__ void build_chain_tables(const CellCtx& C, const Tables& T, bool left, double* keys, uint32_t lane, uint32_t warp, uint32_t nw)
{
    const LuneSmem& S = C.S;
    const uint32_t nc = left ? C.nl : C.nr;
    const double* X = left ? S.Lx : S.Rx; const double* Y = left ? S.Ly : S.Ry;
    const int32_t* I = left ? S.Lidx : S.Ridx;
    uint16_t* sig = C.sig[left]; uint16_t* inv = C.inv[left]; uint16_t* stay = C.stay[left]; EdgeRec* edge = C.edge[left];
    const double fx = C.Ex - C.Sx, fy = C.Ey - C.Sy, sgn = left ? -1.0 : 1.0;
    const double apx = left ? C.Ex : C.Sx, apy = left ? C.Ey : C.Sy;          // the chain's own fan apex (:188, :201 / :261, :274)
    const int ap = left ? C.Ep : C.Sp;
    for (uint32_t h = 1 + warp; h < nc; h += nw)
    {
        const double hx = X[h], hy = Y[h];
        const int hidx = I[h];
        uint16_t* sg = sig + h * (h - 1) / 2;
        for (uint32_t s = lane; s < h; s += 32)
        {
            const double psx = X[s], psy = Y[s];
            keys[s] = pseudo_angle(fx, fy, sgn, hx, hy, psx, psy);
            EdgeRec e;
            e.tri = fabs(signed_area(apx, apy, psx, psy, hx, hy));                                        // :188 / :261
            e.c = (uint16_t) (1u + (uint32_t) tri_count_ad(T, false, I[s], psx, psy, hidx, hx, hy, ap, apx, apy));
            e.cx = (uint16_t) T.C(I[s], ap);                                                              // extra term of a layer-0 seed (:33-36)
            e.pad = 0;
            edge[h * (h - 1) / 2 + s] = e;
        }
        __syncwarp();
        for (uint32_t s = lane; s < h; s += 32)
        {
            const double ks = keys[s];
            uint32_t r = 0;
#pragma unroll 1
            for (uint32_t t = 0; t < h; ++t) { const double kt = keys[t]; r += (kt < ks || (kt == ks && t < s)) ? 1u : 0u; }
            sg[r] = (uint16_t) s;
            inv[h * (h - 1) / 2 + s] = (uint16_t) r;
        }
        __syncwarp();
        for (uint32_t i0 = h + 1; i0 < nc; i0 += 32)
        {
            const uint32_t i = i0 + lane;
            if (i < nc)
            {
                const double qx = X[i], qy = Y[i];
                uint32_t first = h, cnt = 0;
                if (orientation(qx, qy, hx, hy, apx, apy) >= 0)
#pragma unroll 1
                    for (uint32_t r = 0; r < h; ++r)
                    {
                        const uint32_t s = sg[r];
                        if (orientation(qx, qy, hx, hy, X[s], Y[s]) >= 0) { ++cnt; first = min(first, r); }
                    }
                stay[h * nc + i] = (uint16_t) (cnt == 0 ? h : (cnt == h - first ? first : kFB));
            }
        }
        __syncwarp();
    }
}

// The complete transition test of one (source edge, destination) pair, exactly as the reference evaluates it
// (:198-201, :225-229, :271-274, :298-302).  left = the source is a LEFT edge; dR = the destination is on the right chain.
__device__ __forceinline__ bool dest_side_ok(const CellCtx& C, bool left, bool dR, uint32_t a, uint32_t b, double qx, double qy,
                                             double Lax, double Lay, double Rbx, double Rby)
{
    if (dR)
        return left ? (b == 0 || (orientation(Rbx, Rby, C.Sx, C.Sy, qx, qy) >= 0 && orientation(qx, qy, Rbx, Rby, C.Ex, C.Ey) >= 0))   // :301
                    : orientation(qx, qy, Rbx, Rby, C.Sx, C.Sy) >= 0;                                                                 // :201
    return left ? orientation(qx, qy, Lax, Lay, C.Ex, C.Ey) >= 0                                                                       // :274
                : (a == 0 || (orientation(Lax, Lay, C.Ex, C.Ey, qx, qy) >= 0 && orientation(qx, qy, Lax, Lay, C.Sx, C.Sy) >= 0));     // :228
}
__device__ __forceinline__ bool pair_ok(bool left, bool dR, double psx, double psy, double hx, double hy, double refd, double qx, double qy)
{
    if (left != dR) return orientation(qx, qy, hx, hy, psx, psy) >= 0;     // stay on the edge's chain (:201, :274)
    return pseudo_dist(psx, psy, hx, hy, qx, qy) <= refd;                   // switch chains (:226-227, :299-300)
}

// Minimum of two areas.  Areas are non-negative doubles or +inf, whose order is the order of their bit patterns as unsigned
// integers: an integer minimum issues at full rate, DMNMX at a fifth of the FP64 rate (tools/issue_peaks.cu: dmin_per_s).
__device__ __forceinline__ double amin(double a, double b)
{
    const unsigned long long x = (unsigned long long) __double_as_longlong(a), y = (unsigned long long) __double_as_longlong(b);
    return __longlong_as_double((long long) (x < y ? x : y));
}

// S_p[k]: what source r contributes to layer k of any destination it may continue to (:188-195, :203): area + triangle.
__device__ __forceinline__ double src_value(const SrcVal& r, const double* __restrict__ vpool, uint32_t k, bool kact, double amax)
{
    const int e = (int) k - (int) r.ceff - (int) r.kmin;
    double v = INFINITY;
    if (kact && e >= 0 && e < (int) r.len)
    {
        v = __dadd_rn(vpool[r.off + e], r.tri);
        if (!(v <= amax)) v = INFINITY;
    }
    return v;
}

// Minimum over the sources of one group that pass the tests of destination dd, pair by pair (lanes = layers): the path of a
// destination whose passing sources are not a suffix of the sorted group (eps-degenerate inputs only).
__device__ __noinline__ double explicit_min(const CellCtx& C, bool left, uint32_t dd, uint32_t k, bool kact, uint32_t a, uint32_t b,
                                            const SrcGeo* sg, const SrcVal* sv, uint32_t P)
{
    const LuneSmem& S = C.S;
    const uint32_t nDR = C.nr - 1 - b;
    const bool dR = dd < nDR;
    const uint32_t i = dR ? b + 1 + dd : a + 1 + (dd - nDR);
    const double qx = dR ? S.Rx[i] : S.Lx[i], qy = dR ? S.Ry[i] : S.Ly[i];
    const double Lax = S.Lx[a], Lay = S.Ly[a], Rbx = S.Rx[b], Rby = S.Ry[b];
    double v = INFINITY;
    if (!dest_side_ok(C, left, dR, a, b, qx, qy, Lax, Lay, Rbx, Rby)) return v;
    const double hx = left ? Lax : Rbx, hy = left ? Lay : Rby;
#pragma unroll 1
    for (uint32_t p = 0; p < P; ++p)
    {
        const SrcGeo g = sg[p];
        if (pair_ok(left, dR, g.psx, g.psy, hx, hy, g.refd, qx, qy)) v = amin(v, src_value(sv[p], C.sl.vpool, k, kact, C.amax));
    }
    return v;
}

// One warp expands every state of cell (a,b).  bk/bv: this lane's best terminal (layer, area).
// SPILL = false: the warp's workspace is in shared memory (at most kGrp sources per group); returns false, with nothing
// written, when the cell does not fit.  SPILL = true: the same code over the warp's global workspace.
template <bool SPILL>
__device__ __forceinline__ bool process_cell(const CellCtx& C, uint32_t a, uint32_t b, uint32_t lane,
                                             uint32_t& bk, double& bv, unsigned long long& my_cells, unsigned int& my_states)
{
    typedef typename LoType<SPILL>::type lo_t;
    constexpr uint32_t FB = SPILL ? 0xffffu : 0xffu;     // suffix start "not a suffix" / list end, in lo_t
    const uint32_t nl = C.nl, nr = C.nr, W = C.W;
    const LuneSmem& S = C.S;
    const double Lax = S.Lx[a], Lay = S.Ly[a], Rbx = S.Rx[b], Rby = S.Ry[b];
    const bool term_r = b == nr - 1, term_l = a == nl - 1;               // R[nr-1] = s, L[nl-1] = t   (:167-175)
    const bool dead = dist2(Lax, Lay, Rbx, Rby) > C.seg2;                // :183-186 / :256-259
    const uint32_t nDR = nr - 1 - b, nDL = nl - 1 - a, D = nDR + nDL;
    const uint32_t lt = (1u << lane) - 1u;
    const uint32_t dmax = min(C.lim, C.nlayers - 1);
    const double* __restrict__ vpool = C.sl.vpool;
    const uint32_t cell = a * nr + b;

    unsigned char* wbase = SPILL ? C.spill_base + C.spill_bytes * (threadIdx.x >> 5) : C.ws_base + kWarpWsBytes * (threadIdx.x >> 5);
    const uint32_t gcap = SPILL ? C.spill_grp : (uint32_t) kGrp;
    const uint32_t dcap = SPILL ? C.spill_dst : (uint32_t) kDstCap;
    const uint32_t trows = SPILL ? C.spill_grp : (uint32_t) kTabRows;
    double* tab = reinterpret_cast<double*>(wbase);                              // suffix minima of the LEFT group, [row][layer lane]
    SrcGeo* geo = reinterpret_cast<SrcGeo*>(wbase + (size_t) trows * 256);       // [0, gcap): RIGHT group, [gcap, 2 gcap): LEFT group
    SrcVal* val = reinterpret_cast<SrcVal*>(geo + 2 * gcap);
    lo_t* loRs = reinterpret_cast<lo_t*>(val + 2 * gcap);
    lo_t* loLs = loRs + dcap;
    lo_t* nxt = loLs + dcap;
    uint32_t* head = reinterpret_cast<uint32_t*>(nxt + dcap);

    // ---- gather: the pending edges that reach this cell, RIGHT(a; s->b) from cell (a,s), LEFT(b; s->a) from cell (s,b).
    //      The cell's bitmaps name the live ones by sorted position, so the compaction keeps each group in its sorted order.
    //      Blocks of 32 positions, RIGHT group first; the directory words of the next non-empty block are requested before
    //      the current one is worked on (L2 / HBM round trips; the geometry below hides them).
    uint32_t PR = 0, PL = 0;
    uint32_t klo = 0xffffffffu, khi = 0;
    unsigned long long cells_add = 0;
    {
        const uint32_t* bmp = C.sl.live + (size_t) cell * C.WB;
        const uint32_t WR = C.WR, WB = C.WB;
        auto fetch = [&](uint32_t wi, uint32_t m, uint32_t& s, unsigned long long& w, EdgeRec& e)
        {
            if ((m >> lane) & 1u)
            {
                const bool left = wi >= WR;
                const uint32_t h = left ? a : b, u = (left ? wi - WR : wi) * 32 + lane;
                s = C.sig[left][h * (h - 1) / 2 + u];
                // RIGHT(a; s->b) is destination b-s-1 of cell (a,s); LEFT(b; s->a) is destination (nr-1-b) + a-s-1 of cell (s,b)
                w = C.sl.dir[left ? (size_t) (s * nr + b) * W + (nr - 1 - b) + (a - s - 1) : (size_t) (a * nr + s) * W + (b - s - 1)];
                e = C.edge[left][h * (h - 1) / 2 + s];
            }
        };
        uint32_t P = 0;
        bool in_left = false;
#pragma unroll 1
        for (uint32_t w0 = 0; w0 < WB; w0 += 32)
        {   // the cell's bitmap words in one load (the producers set the bits with L2 atomics: read them around L1)
            const uint32_t bw = w0 + lane < WB ? __ldcg(bmp + w0 + lane) : 0u;
            uint32_t nzm = __ballot_sync(kFull, bw != 0u);
            uint32_t wi_n = 0, m_n = 0, s_n = 0; unsigned long long w_n = 0ull; EdgeRec e_n; e_n.tri = 0.0; e_n.c = 0; e_n.cx = 0; e_n.pad = 0;
            bool have_n = nzm != 0u;
            if (have_n)
            {
                const int z = __ffs((int) nzm) - 1; nzm &= nzm - 1u;
                wi_n = w0 + (uint32_t) z; m_n = __shfl_sync(kFull, bw, z);
                fetch(wi_n, m_n, s_n, w_n, e_n);
            }
#pragma unroll 1
            while (have_n)
            {
                const uint32_t q = wi_n, m = m_n, s = s_n;
                const unsigned long long w = w_n; const EdgeRec er = e_n;
                have_n = nzm != 0u;
                if (have_n)
                {   // request the next non-empty block before this one is worked on
                    const int z = __ffs((int) nzm) - 1; nzm &= nzm - 1u;
                    wi_n = w0 + (uint32_t) z; m_n = __shfl_sync(kFull, bw, z);
                    fetch(wi_n, m_n, s_n, w_n, e_n);
                }
                const uint32_t nbR = WR;
            const bool left = q >= nbR;
            if (left && !in_left) { in_left = true; PR = P; P = 0; }
            const uint32_t u = (left ? q - nbR : q) * 32 + lane;
            bool live = (m >> lane) & 1u;
            const DirEnt de = dir_unpack(w);
            if (live)
            {   // pull the layer vector towards L1 while the geometry below is computed
                const char* v0 = reinterpret_cast<const char*>(vpool + de.off);
                const char* v1 = reinterpret_cast<const char*>(vpool + de.off + de.len - 1);
                asm volatile("prefetch.global.L1 [%0];" ::"l"(v0));
                asm volatile("prefetch.global.L1 [%0];" ::"l"(v1));
                if (v1 - v0 > 256) asm volatile("prefetch.global.L1 [%0];" ::"l"(v0 + 128));
            }
            const uint32_t kmax = de.kmin + de.len - 1;
            if (live && (left ? term_l : term_r))
            {   // terminal states only report (:167-175): the deepest layer of this edge is its best
                if (kmax > 0)
                {
                    const double v = vpool[de.off + de.len - 1];
                    if (kmax > bk || (kmax == bk && v < bv)) { bk = kmax; bv = v; }
                }
                live = false;
            }
            if (dead) live = false;
            SrcGeo rg; rg.refd = 0.0; rg.psx = 0.0; rg.psy = 0.0;
            SrcVal rv; rv.tri = er.tri; rv.off = de.off; rv.len = (uint16_t) de.len; rv.kmin = (uint16_t) de.kmin; rv.ceff = 0; rv.pos = (uint16_t) u;
            if (live)
            {
                const double psx = left ? S.Lx[s] : S.Rx[s], psy = left ? S.Ly[s] : S.Ry[s];
                rg.psx = psx; rg.psy = psy;
                const uint32_t ceff = (uint32_t) er.c + (de.kmin == 0 ? (uint32_t) er.cx : 0u);   // only the seeds sit on layer 0 (:33-36)
                rv.ceff = (uint16_t) ceff;
                rg.refd = left ? pseudo_dist(psx, psy, Lax, Lay, Rbx, Rby) : pseudo_dist(psx, psy, Rbx, Rby, Lax, Lay);   // :222 / :295
                // cell updates the reference performs for these states (:198,225,271,298), and the output layer range
                uint32_t nvalid = de.nfin;
                if (kmax + ceff >= C.nlayers)
                {
                    nvalid = 0;
#pragma unroll 1
                    for (uint32_t e = 0; e < de.len; ++e)
                        nvalid += (vpool[de.off + e] < INFINITY && de.kmin + e + ceff < C.nlayers) ? 1u : 0u;
                }
                cells_add += (unsigned long long) nvalid * D;
                const uint32_t lo = de.kmin + ceff, hi = min(kmax + ceff, dmax);
                if (nvalid == 0 || lo > hi || !(rv.tri <= C.amax)) live = false;
                else { klo = min(klo, lo); khi = max(khi, hi); }
            }
            const uint32_t lm = __ballot_sync(kFull, live);
            const uint32_t at = P + __popc(lm & lt);
            P += __popc(lm);
            if (!SPILL && P > gcap) return false;          // does not fit in shared memory: nothing has been written yet
            const uint32_t obase = left ? gcap : 0u;
            if (live) { geo[obase + at] = rg; val[obase + at] = rv; }
            }
        }
        if (in_left) PL = P; else PR = P;
    }
    my_cells += cells_add;
    if (PR + PL == 0 || D == 0) return true;
    klo = __reduce_min_sync(kFull, klo); khi = __reduce_max_sync(kFull, khi);
    __syncwarp();
    const uint32_t ntiles = (khi - klo + kBT) / kBT;
    unsigned long long* mydir = C.sl.dir + (size_t) cell * W;           // this cell's directory words, one per destination

#pragma unroll 1
    for (uint32_t dc0 = 0; dc0 < D; dc0 += dcap)
    {
        const uint32_t nd = min(dcap, D - dc0);
#pragma unroll 1
        for (uint32_t t = lane; t <= PR; t += 32) head[t] = FB;
        __syncwarp();
        uint32_t needL = 0;          // suffix starts of the LEFT group that some leader reads (bit = position; shared-memory mode only)
        // ---- suffix starts: lanes = destinations
#pragma unroll 1
        for (uint32_t d0 = 0; d0 < nd; d0 += 32)
        {
            const uint32_t dl = d0 + lane, dd = dc0 + dl;
            const bool have = dl < nd;
            uint32_t loR = PR, loL = PL;
            if (have)
            {
                const bool dR = dd < nDR;
                const uint32_t i = dR ? b + 1 + dd : a + 1 + (dd - nDR);
                const double qx = dR ? S.Rx[i] : S.Lx[i], qy = dR ? S.Ry[i] : S.Ly[i];
#pragma unroll 1
                for (int g = 0; g < 2; ++g)
                {
                    const bool left = g == 1;
                    const uint32_t P = left ? PL : PR;
                    const uint32_t obase = left ? gcap : 0u;
                    uint32_t lo = P;
                    if (P == 0) { }
                    else if (left != dR)
                    {   // stay: tabulated over ALL s < head (the tests do not depend on the other head); count the live sources before it
                        const uint32_t lf = left ? C.stay[1][a * nl + i] : C.stay[0][b * nr + i];
                        if (lf == kFB) lo = FB;
                        else
                        {   // lower bound of lf in the (ascending) sorted positions of the live sources
                            uint32_t base = 0, len = P;
#pragma unroll 1
                            while (len)
                            {
                                const uint32_t half = len >> 1;
                                if (val[obase + base + half].pos < lf) { base += half + 1; len -= half + 1; } else len = half;
                            }
                            lo = base;
                        }
                    }
                    else if (dest_side_ok(C, left, dR, a, b, qx, qy, Lax, Lay, Rbx, Rby))
                    {   // switch: one pseudo-distance test per live source (:226-227, :299-300)
                        const double hx = left ? Lax : Rbx, hy = left ? Lay : Rby;
                        const SrcGeo* gp = geo + obase;
                        uint32_t first = P, cnt = 0;
#pragma unroll 2
                        for (uint32_t p = 0; p < P; ++p)
                        {
                            const SrcGeo q = gp[p];
                            if (pseudo_dist(q.psx, q.psy, hx, hy, qx, qy) <= q.refd) { ++cnt; first = min(first, p); }
                        }
                        lo = cnt == 0 ? P : (cnt == P - first ? first : FB);
                    }
                    if (left) loL = lo; else loR = lo;
                }
                loRs[dl] = (lo_t) loR; loLs[dl] = (lo_t) loL;
            }
            // destinations with equal suffix starts have equal layer vectors: one of them (the leader) is swept
            const bool fb = loR == FB || loL == FB;
            const uint32_t key = !have ? 0xffffffffu : (fb ? (0x80000000u | dl) : (loR | (loL << 16)));
            const uint32_t peers = __match_any_sync(kFull, key);
            const bool empty = loR == PR && loL == PL;
            const bool leads = have && !empty && (uint32_t) (__ffs((int) peers) - 1) == lane;
            if (leads) nxt[dl] = (lo_t) atomicExch(&head[loR == FB ? PR : loR], dl);
            if (!SPILL) needL |= __reduce_or_sync(kFull, (leads && loL < PL) ? (1u << loL) : 0u);
        }
        __syncwarp();
        // The suffix table holds one row per suffix start that a leader needs (a handful per cell); should they outnumber the
        // rows, the sweep is repeated for the next kTabRows of them.
        const uint32_t npass = SPILL ? 1u : max(1u, ((uint32_t) __popc(needL) + kTabRows - 1u) / kTabRows);
        // ---- the sweep: lanes = layers
#pragma unroll 1
        for (uint32_t tp = 0; tp < ntiles * npass; ++tp)
        {
            const uint32_t t = tp / npass, pass = tp - t * npass;
            const uint32_t k0 = klo + t * kBT, k = k0 + lane;
            const bool kact = k <= khi;
            {   // suffix minima of the LEFT group
                const SrcVal* vl = val + gcap;
                double v = INFINITY;
#pragma unroll 2
                for (int r = (int) PL - 1; r >= 0; --r)
                {
                    v = amin(v, src_value(vl[r], vpool, k, kact, C.amax));
                    if (SPILL) tab[(size_t) r * 32 + lane] = v;
                    else if ((needL >> r) & 1u)
                    {
                        const uint32_t row = (uint32_t) __popc(needL & ((1u << r) - 1u)) - pass * kTabRows;
                        if (row < (uint32_t) kTabRows) tab[row * 32 + lane] = v;
                    }
                }
            }
            // backward sweep over the RIGHT group; the destinations whose suffix starts at r are emitted when r is reached
            double vR = INFINITY;
            double pend = PR ? src_value(val[PR - 1], vpool, k, kact, C.amax) : INFINITY;
#pragma unroll 1
            for (int r = (int) PR; r >= 0; --r)
            {
                if (r < (int) PR)
                {
                    const double cur = pend;
                    if (r > 0) pend = src_value(val[r - 1], vpool, k, kact, C.amax);
                    vR = amin(vR, cur);
                }
#pragma unroll 1
                for (uint32_t dl = head[r]; dl != FB; dl = nxt[dl])
                {
                    const uint32_t dd = dc0 + dl;
                    const uint32_t loR = loRs[dl], loL = loLs[dl];
                    double v = vR;
                    if (loL < PL)
                    {
                        uint32_t row = loL;
                        if (!SPILL)
                        {
                            row = (uint32_t) __popc(needL & ((1u << loL) - 1u)) - pass * kTabRows;
                            if (row >= (uint32_t) kTabRows) continue;             // this leader's row belongs to another pass
                        }
                        v = amin(v, tab[(size_t) row * 32 + lane]);
                    }
                    else if (pass != 0u) continue;                              // no table row needed: emitted in the first pass
                    if (loL == FB) v = amin(v, explicit_min(C, true, dd, k, kact, a, b, geo + gcap, val + gcap, PL));
                    if (loR == FB) v = amin(v, explicit_min(C, false, dd, k, kact, a, b, geo, val, PR));
                    // the destination's layer vector: first .. last finite layer, holes stay +inf.  A cell deeper than one tile
                    // is still swept once: the vector is given room up to the cell's last layer when its first finite layer
                    // shows up, later tiles append to it (the unused tail is never read).
                    DirEnt de; de.off = 0; de.kmin = 0; de.len = 0; de.nfin = 0;
                    if (t > 0) de = dir_unpack(mydir[dd]);
                    const uint32_t fm = __ballot_sync(kFull, v < INFINITY);
                    bool started = de.len > 0;
                    if (fm)
                    {
                        const uint32_t tfirst = k0 + (uint32_t) __ffs((int) fm) - 1u, tlast = k0 + 31u - (uint32_t) __clz((int) fm);
                        if (!started)
                        {
                            const uint32_t need = ntiles == 1 ? tlast - tfirst + 1 : khi - tfirst + 1;
                            uint32_t base = 0;
                            if (lane == 0) base = atomicAdd(C.bump, need);
                            base = __shfl_sync(kFull, base, 0);
                            if ((unsigned long long) base + need <= C.cap) { started = true; de.off = base; de.kmin = tfirst; }
                            else if (lane == 0) *C.overflow = 1u;
                        }
                        if (started) { de.len = tlast - de.kmin + 1; de.nfin += __popc(fm); }
                    }
                    if (started && kact && k >= de.kmin && (ntiles > 1 || k < de.kmin + de.len))
                        C.sl.vpool[de.off + (k - de.kmin)] = v;
                    if (lane == 0 && (t == 0 || fm)) mydir[dd] = dir_pack(de);
                }
            }
            __syncwarp();
        }
        // ---- every destination gets its directory word (the leader's record is shared by its peers), and the consumer
        //      cell of every non-empty one learns that the edge is live
#pragma unroll 1
        for (uint32_t d0 = 0; d0 < nd; d0 += 32)
        {
            const uint32_t dl = d0 + lane, dd = dc0 + dl;
            const bool have = dl < nd;
            const uint32_t loR = have ? (uint32_t) loRs[dl] : PR, loL = have ? (uint32_t) loLs[dl] : PL;
            const bool fb = loR == FB || loL == FB;
            const uint32_t key = !have ? 0xffffffffu : (fb ? (0x80000000u | dl) : (loR | (loL << 16)));
            const uint32_t peers = __match_any_sync(kFull, key);
            const int leader = __ffs((int) peers) - 1;
            const bool empty = loR == PR && loL == PL;
            unsigned long long wv = 0ull;
            if (have && !empty && leader == (int) lane) wv = mydir[dd];
            wv = __shfl_sync(kFull, wv, leader);
            if (have && (empty || leader != (int) lane)) mydir[dd] = empty ? 0ull : wv;
            const uint32_t nf = (uint32_t) (wv >> 52), ln = (uint32_t) (wv >> 40) & 0xfffu;
            if (have && !empty && ln)
            {
                const bool dR = dd < nDR;
                const uint32_t i = dR ? b + 1 + dd : a + 1 + (dd - nDR);
                // consumer: cell (a,i) sees this edge R[b]->R[i] in its RIGHT part, cell (i,b) sees L[a]->L[i] in its LEFT part
                const uint32_t pos = dR ? C.inv[0][i * (i - 1) / 2 + b] : C.inv[1][i * (i - 1) / 2 + a];
                uint32_t* cw = C.sl.live + (size_t) (dR ? a * nr + i : i * nr + b) * C.WB + (dR ? 0u : C.WR) + (pos >> 5);
                atomicOr(cw, 1u << (pos & 31u));
                my_states += nf;
            }
        }
        __syncwarp();
    }
    if (lane == 0) C.sl.alive[cell] = 1;
    return true;
}

struct CellOut { double bv; unsigned long long cells; uint32_t bk; unsigned int states; };
__device__ __noinline__ CellOut process_cell_spilled(const CellCtx& C, uint32_t a, uint32_t b, uint32_t lane, CellOut io)
{   // by value in, by value out: the caller's running totals stay in registers
    process_cell<true>(C, a, b, lane, io.bk, io.bv, io.cells, io.states);
    return io;
}

struct CandResult { uint32_t kbest; uint32_t states; double area; };     // best terminal layer (0 = none) of one candidate

// Shared memory of k_ad_cells: [per-warp workspaces | lune coordinates + indices | done counters].
// The staging arrays of build_lune (projection keys, unsorted indices) are only live while the lists are built, so they
// alias the workspaces.
__host__ __device__ inline size_t cells_ws_bytes(uint32_t nw) { return kWarpWsBytes * nw; }
__host__ __device__ inline size_t cells_stage_bytes(uint32_t capL, uint32_t capR) { return (size_t) (capL + capR) * (sizeof(double) + sizeof(int32_t)); }
__host__ __device__ inline size_t cells_front_bytes(uint32_t capL, uint32_t capR, uint32_t nw)
{
    const size_t a = cells_ws_bytes(nw), b = (cells_stage_bytes(capL, capR) + 15) & ~(size_t) 15;
    return a > b ? a : b;
}
__host__ __device__ inline size_t cells_smem_bytes(uint32_t capL, uint32_t capR, uint32_t nw)
{
    return cells_front_bytes(capL, capR, nw) + (size_t) (capL + capR) * (2 * sizeof(double) + sizeof(int32_t) + sizeof(uint32_t)) + 16;
}

__device__ __forceinline__ LuneSmem carve_cells_smem(unsigned char* smem_raw, uint32_t capL, uint32_t capR, uint32_t nw, uint32_t** done)
{
    LuneSmem S;
    S.tproj = reinterpret_cast<double*>(smem_raw);                                  // aliases the workspaces
    S.tidx = reinterpret_cast<int32_t*>(smem_raw + sizeof(double) * (capL + capR));
    double* d = reinterpret_cast<double*>(smem_raw + cells_front_bytes(capL, capR, nw));
    S.Lx = d; d += capL; S.Ly = d; d += capL; S.Rx = d; d += capR; S.Ry = d; d += capR;
    int32_t* q = reinterpret_cast<int32_t*>(d);
    S.Lidx = q; q += capL; S.Ridx = q; q += capR;
    *done = reinterpret_cast<uint32_t*>(q);
    S.lflag = nullptr;
    return S;
}

// Runs the whole layered DP of ONE candidate (already described by pr) inside the calling CTA's slice.
// Cells are handed out in diagonal-major order; a warp starts cell (a,b) as soon as row a is finished up to b and
// column b up to a (its only inputs), so no CTA-wide barrier separates the anti-diagonals.
__device__ void run_candidate(const double* __restrict__ x, const double* __restrict__ y, uint32_t n, const Tables& T, const PairRec& pr,
                              CellCtx& C, uint32_t* s_cnt, uint32_t* s_wk, double* s_wv, uint64_t kmax, uint32_t nlayers_ref,
                              unsigned long long& my_cells, unsigned int& my_states, CandResult& out, unsigned int* s_cstates)
{
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const unsigned int states0 = my_states;
    if (threadIdx.x == 0) *s_cstates = 0u;
    {
        LuneSmem S = C.S;
        build_lune(x, y, n, pr, S, s_cnt);
    }
    const uint32_t nl = pr.nl, nr = pr.nr;
    const uint64_t inner = (uint64_t) nl + nr - 4;
    if (threadIdx.x == 0)
    {
        C.nl = nl; C.nr = nr;
        C.lim = (uint32_t) (kmax - 2 < inner ? kmax - 2 : inner);           // :137
        C.nlayers = nlayers_ref;                                          // :483
        C.Sx = C.S.Lx[0]; C.Sy = C.S.Ly[0]; C.Ex = C.S.Lx[nl - 1]; C.Ey = C.S.Ly[nl - 1];
        C.Sp = C.S.Lidx[0]; C.Ep = C.S.Lidx[nl - 1];
        C.seg2 = dist2(C.Sx, C.Sy, C.Ex, C.Ey);
        C.W = nl + nr - 2; C.WR = (nr + 31) / 32; C.WB = live_words(nl, nr);
    }
    __syncthreads();
    // sorted orders and "stay" suffix starts of both chains (the key scratch is the warp's own workspace)
    {
        const uint32_t nmax = nl > nr ? nl : nr;
        double* keys = nmax * sizeof(double) <= kWarpWsBytes ? reinterpret_cast<double*>(C.ws_base + kWarpWsBytes * warp)
                                                          : reinterpret_cast<double*>(C.spill_base + C.spill_bytes * warp);
        build_chain_tables(C, T, false, keys, lane, warp, nw);
        build_chain_tables(C, T, true, keys, lane, warp, nw);
    }
    // reset the slice: no cell has produced anything, no edge is live
    for (uint32_t t = threadIdx.x; t < nl * nr; t += blockDim.x) C.sl.alive[t] = t == 0 ? 1 : 0;
    {
        const size_t nwords = (size_t) nl * nr * C.WB;                 // slices are 256-byte aligned and sized for the widest lune
        uint4* z = reinterpret_cast<uint4*>(C.sl.live);
        for (size_t t = threadIdx.x; t < (nwords + 3) / 4; t += blockDim.x) z[t] = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncthreads();
    // cell (0,0) "produces" the layer-0 seeds (:149-160): RIGHT(0; 0->i) and LEFT(0; 0->i) all share one record (layer 0, area 0)
    const uint32_t D0 = (nr - 1) + (nl - 1);
    for (uint32_t t = threadIdx.x; t < D0; t += blockDim.x)
    {
        DirEnt de; de.off = 0; de.kmin = 0; de.len = 1; de.nfin = 1;
        C.sl.dir[t] = dir_pack(de);
        const bool dR = t < nr - 1;
        const uint32_t i = dR ? t + 1 : t - (nr - 1) + 1;
        const uint32_t pos = C.inv[dR ? 0 : 1][i * (i - 1) / 2];
        atomicOr(C.sl.live + (size_t) (dR ? i : i * nr) * C.WB + (dR ? 0u : C.WR) + (pos >> 5), 1u << (pos & 31u));
    }
    for (uint32_t t = threadIdx.x; t < nl; t += blockDim.x) C.rowdone[t] = t == 0 ? 1u : 0u;
    for (uint32_t t = threadIdx.x; t < nr; t += blockDim.x) C.coldone[t] = t == 0 ? 1u : 0u;
    if (threadIdx.x == 0) { C.sl.vpool[0] = 0.0; *C.bump = 1u; *C.overflow = 0u; *C.next = 0u; }
    my_states += (threadIdx.x == 0) ? D0 : 0u;
    uint32_t bk = 0; double bv = INFINITY;
    __syncthreads();
    const uint32_t total = nl * nr - 1;                 // every cell but (0,0)
    uint32_t d = 1, dstart = 0;                         // current diagonal and the order index of its first cell
    for (;;)
    {
        uint32_t t = 0;
        if (lane == 0) t = atomicAdd(C.next, 1u);
        t = __shfl_sync(kFull, t, 0);
        if (t >= total) break;
        uint32_t alo, cnt;
        for (;;)
        {
            alo = d > nr - 1 ? d - (nr - 1) : 0u;
            cnt = min(d, nl - 1) - alo + 1;
            if (t < dstart + cnt) break;
            dstart += cnt; ++d;
        }
        const uint32_t a = alo + (t - dstart), b = d - a;
        if (lane == 0) while (C.rowdone[a] < b || C.coldone[b] < a) __nanosleep(32);
        __syncwarp();
        __threadfence_block();
        if (!process_cell<false>(C, a, b, lane, bk, bv, my_cells, my_states))
        {
            CellOut io; io.bk = bk; io.bv = bv; io.cells = my_cells; io.states = my_states;
            io = process_cell_spilled(C, a, b, lane, io);
            bk = io.bk; bv = io.bv; my_cells = io.cells; my_states = io.states;
        }
        __threadfence_block();
        __syncwarp();
        if (lane == 0) { C.rowdone[a] = b + 1; C.coldone[b] = a + 1; }
    }
    // best terminal state of the candidate: deepest layer, then smallest area (:169-173)
    for (int o = 16; o; o >>= 1)
    {
        const uint32_t ok = __shfl_xor_sync(kFull, bk, o); const double ov = __shfl_xor_sync(kFull, bv, o);
        if (ok > bk || (ok == bk && ov < bv)) { bk = ok; bv = ov; }
    }
    unsigned int dst = my_states - states0;          // states this candidate created (the reference's hash entries, :519-524)
    for (int o = 16; o; o >>= 1) dst += __shfl_xor_sync(kFull, dst, o);
    if (lane == 0) { s_wk[warp] = bk; s_wv[warp] = bv; atomicAdd(s_cstates, dst); }
    __syncthreads();
    out.states = *s_cstates;
    uint32_t k = 0; double v = INFINITY;
    for (int w = 0; w < (int) (blockDim.x >> 5); ++w) if (s_wk[w] > k || (s_wk[w] == k && s_wv[w] < v)) { k = s_wk[w]; v = s_wv[w]; }
    out.kbest = k; out.area = v;
    __syncthreads();
}

__device__ __forceinline__ void setup_ctx(CellCtx& C, unsigned char* smem_raw, uint32_t capL, uint32_t capR, unsigned char* arena,
                                          const CellLayout& Lo, uint32_t slice, unsigned char* aux, const AuxLayout& Ax, double amax,
                                          unsigned int* next, unsigned int* bump, unsigned int* overflow)
{
    uint32_t* done = nullptr;
    const uint32_t nw = blockDim.x >> 5;
    C.S = carve_cells_smem(smem_raw, capL, capR, nw, &done);
    C.ws_base = smem_raw;
    unsigned char* ax = aux + (size_t) blockIdx.x * Ax.bytes;
    for (int ch = 0; ch < 2; ++ch)
    {
        C.sig[ch] = reinterpret_cast<uint16_t*>(ax + Ax.off_sig[ch]); C.inv[ch] = reinterpret_cast<uint16_t*>(ax + Ax.off_inv[ch]);
        C.stay[ch] = reinterpret_cast<uint16_t*>(ax + Ax.off_stay[ch]); C.edge[ch] = reinterpret_cast<EdgeRec*>(ax + Ax.off_edge[ch]);
    }
    C.spill_base = ax + Ax.off_spill; C.spill_bytes = Ax.spill_bytes;
    C.spill_grp = Ax.grp_cap; C.spill_dst = Ax.dst_cap;
    C.rowdone = done; C.coldone = done + capL;
    C.sl = carve_cells(arena, Lo, slice);
    C.cap = Lo.cap; C.amax = amax; C.next = next; C.bump = bump; C.overflow = overflow;
}

// One candidate pair per CTA, pulled from a device queue.  `order` (optional) lists the queue in processing order.
// Two instantiations of the same body: the 4-warp CTA (4 per SM) and the wide CTA (8 or 16 warps); 128 registers either way.
template <int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB)
k_ad_cells(const double* __restrict__ x, const double* __restrict__ y, uint32_t n, Tables T,
           const PairRec* __restrict__ pairs, const uint32_t* __restrict__ order, uint32_t npairs, AdCounters* __restrict__ counters,
           unsigned char* arena, CellLayout Lo, unsigned char* aux, AuxLayout Ax, uint32_t capL, uint32_t capR,
           uint64_t kmax, uint32_t nlayers_ref, double amax, CtaBest* __restrict__ cta_best,
           uint32_t* __restrict__ ovf_list, unsigned int* __restrict__ ovf_count, uint32_t round)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t s_cnt[2];
    __shared__ unsigned int s_pair, s_bump, s_overflow, s_next, s_cstates;
    __shared__ uint32_t s_wk[kMaxCellWarps];
    __shared__ double s_wv[kMaxCellWarps];
    __shared__ CtaBest s_best;
    __shared__ CellCtx C;
    uint32_t cur = 0;                      // each CTA owns slices 2*blockIdx and 2*blockIdx+1: the one being written and the
                                           // one that holds the CTA's best candidate so far (kept for the witness walk)
    if (threadIdx.x == 0) setup_ctx(C, smem_raw, capL, capR, arena, Lo, 2 * blockIdx.x + cur, aux, Ax, amax, &s_next, &s_bump, &s_overflow);
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31;
    unsigned long long my_cells = 0; unsigned int my_states = 0;
    if (threadIdx.x == 0) { s_best.found = 0; s_best.count = 0; s_best.area = 0.0; s_best.key = ~0ull; s_best.s = s_best.t = 0;
                            s_best.nl = s_best.nr = 0; s_best.slice = 0; s_best.round = round; s_best.states = 0; s_best.pad = 0; }
    for (;;)
    {
        __syncthreads();
        if (threadIdx.x == 0) s_pair = atomicAdd(&counters->queue, 1u);
        __syncthreads();
        const uint32_t qi = s_pair;
        if (qi >= npairs) break;
        const uint32_t pi = order ? order[qi] : qi;
        const PairRec pr = pairs[pi];
        CandResult res;
        unsigned long long cells0 = my_cells; unsigned int states0 = my_states;
        run_candidate(x, y, n, T, pr, C, s_cnt, s_wk, s_wv, kmax, nlayers_ref, my_cells, my_states, res, &s_cstates);
        if (s_overflow)
        {   // the record pool of this slice was too small: the host re-runs the candidate in a larger slice
            my_cells = cells0; my_states = states0;
            if (threadIdx.x == 0) ovf_list[atomicAdd(ovf_count, 1u)] = pi;
            continue;
        }
        // candidate result (AntipodalOptimized.cpp:328-331, 512-517): count desc, area asc, first pair in (i<j) order
        bool better = false;
        if (res.kbest > 0)
        {
            const uint32_t cnt = res.kbest + 2;
            const unsigned long long key = (unsigned long long) pr.s * n + pr.t;
            better = !s_best.found || cnt > s_best.count ||
                     (cnt == s_best.count && (res.area < s_best.area || (res.area == s_best.area && key < s_best.key)));
            __syncthreads();                   // every thread has read s_best
            if (better && threadIdx.x == 0)
            {
                s_best.found = 1; s_best.count = cnt; s_best.area = res.area; s_best.key = key; s_best.s = pr.s; s_best.t = pr.t;
                s_best.nl = pr.nl; s_best.nr = pr.nr; s_best.slice = 2 * blockIdx.x + cur; s_best.round = round; s_best.states = res.states;
            }
        }
        if (better)
        {   // keep this candidate's records: continue in the other slice
            cur ^= 1u;
            __syncthreads();
            if (threadIdx.x == 0) C.sl = carve_cells(arena, Lo, 2 * blockIdx.x + cur);
        }
    }
    for (int o = 16; o; o >>= 1) { my_cells += __shfl_xor_sync(kFull, my_cells, o); my_states += __shfl_xor_sync(kFull, my_states, o); }
    if (lane == 0) { atomicAdd(&counters->cells, my_cells); atomicAdd(&counters->states, (unsigned long long) my_states); }
    __syncthreads();
    if (threadIdx.x == 0) cta_best[blockIdx.x] = s_best;
}

__device__ __forceinline__ bool cta_best_better(const CtaBest& c, const CtaBest& b)
{
    if (!c.found) return false;
    return !b.found || c.count > b.count || (c.count == b.count && (c.area < b.area || (c.area == b.area && c.key < b.key)));
}

// Best candidate over the per-CTA records (count desc, area asc, pair key asc: AntipodalOptimized.cpp:512-517).  One CTA: every
// thread folds a strided share, then a tree over shared memory (the order is total, so any association gives the same winner).
__global__ void __launch_bounds__(256)
k_ad_reduce(const CtaBest* __restrict__ in, uint32_t count, CtaBest* __restrict__ out, int merge_existing)
{
    __shared__ CtaBest s_b[256];
    CtaBest b;
    b.found = 0; b.count = 0; b.area = 0.0; b.key = ~0ull; b.s = b.t = 0; b.nl = b.nr = 0; b.slice = 0; b.round = 0; b.states = 0; b.pad = 0;
    if (merge_existing && threadIdx.x == 0) b = *out;
    for (uint32_t i = threadIdx.x; i < count; i += blockDim.x)
    {
        const CtaBest c = in[i];
        if (cta_best_better(c, b)) b = c;
    }
    s_b[threadIdx.x] = b;
    __syncthreads();
    for (uint32_t o = blockDim.x >> 1; o; o >>= 1)
    {
        if (threadIdx.x < o && cta_best_better(s_b[threadIdx.x + o], s_b[threadIdx.x])) s_b[threadIdx.x] = s_b[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) *out = s_b[0];
}

// --------------------------------------------------------------------------------------------- witness recovery

struct AdHull { uint32_t len; uint32_t ok; uint32_t nl, nr; };

// Area pattern of the reference state Entry{o, s, ss} on the given side at layer k, or +inf if it was never reached.
__device__ unsigned long long state_value(const CellSlice& sl, uint32_t nl, uint32_t nr, bool right, uint32_t o, uint32_t s, uint32_t ss, uint32_t k)
{
    // RIGHT(o; s->ss) is destination ss-s-1 of cell (o, s); LEFT(o; s->ss) is destination (nr-1-o) + ss-s-1 of cell (s, o)
    const uint32_t cell = right ? o * nr + s : s * nr + o;
    if (!sl.alive[cell]) return kInfBits;
    const DirEnt de = dir_unpack(sl.dir[(size_t) cell * (nl + nr - 2) + (right ? ss - s - 1 : (nr - 1 - o) + (ss - s - 1))]);
    if (de.len == 0 || k < de.kmin || k >= de.kmin + de.len) return kInfBits;
    return area_bits(sl.vpool[de.off + (k - de.kmin)]);      // holes hold +inf = kInfBits
}

// Re-derives the chain of states behind the best terminal state of layer kopt and emits the hull in the
// reference's order (locReconstructHull, AntipodalOptimized.cpp:335-436): [t, left chain backwards.., s, right chain backwards..].
// Runs after k_ad_cells processed the winning candidate alone in slice 0.
__global__ void __launch_bounds__(kCellThreads)
k_ad_backtrack(const double* __restrict__ x, const double* __restrict__ y, uint32_t n, Tables T, PairRec pr,
               unsigned char* arena, CellLayout Lo, uint32_t slice, uint32_t kopt,
               int32_t* __restrict__ left, int32_t* __restrict__ right, int32_t* __restrict__ hull, AdHull* __restrict__ out)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t s_cnt[2];
    LuneSmem S = carve_smem(smem_raw, pr.nl, pr.nr);
    build_lune(x, y, n, pr, S, s_cnt);
    const CellSlice sl = carve_cells(arena, Lo, slice);
    const uint32_t nl = pr.nl, nr = pr.nr;
    // best terminal state of layer kopt (:348-355): smallest area, first in (side, o, s) order; all threads scan
    __shared__ unsigned long long s_v[kCellThreads];
    __shared__ uint32_t s_i[kCellThreads];
    {
        unsigned long long bvv = kInfBits; uint32_t bi = 0xffffffffu;
        const uint32_t nR = nl * (nr - 1), nLs = nr * (nl - 1);
        for (uint32_t t = threadIdx.x; t < nR + nLs; t += blockDim.x)
        {
            const bool rt = t < nR;
            const uint32_t u = rt ? t : t - nR;
            const uint32_t o = rt ? u / (nr - 1) : u / (nl - 1), s = rt ? u % (nr - 1) : u % (nl - 1);
            const unsigned long long v = state_value(sl, nl, nr, rt, o, s, rt ? nr - 1 : nl - 1, kopt);
            if (v < bvv || (v == bvv && t < bi)) { bvv = v; bi = t; }
        }
        s_v[threadIdx.x] = bvv; s_i[threadIdx.x] = bi;
    }
    __syncthreads();
    if (threadIdx.x >= 32) return;            // one warp walks: lanes test 32 predecessor candidates at a time
    const uint32_t lane = threadIdx.x;
    const double Sx = S.Lx[0], Sy = S.Ly[0], Ex = S.Lx[nl - 1], Ey = S.Ly[nl - 1];
    const int Sp = S.Lidx[0], Ep = S.Lidx[nl - 1];
    const double seg2 = dist2(Sx, Sy, Ex, Ey);
    bool e_right = true; uint32_t eo = 0, es = 0, ess = 0; unsigned long long ev = kInfBits;
    {
        uint32_t bi = 0xffffffffu;
        for (uint32_t t = 0; t < blockDim.x; ++t) if (s_v[t] < ev || (s_v[t] == ev && s_i[t] < bi)) { ev = s_v[t]; bi = s_i[t]; }
        if (ev != kInfBits)
        {
            const uint32_t nR = nl * (nr - 1);
            e_right = bi < nR;
            const uint32_t u = e_right ? bi : bi - nR;
            eo = e_right ? u / (nr - 1) : u / (nl - 1); es = e_right ? u % (nr - 1) : u % (nl - 1); ess = e_right ? nr - 1 : nl - 1;
        }
    }
    uint32_t a = 0, b = 0, ok = ev != kInfBits ? 1u : 0u;
    if (e_right) left[a++] = Ep; else right[b++] = Sp;     // :359-366
    uint32_t k = kopt;
    while (ok && k > 0)
    {
        if (a + b >= nl + nr) { ok = 0; break; }     // exactly degenerate input: a step that consumes no point must not run away
        const bool rt = e_right;
        const double* ownx = rt ? S.Rx : S.Lx; const double* owny = rt ? S.Ry : S.Ly;
        const double* othx = rt ? S.Lx : S.Rx; const double* othy = rt ? S.Ly : S.Ry;
        const int32_t* owni = rt ? S.Ridx : S.Lidx; const int32_t* othi = rt ? S.Lidx : S.Ridx;
        const double ax = rt ? Sx : Ex, ay = rt ? Sy : Ey; const int ap = rt ? Sp : Ep;        // own fan apex
        const double bx = rt ? Ex : Sx, by = rt ? Ey : Sy; const int bp = rt ? Ep : Sp;        // other fan apex
        int32_t* own_out = rt ? right : left; uint32_t& own_cnt = rt ? b : a;
        int32_t* oth_out = rt ? left : right; uint32_t& oth_cnt = rt ? a : b;
        const uint32_t o = eo, s = es, ss = ess;
        bool found = false, via_switch = false; uint32_t prev = 0, cused = 0; unsigned long long pv = kInfBits;
        // (a) the state was produced by staying on its own chain: source (o, p, s) on the same side; the first p (then z) that
        //     reproduces the stored value, as a sequential scan would find it
        if (dist2(othx[o], othy[o], ownx[s], owny[s]) <= seg2)
            for (uint32_t p0 = 0; p0 < s && !found; p0 += 32)
            {
                const uint32_t p = p0 + lane;
                bool hit = false; uint32_t hc = 0; unsigned long long hv = kInfBits;
                if (p < s && orientation(ownx[ss], owny[ss], ownx[s], owny[s], ownx[p], owny[p]) >= 0 &&
                    orientation(ownx[ss], owny[ss], ownx[s], owny[s], ax, ay) >= 0)
                {
                    const double tri = fabs(signed_area(ax, ay, ownx[p], owny[p], ownx[s], owny[s]));
                    for (int z = 0; z < 2 && !hit; ++z)
                    {
                        const uint32_t c = 1u + (uint32_t) tri_count_ad(T, z == 1, owni[p], ownx[p], owny[p], owni[s], ownx[s], owny[s], ap, ax, ay);
                        if (c > k) continue;
                        const uint32_t ks = k - c;
                        if ((z == 1) != (ks == 0)) continue;
                        const unsigned long long sv = state_value(sl, nl, nr, rt, o, p, s, ks);
                        if (sv == kInfBits) continue;
                        if (area_bits(__dadd_rn(bits_area(sv), tri)) != ev) continue;
                        hit = true; hc = c; hv = sv;
                    }
                }
                const uint32_t hm = __ballot_sync(0xffffffffu, hit);
                if (hm)
                {
                    const int w = __ffs((int) hm) - 1;
                    found = true; via_switch = false; prev = p0 + (uint32_t) w;
                    cused = __shfl_sync(0xffffffffu, hc, w); pv = __shfl_sync(0xffffffffu, hv, w);
                }
            }
        // (b) it was produced by a switch from the other chain: source (o' = s, s' = p, ss' = o) on the other side
        if (!found && dist2(ownx[s], owny[s], othx[o], othy[o]) <= seg2)
            for (uint32_t p0 = 0; p0 < o && !found; p0 += 32)
            {
                const uint32_t p = p0 + lane;
                bool hit = false; uint32_t hc = 0; unsigned long long hv = kInfBits;
                if (p < o)
                {
                    const double rx = othx[p], ry = othy[p], px = othx[o], py = othy[o];
                    const double refd = pseudo_dist(rx, ry, px, py, ownx[s], owny[s]);
                    const double d = pseudo_dist(rx, ry, px, py, ownx[ss], owny[ss]);
                    if (d <= refd && (s == 0 || (orientation(ownx[s], owny[s], ax, ay, ownx[ss], owny[ss]) >= 0 &&
                                                 orientation(ownx[ss], owny[ss], ownx[s], owny[s], bx, by) >= 0)))
                    {
                        const double tri = fabs(signed_area(bx, by, rx, ry, px, py));
                        for (int z = 0; z < 2 && !hit; ++z)
                        {
                            const uint32_t c = 1u + (uint32_t) tri_count_ad(T, z == 1, othi[p], rx, ry, othi[o], px, py, bp, bx, by);
                            if (c > k) continue;
                            const uint32_t ks = k - c;
                            if ((z == 1) != (ks == 0)) continue;
                            const unsigned long long sv = state_value(sl, nl, nr, !rt, s, p, o, ks);
                            if (sv == kInfBits) continue;
                            if (area_bits(__dadd_rn(bits_area(sv), tri)) != ev) continue;
                            hit = true; hc = c; hv = sv;
                        }
                    }
                }
                const uint32_t hm = __ballot_sync(0xffffffffu, hit);
                if (hm)
                {
                    const int w = __ffs((int) hm) - 1;
                    found = true; via_switch = true; prev = p0 + (uint32_t) w;
                    cused = __shfl_sync(0xffffffffu, hc, w); pv = __shfl_sync(0xffffffffu, hv, w);
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
    __syncwarp();
    if (lane != 0) return;
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
struct AdSmall { PairStats stats; AdCounters counters; CtaBest best; AdHull hull; uint32_t lune_cnt[2]; unsigned int ovf_count[2]; };

constexpr uint64_t kMinPoolRecords = 1u << 16;        // smallest record pool a slice may have
constexpr uint64_t kDefaultPoolRecords = 4u << 20;    // first-try pool per slice (overflowing candidates are re-run larger)

// The three CTA widths of k_ad_cells (warps per candidate): 4 (5 CTAs per SM), 10 (2 per SM), 20 (1 per SM).
#define ADK_CELLS_DISPATCH(nw, EXPR) \
    do { if ((nw) == (uint32_t) kCellWarps) { auto kern = k_ad_cells<kCellThreads, kCellCtas>; EXPR; } \
         else if ((nw) == (uint32_t) kMidCellWarps) { auto kern = k_ad_cells<kMidCellWarps * 32, 2>; EXPR; } \
         else { auto kern = k_ad_cells<kMaxCellWarps * 32, 1>; EXPR; } } while (0)

// Launches the cell kernel over `count` queue entries (order == nullptr: pairs[0..count)) with `slices` CTAs.
int ad_launch_cells(Ctx* c, const PairRec* d_pairs, const uint32_t* d_order, uint32_t count, uint32_t slices, const CellLayout& Lo,
                    const AuxLayout& Ax, uint32_t capL, uint32_t capR, uint64_t kmax, uint32_t nlayers_ref, double amax, AdSmall* sm,
                    CtaBest* d_cta_best, uint32_t* d_ovf_list, unsigned int* d_ovf_count, uint32_t round, uint32_t nw)
{
    const size_t smem = cells_smem_bytes(capL, capR, nw);
    if (smem > 200 * 1024) { c->err = "area+diameter path: lune lists do not fit in shared memory"; return ADK_ERR_NOMEM; }
    ADK_CELLS_DISPATCH(nw, ADK_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem)));
    ADK_CUDA(c, cudaMemsetAsync(&sm->counters.queue, 0, sizeof(unsigned int), c->stream));
    ADK_CUDA(c, cudaMemsetAsync(d_ovf_count, 0, sizeof(unsigned int), c->stream));
    Tables T{c->below.as<int32_t>(), c->coll.as<int32_t>(), c->rb.as<int32_t>(), c->n};
    ADK_CELLS_DISPATCH(nw, (kern<<<slices, nw * 32, smem, c->stream>>>(c->x.as<double>(), c->y.as<double>(), c->n, T, d_pairs, d_order, count,
                                                          &sm->counters, c->arena.as<unsigned char>(), Lo, c->ad_aux.as<unsigned char>(), Ax, capL, capR, kmax, nlayers_ref,
                                                          amax, d_cta_best, d_ovf_list, d_ovf_count, round)));
    ++c->launches;
    ADK_CUDA(c, cudaGetLastError());
    return ADK_OK;
}

// Resident CTAs of k_ad_cells per SM for the given lune capacity.
int ad_cells_occupancy(Ctx* c, uint32_t capL, uint32_t capR, uint32_t nw, int* per_sm)
{
    const size_t smem = cells_smem_bytes(capL, capR, nw);
    if (smem > 200 * 1024)
    {
        if (nw > (uint32_t) kCellWarps) { *per_sm = 0; return ADK_OK; }        // this CTA width is not available for such lunes
        c->err = "area+diameter path: lune lists do not fit in shared memory"; return ADK_ERR_NOMEM;
    }
    ADK_CELLS_DISPATCH(nw, ADK_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem)));
    int occ = 0;
    ADK_CELLS_DISPATCH(nw, ADK_CUDA(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, (int) nw * 32, smem)));
    *per_sm = occ > 0 ? occ : (nw > (uint32_t) kCellWarps ? 0 : 1);
    return ADK_OK;
}

// Runs every queued candidate; candidates whose record pool overflowed are re-run in fewer, larger slices.
int ad_run_queue(Ctx* c, const PairRec* d_pairs, const uint32_t* d_order0, uint32_t nqueued, const PairStats& ps, uint64_t kmax, uint32_t nlayers_ref, double amax,
                 AdSmall* sm, CellLayout* used_layout, uint32_t* last_round)
{
    if (ps.max_dirs >= 0xffffffffull) { c->err = "area+diameter path: candidate too large for 32-bit directory offsets"; return ADK_ERR_NOMEM; }
    const size_t fixed = cell_layout(ps.max_cells, ps.max_dirs, ps.max_nl, ps.max_nr, 0).off_v + 8192;
    const AuxLayout Ax = aux_layout(ps.max_nl, ps.max_nr);
    // Warps per candidate.  Plenty of candidates per CTA: 4 warps each, 4 CTAs per SM (best throughput).  Few, heavy ones
    // (a shard of a multi-GPU solve, the re-run of a winner): the longest candidate is the critical path, so wider CTAs
    // take fewer candidates at a time and finish each one sooner.  Both sides are estimated in the same cost proxy.
    int occ = 1;
    uint32_t nw = (uint32_t) kCellWarps;
    {
        const uint32_t widths[3] = {(uint32_t) kCellWarps, (uint32_t) kMidCellWarps, (uint32_t) kMaxCellWarps};
        const double eff[3] = {1.0, 0.88, 0.67};                       // measured throughput per warp of 4 / 10 / 20 warps per candidate (386 / 438 / 573 ms per sweep)
        double best_t = 0.0;
        for (int v = 0; v < 3; ++v)
        {
            const uint32_t w = widths[v];
            int o = 0;
            ADK_TRY(ad_cells_occupancy(c, ps.max_nl, ps.max_nr, w, &o));
            if (o <= 0) break;
            const double speed = (double) w / kCellWarps * eff[v];       // of one CTA, relative to a 4-warp CTA
            const double ctas = (double) std::min<uint64_t>(nqueued, (uint64_t) c->sm_count * (uint64_t) o);
            const double t = std::max((double) ps.cost_sum / (ctas * speed), (double) ps.cost_max / speed);
            if (c->force_cell_warps ? c->force_cell_warps == w : (v == 0 || t < 0.9 * best_t)) { best_t = v == 0 ? t : std::min(t, best_t); nw = w; occ = o; }
            if (c->debug) fprintf(stderr, "[adk] queue %u  w=%u occ=%d  load=%.3g critical=%.3g -> t=%.3g  (chosen w=%u)\n", nqueued, w, o,
                                             (double) ps.cost_sum / (ctas * speed), (double) ps.cost_max / speed, t, nw);
        }
    }
    // per-CTA tables + spill workspaces (not kept between candidates), sized for the widest lune of this queue
    ADK_TRY(ensure(c, c->ad_aux, Ax.bytes * (size_t) std::min<uint64_t>(nqueued, (uint64_t) c->sm_count * (uint64_t) occ)));
    ADK_TRY(ensure(c, c->ad_ovf, sizeof(uint32_t) * 2 * (size_t) nqueued + 64));
    uint32_t* d_ovf[2] = {c->ad_ovf.as<uint32_t>(), c->ad_ovf.as<uint32_t>() + nqueued};
    const uint32_t* d_order = d_order0;
    uint32_t count = nqueued;
    uint64_t pool = kDefaultPoolRecords, prev_cap = 0;
    int merge = 0;
    for (int round = 0; count > 0; ++round)
    {
        const uint64_t budget = workspace_budget(c, c->arena);
        const uint64_t min_bytes = fixed + 10 * kMinPoolRecords, want_bytes = fixed + 10 * pool;
        if (budget < 2 * min_bytes) { c->err = "area+diameter path: one candidate's workspace exceeds the memory budget"; return ADK_ERR_NOMEM; }
        // every CTA owns two slices (the one it writes and the one holding its best candidate so far)
        uint64_t slices = std::min<uint64_t>(count, (uint64_t) c->sm_count * (uint64_t) occ);
        if (round > 0) slices = std::min<uint64_t>(slices, std::max<uint64_t>(1, budget / (2 * want_bytes)));   // fewer, larger slices
        slices = std::max<uint64_t>(1, std::min<uint64_t>(slices, budget / (2 * min_bytes)));
        size_t slice_bytes = (size_t) std::min<uint64_t>(budget / (2 * slices), want_bytes) & ~(size_t) 255;
        const CellLayout Lo = cell_layout(ps.max_cells, ps.max_dirs, ps.max_nl, ps.max_nr, slice_bytes);
        if (round > 0 && Lo.cap <= prev_cap)
        {   // the overflowing candidate already had everything the budget allows
            c->err = "area+diameter path: one candidate's record pool exceeds the memory budget"; return ADK_ERR_NOMEM;
        }
        ADK_TRY(ensure(c, c->arena, slice_bytes * 2 * (size_t) slices));
        ADK_TRY(ensure(c, c->ad_cta, sizeof(CtaBest) * slices));
        ADK_TRY(ad_launch_cells(c, d_pairs, d_order, count, (uint32_t) slices, Lo, Ax, ps.max_nl, ps.max_nr, kmax, nlayers_ref, amax, sm,
                                c->ad_cta.as<CtaBest>(), d_ovf[round & 1], &sm->ovf_count[round & 1], (uint32_t) round, nw));
        k_ad_reduce<<<1, 256, 0, c->stream>>>(c->ad_cta.as<CtaBest>(), (uint32_t) slices, &sm->best, merge);
        ++c->launches;
        ADK_CUDA(c, cudaGetLastError());
        merge = 1;
        unsigned int novf = 0;
        ADK_CUDA(c, cudaMemcpyAsync(&novf, &sm->ovf_count[round & 1], sizeof(novf), cudaMemcpyDeviceToHost, c->stream));
        ADK_CUDA(c, cudaStreamSynchronize(c->stream));
        if (used_layout) *used_layout = Lo;
        if (last_round) *last_round = (uint32_t) round;
        if (novf == 0) break;
        d_order = d_ovf[round & 1]; count = novf;
        prev_cap = Lo.cap; pool = (uint64_t) Lo.cap * 8;
    }
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
    uint32_t* dense = nullptr; uint32_t* rowhist = nullptr;
    if (nshards > 1)
    {
        if (n > 0xffffu) { c->err = "area+diameter shards: n too large for the dense candidate table"; return ADK_ERR_ARG; }
        ADK_TRY(ensure(c, c->scratch1, sizeof(uint32_t) * (size_t) n * n));
        ADK_TRY(ensure(c, c->scratch2, sizeof(uint32_t) * (size_t) n * kCostBuckets));
        dense = c->scratch1.as<uint32_t>(); rowhist = c->scratch2.as<uint32_t>();
    }
    k_ad_pairs<<<n, 256, 0, c->stream>>>(c->x.as<double>(), c->y.as<double>(), n, dmax2, kmax, c->scratch0.as<PairRec>(), &sm->stats,
                                         dense, rowhist);
    ++c->launches;
    ADK_CUDA(c, cudaGetLastError());
    if (nshards > 1)
    {
        k_ad_deal<<<n, 128, 0, c->stream>>>(n, kmax, shard, nshards, dense, rowhist, c->scratch0.as<PairRec>(), &sm->stats);
        ++c->launches;
        ADK_CUDA(c, cudaGetLastError());
    }
    PairStats ps;
    ADK_CUDA(c, cudaMemcpyAsync(&ps, &sm->stats, sizeof(ps), cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaStreamSynchronize(c->stream));
    float dp_ms = 0.f;
    CtaBest hb; std::memset(&hb, 0, sizeof(hb));
    AdCounters hc; std::memset(&hc, 0, sizeof(hc));
    CellLayout keep_layout; std::memset(&keep_layout, 0, sizeof(keep_layout));
    uint32_t keep_round = 0;
    c->ad_keep_valid = false;
    if (ps.nqueued > 0)
    {
        ADK_CUDA(c, cudaEventRecord(c->ev2, c->stream));
        ADK_TRY(ensure(c, c->ad_order, sizeof(uint32_t) * (size_t) ps.nqueued));
        k_ad_order<<<std::min<uint32_t>((ps.nqueued + 255) / 256, 64), 256, 0, c->stream>>>(c->scratch0.as<PairRec>(), ps.nqueued, &sm->stats,
                                                                                          c->ad_order.as<uint32_t>());
        ++c->launches;
        ADK_CUDA(c, cudaGetLastError());
        ADK_TRY(ad_run_queue(c, c->scratch0.as<PairRec>(), c->ad_order.as<uint32_t>(), ps.nqueued, ps, kmax, nlayers_ref, amax, sm,
                             &keep_layout, &keep_round));
        ADK_CUDA(c, cudaEventRecord(c->ev3, c->stream));
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
        if (hb.round == keep_round)
        {   // the winner's records are still in its slice: ad_finish can walk them without re-running the candidate
            static_assert(sizeof(CellLayout) <= sizeof(c->ad_keep_layout), "ad_keep_layout too small");
            std::memcpy(c->ad_keep_layout, &keep_layout, sizeof(CellLayout));
            c->ad_keep_valid = true; c->ad_keep_slice = hb.slice; c->ad_keep_s = hb.s; c->ad_keep_t = hb.t;
            c->ad_keep_nl = hb.nl; c->ad_keep_nr = hb.nr; c->ad_keep_count = hb.count; c->ad_keep_states = hb.states; c->ad_keep_kmax = kmax; c->ad_keep_amax = amax;
        }
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
    ADK_TRY(ensure(c, c->scratch6, sizeof(int32_t) * 3 * ((size_t) n + 4) + sizeof(PairRec) + 64));
    int32_t* d_left = c->scratch6.as<int32_t>();
    int32_t* d_right = d_left + (n + 4);
    int32_t* d_hull = d_right + (n + 4);
    PairRec* d_pair = reinterpret_cast<PairRec*>(d_hull + (n + 4));
    PairRec pr;
    CellLayout Lo;
    uint32_t slice = 0;
    const uint32_t kopt = best->count - 2;
    if (c->ad_keep_valid && c->ad_keep_s == best->w0 && c->ad_keep_t == best->w1 && c->ad_keep_count == best->count &&
        c->ad_keep_kmax == kmax && c->ad_keep_amax == amax)
    {   // the partial solve on this context left the winner's records in place
        pr = PairRec{best->w0, best->w1, c->ad_keep_nl, c->ad_keep_nr};
        std::memcpy(&Lo, c->ad_keep_layout, sizeof(CellLayout));
        slice = c->ad_keep_slice;
        out->entries_required = (int64_t) c->ad_keep_states;      // carried to the caller: the re-run the reference does (:529-535)
    }
    else
    {   // the winning candidate alone (it was found on another GPU, or evicted): size it, re-run it (:529-535)
        const size_t smem_ll = lune_smem_bytes(n + 2, n + 2, 1);
        if (smem_ll > 200 * 1024) { c->err = "ad_finish: point set too large for the lune kernel"; return ADK_ERR_NOMEM; }
        ADK_CUDA(c, cudaFuncSetAttribute(k_ad_lune_lists, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem_ll));
        k_ad_lune_lists<<<1, kAdThreads, smem_ll, c->stream>>>(c->x.as<double>(), c->y.as<double>(), n, best->w0, best->w1, d_left, d_right, sm->lune_cnt);
        ++c->launches;
        ADK_CUDA(c, cudaGetLastError());
        uint32_t cnt[2];
        ADK_CUDA(c, cudaMemcpyAsync(cnt, sm->lune_cnt, sizeof(cnt), cudaMemcpyDeviceToHost, c->stream));
        ADK_CUDA(c, cudaStreamSynchronize(c->stream));
        pr = PairRec{best->w0, best->w1, cnt[0], cnt[1]};
        const uint64_t inner = (uint64_t) pr.nl + pr.nr - 4;
        const uint32_t lim = (uint32_t) std::min<uint64_t>(kmax - 2, inner);
        if (kopt > lim) { c->err = "ad_finish: witness count exceeds the candidate's capacity"; return ADK_ERR_ARG; }
        PairStats ps; std::memset(&ps, 0, sizeof(ps));
        ps.nqueued = 1; ps.max_nl = pr.nl; ps.max_nr = pr.nr; ps.max_lim = lim;
        ps.max_cells = (unsigned long long) pr.nl * pr.nr; ps.max_dirs = ps.max_cells * (pr.nl + pr.nr - 2);
        ps.cost_sum = ps.cost_max = ps.max_cells * (pr.nl + pr.nr);
        ADK_CUDA(c, cudaMemcpyAsync(d_pair, &pr, sizeof(pr), cudaMemcpyHostToDevice, c->stream));
        ADK_CUDA(c, cudaMemsetAsync(&sm->counters, 0, sizeof(AdCounters), c->stream));
        c->ad_keep_valid = false;
        ADK_TRY(ad_run_queue(c, d_pair, nullptr, 1, ps, kmax, nlayers_ref, amax, sm, &Lo, nullptr));
        slice = 0;                               // the only candidate of the only CTA: its first slice
        AdCounters rc;
        ADK_CUDA(c, cudaMemcpyAsync(&rc, &sm->counters, sizeof(rc), cudaMemcpyDeviceToHost, c->stream));
        ADK_CUDA(c, cudaStreamSynchronize(c->stream));
        out->entries_required = (int64_t) rc.states;
    }
    Tables T{c->below.as<int32_t>(), c->coll.as<int32_t>(), c->rb.as<int32_t>(), n};
    const size_t smem = lune_smem_bytes(pr.nl, pr.nr, 1);
    ADK_CUDA(c, cudaFuncSetAttribute(k_ad_backtrack, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    k_ad_backtrack<<<1, kCellThreads, smem, c->stream>>>(c->x.as<double>(), c->y.as<double>(), n, T, pr, c->arena.as<unsigned char>(), Lo, slice, kopt,
                                                         d_left, d_right, d_hull, &sm->hull);
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
