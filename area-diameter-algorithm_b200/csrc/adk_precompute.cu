// adk_precompute.cu — geometry precompute shared by both DP paths (sm_100a).
//
//  k_angles      : angle / squared distance of every point seen from every centre          (CustomMath.cpp:56-65, 15-20)
//  k_order_sort  : per-centre neighbour orders: segmented radix sort + exact-comparator fix-up of the near-tie windows
//                    mode 0: clockwise list                                                   (Utils.cpp:295-316)
//                    mode 1: direction-mod-pi list = the one cyclic sequence every slope list of
//                            Eppstein.cpp:62-114 is a rotation of
//  k_count_tables: the three integer tables of Utils.h:28-33 from their definitions (Utils.cpp:145-223),
//                  one warp per point pair, lanes over the third point
//  k_yrank       : anchor order (Eppstein.cpp:158-159) and |Above(i)|
//
// All inputs are SoA double x[n], y[n]; reads are coalesced along the third point.
#include "adk_internal.h"
#include "adk_device.cuh"

#include <cstring>
#include <string>

namespace adk
{

__global__ void k_angles(const double* __restrict__ x, const double* __restrict__ y, uint32_t n,
                         double* __restrict__ ang, double* __restrict__ d2c)
{
    const uint32_t c = blockIdx.x;
    const double cx = x[c], cy = y[c];
    for (uint32_t p = threadIdx.x; p < n; p += blockDim.x)
    {
        double a = 0.0, d = 0.0;
        if (p != c) { a = angle_of(cx, cy, x[p], y[p]); d = dist2(cx, cy, x[p], y[p]); }
        ang[(size_t) c * n + p] = a;
        d2c[(size_t) c * n + p] = d;
    }
}

// Angular order around every centre (north_star stage 2): a SEGMENTED RADIX SORT, one segment = one centre = one CTA.
//   1. key = order-preserving 64-bit image of the angle (angles are >= 0, so the bit pattern is monotone), inverted because the
//      reference sorts angle-DESCENDING (Utils.cpp:295-316); payload = the point index.  LSD radix sort, 8 passes of 8 bits, in shared
//      memory: per pass a histogram per warp, one scan over (digit, warp), and a stable scatter in which the lanes of a warp rank
//      themselves inside their digit with __match_any_sync.  Stable, so equal angles keep index order (the comparator's last rule).
//   2. fix-up with the EXACT comparator: the reference orders two points whose angles differ by < 1e-6 AND that are eps-collinear with
//      the centre by distance instead (nearer first).  In the sorted list the points within 1e-6 of p form a window around p, so
//      the number of points the comparator puts before p is (window start) + (exact tests inside the window) — the same count the
//      round-1 kernel formed by testing p against all n points.
//   3. that comparator is not transitive on eps-collinear triples (A before B before C before A is possible), so two points can
//      count the same number of predecessors: whenever that happens the final positions are taken from the total order
//      (count, point index) instead, which is a permutation whatever the comparator does (and equals the counts when they are distinct).
// O(n) work per point instead of O(n^2): n = 4096 sorts in well under 2 ms; the round-1 rank counting was ~70 G comparisons there.
constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;

template <int MODE>
__global__ void __launch_bounds__(kSortThreads)
k_order_sort(const double* __restrict__ x, const double* __restrict__ y, uint32_t n,
             const double* __restrict__ ang, const double* __restrict__ d2c,
             int32_t* __restrict__ order, uint16_t* __restrict__ rank)
{
    extern __shared__ __align__(16) unsigned char sm_raw[];
    const uint32_t m = n - 1;                                      // keys per segment
    // [keyA m x u64][keyB m x u64][idxA m x u16][idxB m x u16][sx n][sy n][sd n] ; counts / histogram alias the spare sort buffer
    unsigned long long* keyA = reinterpret_cast<unsigned long long*>(sm_raw);
    unsigned long long* keyB = keyA + m;
    double* sx = reinterpret_cast<double*>(keyB + m);
    double* sy = sx + n;
    double* sd = sy + n;
    uint16_t* idxA = reinterpret_cast<uint16_t*>(sd + n);
    uint16_t* idxB = idxA + m;
    __shared__ uint32_t s_hist[kSortWarps][256];
    __shared__ uint32_t s_digit_base[256];
    __shared__ uint32_t s_warp_tot[kSortWarps];
    __shared__ int s_dup;
    const uint32_t c = blockIdx.x;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_dup = 0;
    for (uint32_t p = threadIdx.x; p < n; p += blockDim.x)
    {
        sx[p] = x[p]; sy[p] = y[p]; sd[p] = d2c[(size_t) c * n + p];
        if (p == c) continue;
        double a = ang[(size_t) c * n + p];
        if (MODE == 1 && a >= kPi) a = __dsub_rn(a, kPi);
        const uint32_t t = p < c ? p : p - 1;
        keyA[t] = ~(unsigned long long) __double_as_longlong(a);      // ascending in the inverted bits = descending in the angle
        idxA[t] = (uint16_t) p;
    }
    __syncthreads();
    // ---- 1. LSD radix sort of (key, index), 8 bits per pass; warp w owns the contiguous slice [w * per, (w + 1) * per) of the input
    const uint32_t per = ((m + kSortWarps - 1) / kSortWarps + 31) & ~31u;
    const uint32_t w0 = min(m, warp * per), w1 = min(m, w0 + per);
    unsigned long long* kin = keyA; unsigned long long* kout = keyB;
    uint16_t* iin = idxA; uint16_t* iout = idxB;
    for (uint32_t shift = 0; shift < 64; shift += 8)
    {
        for (uint32_t t = threadIdx.x; t < kSortWarps * 256; t += blockDim.x) (&s_hist[0][0])[t] = 0;
        __syncthreads();
        for (uint32_t t = w0 + lane; t < w1; t += 32) atomicAdd(&s_hist[warp][(uint32_t) (kin[t] >> shift) & 255u], 1u);
        __syncthreads();
        {   // exclusive scan over (digit major, warp minor): thread d owns digit d
            const uint32_t d = threadIdx.x;                       // kSortThreads == 256 digits
            uint32_t col[kSortWarps], tot = 0;
            for (int w = 0; w < kSortWarps; ++w) { col[w] = s_hist[w][d]; tot += col[w]; }
            uint32_t incl = tot;
            for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o); if ((int) lane >= o) incl += v; }
            if (lane == 31) s_warp_tot[warp] = incl;
            __syncthreads();
            uint32_t base = incl - tot;
            for (uint32_t w = 0; w < warp; ++w) base += s_warp_tot[w];
            for (int w = 0; w < kSortWarps; ++w) { s_hist[w][d] = base; base += col[w]; }
        }
        __syncthreads();
        for (uint32_t t0 = w0; t0 < w1; t0 += 32)
        {
            const uint32_t t = t0 + lane;
            const bool have = t < w1;
            const unsigned long long k = have ? kin[t] : 0ull;
            const uint32_t d = have ? (uint32_t) (k >> shift) & 255u : 256u + lane;      // idle lanes: distinct pseudo digits
            const uint32_t peers = __match_any_sync(0xffffffffu, d);
            const uint32_t r = __popc(peers & ((1u << lane) - 1u));
            uint32_t pos = 0;
            if (have) pos = s_hist[warp][d] + r;
            __syncwarp();
            if (have && r == 0) s_hist[warp][d] += __popc(peers);                        // the group's first lane moves the cursor on
            __syncwarp();
            if (have) { kout[pos] = k; iout[pos] = iin[t]; }
        }
        __syncthreads();
        unsigned long long* tk = kin; kin = kout; kout = tk;
        uint16_t* ti = iin; iin = iout; iout = ti;
    }
    // after 8 passes the sorted data is back in (keyA, idxA); keyB is free: counts and their histogram live there
    uint32_t* scnt = reinterpret_cast<uint32_t*>(keyB);          // predecessor count of the point at sorted position t
    uint32_t* shist = scnt + m;                                   // points per count   (2 x 4 m bytes <= 8 m bytes)
    for (uint32_t t = threadIdx.x; t < m; t += blockDim.x) shist[t] = 0;
    __syncthreads();
    // ---- 2. exact comparator inside the near-tie window (ArePointsClockwise(centre, q, p), Utils.cpp:295-309)
    const double cx = sx[c], cy = sy[c];
    for (uint32_t t = threadIdx.x; t < m; t += blockDim.x)
    {
        const uint32_t p = iin[t];
        const double kp = __longlong_as_double((long long) ~kin[t]);
        const double dp = sd[p], px = sx[p], py = sy[p];
        uint32_t lo = t;                                          // window = sorted positions whose key is within the tolerance of kp
        while (lo > 0 && fabs(__dsub_rn(__longlong_as_double((long long) ~kin[lo - 1]), kp)) < kAngleTol) --lo;
        uint32_t cnt = lo;                                        // everything before the window has a clearly larger angle
        for (uint32_t u = lo; u < m; ++u)
        {
            if (u == t) continue;
            const double kq = __longlong_as_double((long long) ~kin[u]);
            if (!(fabs(__dsub_rn(kq, kp)) < kAngleTol)) { if (u > t) break; else continue; }
            const uint32_t q = iin[u];
            bool before;
            if (near_zero(signed_area(cx, cy, sx[q], sy[q], px, py)))
            {
                const double dq = sd[q];
                before = dq != dp ? dq < dp : q < p;
            }
            else before = kq != kp ? kq > kp : q < p;
            cnt += before ? 1u : 0u;
        }
        scnt[t] = cnt;
        if (atomicAdd(&shist[cnt], 1u) != 0u) s_dup = 1;
    }
    __syncthreads();
    // ---- 3. placement
    const bool dup = s_dup != 0;
    for (uint32_t t = threadIdx.x; t < m; t += blockDim.x)
    {
        const uint32_t p = iin[t];
        uint32_t pos = scnt[t];
        if (dup)
        {
            const uint32_t cp = pos;
            pos = 0;
            for (uint32_t u = 0; u < m; ++u)
            {
                const uint32_t cq = scnt[u], q = iin[u];
                pos += (cq < cp || (cq == cp && q < p)) ? 1u : 0u;
            }
        }
        order[(size_t) c * m + pos] = (int32_t) p;
        if (rank) rank[(size_t) c * n + p] = (uint16_t) pos;
    }
}

// One warp per unordered pair (a, b), a = blockIdx.x, b strided over the CTA's warps.
__global__ void k_count_tables(const double* __restrict__ x, const double* __restrict__ y, uint32_t n,
                               int32_t* __restrict__ below, int32_t* __restrict__ coll, int32_t* __restrict__ rb)
{
    const uint32_t a = blockIdx.x;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const double ax = x[a], ay = y[a];
    if (warp == 0)
    {   // rightBelow: same-x points that precede a in (x, y) order (Utils.cpp:92-95)
        int cnt = 0;
        for (uint32_t p = lane; p < n; p += 32) cnt += (p != a && x[p] == ax && y[p] < ay) ? 1 : 0;
        for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        if (lane == 0) rb[a] = cnt;
        if (lane == 0) { below[(size_t) a * n + a] = 0; coll[(size_t) a * n + a] = 0; }
    }
    for (uint32_t b = a + 1 + warp; b < n; b += nwarps)
    {
        const double bx = x[b], by = y[b];
        // r = the later point in (x, y) order (the sweep's pi), l = the other (Utils.cpp:86-102)
        const bool a_first = xy_less(ax, ay, bx, by);
        const double rx = a_first ? bx : ax, ry = a_first ? by : ay;
        const double lx = a_first ? ax : bx, ly = a_first ? ay : by;
        const double d2 = dist2(rx, ry, lx, ly);
        const double ex = __dsub_rn(lx, rx), ey = __dsub_rn(ly, ry);
        int cc = 0, bb = 0;
        for (uint32_t p = lane; p < n; p += 32)
        {
            if (p == a || p == b) continue;
            const double px = x[p], py = y[p];
            const double s = signed_area(rx, ry, lx, ly, px, py);
            if (near_zero(s))
            {   // collinear and strictly inside the segment (Utils.cpp:148-162)
                const double dot = __dadd_rn(__dmul_rn(ex, __dsub_rn(px, rx)), __dmul_rn(ey, __dsub_rn(py, ry)));
                cc += (dot > 0.0 && dot < d2) ? 1 : 0;
            }
            else if (s > 0.0 && px > lx && px < rx) ++bb;   // strictly below the segment (Utils.cpp:176-187)
        }
        for (int o = 16; o; o >>= 1)
        {
            cc += __shfl_xor_sync(0xffffffffu, cc, o);
            bb += __shfl_xor_sync(0xffffffffu, bb, o);
        }
        if (lane == 0)
        {
            if (ax == bx) bb = 0;   // never written by the sweep (Utils.cpp:109,118)
            below[(size_t) a * n + b] = bb; below[(size_t) b * n + a] = bb;
            coll[(size_t) a * n + b] = cc; coll[(size_t) b * n + a] = cc;
        }
    }
}

// yorder[r] = the point with the r-th smallest y (ties by index); above[i] = #{p != i : y_p >= y_i}
__global__ void k_yrank(const double* __restrict__ y, uint32_t n, int32_t* __restrict__ yorder, int32_t* __restrict__ above)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double yi = y[i];
    uint32_t r = 0, lower = 0;
    for (uint32_t p = 0; p < n; ++p)
    {
        const double yp = y[p];
        lower += yp < yi ? 1u : 0u;
        r += (yp < yi || (yp == yi && p < i)) ? 1u : 0u;
    }
    yorder[r] = (int32_t) i;
    above[i] = (int32_t) (n - 1 - lower);
}

__global__ void k_pit_query(const double* __restrict__ x, const double* __restrict__ y, Tables T,
                            const uint32_t* __restrict__ tri, uint32_t m, int32_t* __restrict__ out)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= m) return;
    const int a = (int) tri[3 * t], b = (int) tri[3 * t + 1], c = (int) tri[3 * t + 2];
    out[t] = points_in_triangle(T, a, x[a], y[a], b, x[b], y[b], c, x[c], y[c]);
}

// ------------------------------------------------------------------------------------------------ host side

// new_solve: the call starts a solve (a whole solve or a partial), as opposed to the finish that follows a partial.
int upload_points(Ctx* c, const double* x, const double* y, uint32_t n, bool device_ptrs, bool new_solve)
{
    if (!x || !y || n == 0 || n > ADK_MAX_POINTS) { c->err = "bad point arrays / n out of range"; return ADK_ERR_ARG; }
    // The same point set as the resident one (a diameter sweep, a partial followed by its finish) keeps the count tables,
    // the orders and the retained winning candidate.  "Same" is decided on the coordinates themselves (a host copy of the
    // resident set), never on a hash.  With reuse switched off (adk_set_reuse) every new solve recomputes everything, as the
    // reference does (Eppstein.cpp:143-155, AntipodalOptimized.cpp:469-481); a finish still finds its partial's products.
    bool same = false;
    if (c->points_valid && c->n == n && (c->reuse || !new_solve))
    {
        if (device_ptrs && !new_solve && x == c->dev_x && y == c->dev_y) same = true;     // the buffers the partial just read
        else
        {
            const double* hx = x; const double* hy = y;
            if (device_ptrs)
            {
                ADK_TRY(ensure_pinned(c, sizeof(double) * 2 * (size_t) n));
                double* h = static_cast<double*>(c->h_pinned);
                ADK_CUDA(c, cudaMemcpyAsync(h, x, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream));
                ADK_CUDA(c, cudaMemcpyAsync(h + n, y, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream));
                ADK_CUDA(c, cudaStreamSynchronize(c->stream));
                hx = h; hy = h + n;
            }
            same = c->host_xy.size() == 2 * (size_t) n && std::memcmp(c->host_xy.data(), hx, sizeof(double) * n) == 0 &&
                   std::memcmp(c->host_xy.data() + n, hy, sizeof(double) * n) == 0;
        }
    }
    ADK_TRY(ensure(c, c->x, sizeof(double) * n));
    ADK_TRY(ensure(c, c->y, sizeof(double) * n));
    const cudaMemcpyKind kind = device_ptrs ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    ADK_CUDA(c, cudaMemcpyAsync(c->x.p, x, sizeof(double) * n, kind, c->stream));
    ADK_CUDA(c, cudaMemcpyAsync(c->y.p, y, sizeof(double) * n, kind, c->stream));
    c->dev_x = device_ptrs ? x : nullptr; c->dev_y = device_ptrs ? y : nullptr;
    if (same) return ADK_OK;       // the copy is still made (the caller's buffer is the input of this call); products stay valid
    c->n = n;
    c->have_cw = c->have_modpi = c->have_tables = c->have_yorder = false;
    c->ad_keep_valid = false;
    c->area_keep_valid = false;
    c->points_valid = true;
    if (!device_ptrs)
    {
        c->host_xy.resize(2 * (size_t) n);
        std::memcpy(c->host_xy.data(), x, sizeof(double) * n);
        std::memcpy(c->host_xy.data() + n, y, sizeof(double) * n);
    }
    else if (c->reuse)
    {   // keep the host copy current for later comparisons (one small read-back per NEW device-resident point set)
        ADK_TRY(ensure_pinned(c, sizeof(double) * 2 * (size_t) n));
        double* h = static_cast<double*>(c->h_pinned);
        ADK_CUDA(c, cudaMemcpyAsync(h, c->x.p, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream));
        ADK_CUDA(c, cudaMemcpyAsync(h + n, c->y.p, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream));
        ADK_CUDA(c, cudaStreamSynchronize(c->stream));
        c->host_xy.assign(h, h + 2 * (size_t) n);
    }
    else c->host_xy.clear();
    return ADK_OK;
}

int run_angles_and_orders(Ctx* c, bool need_cw, bool need_modpi)
{
    const uint32_t n = c->n;
    if (n < 2) return ADK_OK;
    if ((need_cw && !c->have_cw) || (need_modpi && !c->have_modpi))
    {
        ADK_TRY(ensure(c, c->ang, sizeof(double) * (size_t) n * n));
        ADK_TRY(ensure(c, c->d2c, sizeof(double) * (size_t) n * n));
        k_angles<<<n, 256, 0, c->stream>>>(c->x.as<double>(), c->y.as<double>(), n, c->ang.as<double>(), c->d2c.as<double>());
        ++c->launches;
        ADK_CUDA(c, cudaGetLastError());
    }
    const size_t smem = (sizeof(unsigned long long) * 2 + sizeof(uint16_t) * 2) * (size_t) (n - 1) + sizeof(double) * 3 * (size_t) n + 64;
    if (need_cw && !c->have_cw)
    {
        ADK_TRY(ensure(c, c->cw, sizeof(int32_t) * (size_t) n * (n - 1)));
        ADK_CUDA(c, cudaFuncSetAttribute(k_order_sort<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        k_order_sort<0><<<n, kSortThreads, smem, c->stream>>>(c->x.as<double>(), c->y.as<double>(), n, c->ang.as<double>(),
                                                     c->d2c.as<double>(), c->cw.as<int32_t>(), nullptr);
        ++c->launches;
        ADK_CUDA(c, cudaGetLastError());
        c->have_cw = true;
    }
    if (need_modpi && !c->have_modpi)
    {
        ADK_TRY(ensure(c, c->modpi, sizeof(int32_t) * (size_t) n * (n - 1)));
        ADK_TRY(ensure(c, c->rank, sizeof(uint16_t) * (size_t) n * n));
        ADK_CUDA(c, cudaFuncSetAttribute(k_order_sort<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        k_order_sort<1><<<n, kSortThreads, smem, c->stream>>>(c->x.as<double>(), c->y.as<double>(), n, c->ang.as<double>(),
                                                     c->d2c.as<double>(), c->modpi.as<int32_t>(), c->rank.as<uint16_t>());
        ++c->launches;
        ADK_CUDA(c, cudaGetLastError());
        c->have_modpi = true;
    }
    ADK_CUDA(c, cudaGetLastError());
    return ADK_OK;
}

int run_count_tables(Ctx* c)
{
    const uint32_t n = c->n;
    if (c->have_tables) return ADK_OK;
    ADK_TRY(ensure(c, c->below, sizeof(int32_t) * (size_t) n * n));
    ADK_TRY(ensure(c, c->coll, sizeof(int32_t) * (size_t) n * n));
    ADK_TRY(ensure(c, c->rb, sizeof(int32_t) * n));
    k_count_tables<<<n, 256, 0, c->stream>>>(c->x.as<double>(), c->y.as<double>(), n, c->below.as<int32_t>(),
                                             c->coll.as<int32_t>(), c->rb.as<int32_t>());
    ++c->launches;
    ADK_CUDA(c, cudaGetLastError());
    c->have_tables = true;
    return ADK_OK;
}

int run_yorder(Ctx* c)
{
    const uint32_t n = c->n;
    if (c->have_yorder) return ADK_OK;
    ADK_TRY(ensure(c, c->yorder, sizeof(int32_t) * n));
    ADK_TRY(ensure(c, c->above, sizeof(int32_t) * n));
    k_yrank<<<(n + 127) / 128, 128, 0, c->stream>>>(c->y.as<double>(), n, c->yorder.as<int32_t>(), c->above.as<int32_t>());
    ++c->launches;
    ADK_CUDA(c, cudaGetLastError());
    c->have_yorder = true;
    return ADK_OK;
}

int run_pit_query(Ctx* c, const uint32_t* h_triples, uint32_t m, int32_t* h_out)
{
    if (m == 0) return ADK_OK;
    ADK_TRY(run_count_tables(c));
    ADK_TRY(ensure(c, c->scratch0, sizeof(uint32_t) * 3 * (size_t) m));
    ADK_TRY(ensure(c, c->scratch1, sizeof(int32_t) * (size_t) m));
    ADK_CUDA(c, cudaMemcpyAsync(c->scratch0.p, h_triples, sizeof(uint32_t) * 3 * (size_t) m, cudaMemcpyHostToDevice, c->stream));
    Tables T{c->below.as<int32_t>(), c->coll.as<int32_t>(), c->rb.as<int32_t>(), c->n};
    k_pit_query<<<(m + 127) / 128, 128, 0, c->stream>>>(c->x.as<double>(), c->y.as<double>(), T, c->scratch0.as<uint32_t>(), m,
                                                        c->scratch1.as<int32_t>());
    ++c->launches;
    ADK_CUDA(c, cudaGetLastError());
    ADK_CUDA(c, cudaMemcpyAsync(h_out, c->scratch1.p, sizeof(int32_t) * (size_t) m, cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaStreamSynchronize(c->stream));
    return ADK_OK;
}

// ------------------------------------------------------------------------------------------------ no-allocation variant
//
// adk_cw_order_noalloc (include/adk.h): the clockwise-order stage for callers that must not allocate (CUDA graph capture, tight
// serving loops).  It has no counterpart in the reference.
// The stage needs scratch: k_angles writes the n x n angle and squared-distance tables that k_order_sort reads.  Here that scratch
// is this file-scope array instead of the context's grow-only buffers: 2 x kNoAllocMaxN^2 doubles = 4 MiB, held by every GPU that
// loads the library until the process exits.  EVERY call shares it, so calls that overlap on the GPU (two streams, two contexts,
// two threads) race; it keeps the previous call's contents, which is harmless: k_angles writes every entry of both tables for the
// call's own n (row stride n) before k_order_sort reads any, so nothing has to start at zero.
constexpr uint32_t kNoAllocMaxN = ADK_NOALLOC_MAX_POINTS;
static __device__ __align__(256) double g_noalloc_scratch[2 * (size_t) kNoAllocMaxN * kNoAllocMaxN];

// In host code the array's name is only a host-side stand-in for the symbol, not a device address: the address is asked for
// once per device, with that device current, and cached.
static double* g_noalloc_ptr[64] = {nullptr};

int cw_order_noalloc(Ctx* c, const double* d_x, const double* d_y, uint32_t n, int32_t* d_order)
{
    if (!d_x || !d_y || !d_order) { c->err = "adk_cw_order_noalloc: null pointer"; return ADK_ERR_ARG; }
    if (n < 2) { c->err = "adk_cw_order_noalloc: n < 2"; return ADK_ERR_ARG; }
    if (n > kNoAllocMaxN)
    {
        c->err = "adk_cw_order_noalloc: n = " + std::to_string(n) + " exceeds the static scratch (ADK_NOALLOC_MAX_POINTS = " +
                 std::to_string(kNoAllocMaxN) + "); use adk_cw_order";
        return ADK_ERR_ARG;
    }
    if (c->device < 0 || c->device >= 64) { c->err = "adk_cw_order_noalloc: device ordinal out of range"; return ADK_ERR_ARG; }
    double* scratch = g_noalloc_ptr[c->device];
    if (!scratch)
    {   // set-up, once per device; the caller (ADK_ENTER) has made the context's device current
        ADK_CUDA(c, cudaGetSymbolAddress(reinterpret_cast<void**>(&scratch), g_noalloc_scratch));
        g_noalloc_ptr[c->device] = scratch;
    }
    double* ang = scratch;
    double* d2c = scratch + (size_t) n * n;
    // 44 n + 64 bytes of dynamic shared memory: 22.6 KB at the maximum n, below the 48 KB that need no opt-in
    const size_t smem = (sizeof(unsigned long long) * 2 + sizeof(uint16_t) * 2) * (size_t) (n - 1) + sizeof(double) * 3 * (size_t) n + 64;
    k_angles<<<n, 256, 0, c->stream>>>(d_x, d_y, n, ang, d2c);
    k_order_sort<0><<<n, kSortThreads, smem, c->stream>>>(d_x, d_y, n, ang, d2c, d_order, nullptr);
    c->launches += 2;
    ADK_CUDA(c, cudaGetLastError());
    return ADK_OK;
}

}  // namespace adk
