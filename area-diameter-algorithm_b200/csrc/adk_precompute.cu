// adk_precompute.cu — geometry precompute shared by both DP paths (sm_100a).
//
//  k_angles      : angle / squared distance of every point seen from every centre          (CustomMath.cpp:56-65, 15-20)
//  k_order_rank  : per-centre neighbour orders by rank-counting with the reference's comparator
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

// One CTA per centre.  rank(p) = #neighbours that the comparator puts before p; O(n^2) per centre but
// exact for the reference's (angle, then "near-tie and collinear -> nearer first") order.  That comparator is not transitive
// on eps-collinear triples (A before B before C before A is possible), so two points can count the same number of
// predecessors: whenever that happens the final positions are taken from the total order (count, point index) instead, which
// is a permutation whatever the comparator does (and equals the counts when they are distinct).
template <int MODE>
__global__ void k_order_rank(const double* __restrict__ x, const double* __restrict__ y, uint32_t n,
                             const double* __restrict__ ang, const double* __restrict__ d2c,
                             int32_t* __restrict__ order, uint16_t* __restrict__ rank)
{
    extern __shared__ double sm[];
    double* sk = sm;            // sort key (angle, or angle mod pi)
    double* sd = sm + n;        // squared distance from the centre
    double* sx = sm + 2 * n;
    double* sy = sm + 3 * n;
    uint32_t* scnt = reinterpret_cast<uint32_t*>(sm + 4 * n);     // predecessor count of every point
    uint32_t* shist = scnt + n;                                   // points per count
    __shared__ int s_dup;
    const uint32_t c = blockIdx.x;
    if (threadIdx.x == 0) s_dup = 0;
    for (uint32_t p = threadIdx.x; p < n; p += blockDim.x)
    {
        double a = ang[(size_t) c * n + p];
        if (MODE == 1 && a >= kPi) a = __dsub_rn(a, kPi);
        sk[p] = a; sd[p] = d2c[(size_t) c * n + p]; sx[p] = x[p]; sy[p] = y[p];
        scnt[p] = 0xffffffffu; shist[p] = 0;
    }
    __syncthreads();
    const double cx = sx[c], cy = sy[c];
    for (uint32_t p = threadIdx.x; p < n; p += blockDim.x)
    {
        if (p == c) continue;
        const double kp = sk[p], dp = sd[p], px = sx[p], py = sy[p];
        uint32_t cnt = 0;
        for (uint32_t q = 0; q < n; ++q)
        {
            if (q == c || q == p) continue;
            const double kq = sk[q];
            bool before;
            // ArePointsClockwise(centre, q, p): Utils.cpp:295-309
            if (fabs(__dsub_rn(kq, kp)) < kAngleTol && near_zero(signed_area(cx, cy, sx[q], sy[q], px, py)))
            {
                const double dq = sd[q];
                before = dq != dp ? dq < dp : q < p;
            }
            else before = kq != kp ? kq > kp : q < p;
            cnt += before ? 1u : 0u;
        }
        scnt[p] = cnt;
        if (atomicAdd(&shist[cnt], 1u) != 0u) s_dup = 1;
    }
    __syncthreads();
    const bool dup = s_dup != 0;
    for (uint32_t p = threadIdx.x; p < n; p += blockDim.x)
    {
        if (p == c) continue;
        uint32_t pos = scnt[p];
        if (dup)
        {
            const uint32_t cp = pos;
            pos = 0;
            for (uint32_t q = 0; q < n; ++q)
            {
                if (q == c) continue;
                const uint32_t cq = scnt[q];
                pos += (cq < cp || (cq == cp && q < p)) ? 1u : 0u;
            }
        }
        order[(size_t) c * (n - 1) + pos] = (int32_t) p;
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
    const size_t smem = sizeof(double) * 4 * n + sizeof(uint32_t) * 2 * n;
    if (need_cw && !c->have_cw)
    {
        ADK_TRY(ensure(c, c->cw, sizeof(int32_t) * (size_t) n * (n - 1)));
        ADK_CUDA(c, cudaFuncSetAttribute(k_order_rank<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        k_order_rank<0><<<n, 256, smem, c->stream>>>(c->x.as<double>(), c->y.as<double>(), n, c->ang.as<double>(),
                                                     c->d2c.as<double>(), c->cw.as<int32_t>(), nullptr);
        ++c->launches;
        ADK_CUDA(c, cudaGetLastError());
        c->have_cw = true;
    }
    if (need_modpi && !c->have_modpi)
    {
        ADK_TRY(ensure(c, c->modpi, sizeof(int32_t) * (size_t) n * (n - 1)));
        ADK_TRY(ensure(c, c->rank, sizeof(uint16_t) * (size_t) n * n));
        ADK_CUDA(c, cudaFuncSetAttribute(k_order_rank<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        k_order_rank<1><<<n, 256, smem, c->stream>>>(c->x.as<double>(), c->y.as<double>(), n, c->ang.as<double>(),
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

}  // namespace adk
