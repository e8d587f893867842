// adk_tiled.cu — the tiled area-only workflow with the windows generated ON THE DEVICE.
//
// Replaces create_patch_dict + the window loop of run_area_only_tiled (external/Area-Selector,
// src/convex_area_selector/convex_polygon_finder.py:139-219 and run/quality_of_results_comparison.py:77-103): the slide is cut into
// square windows (patch_size, step_size) on the np.arange grid that starts at min - 0.01, the windows are visited by population,
// largest first (ties in x-major grid order: Python's sorted() is stable), each one is solved with the area path, and the walk stops
// at the first window that holds fewer points than the best polygon found so far.
//   k_tile_population : one warp per window, inclusive bounds (Series.between), ballot / popc over coalesced SoA coordinates
//   k_tile_order      : visiting order of the non-empty windows (population descending, grid index ascending), rank by counting
//   k_tile_members    : the members of one window, in point order, as SoA coordinates + indices — the instance the area path reads
// The window instances never visit the host: only the populations, the order and the winning window's member list do.
#include "adk_internal.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>

namespace adk
{
namespace
{
__global__ void k_tile_population(const double* __restrict__ x, const double* __restrict__ y, uint32_t n,
                                  const double* __restrict__ xr, uint32_t nxw, const double* __restrict__ yr, uint32_t nyw,
                                  double patch, uint32_t* __restrict__ pop)
{
    const uint32_t lane = threadIdx.x & 31, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (warp >= nxw * nyw) return;
    const double x0 = xr[warp / nyw], y0 = yr[warp % nyw];           // x-major: itertools.product(x_range, y_range)
    const double x1 = x0 + patch, y1 = y0 + patch;
    uint32_t cnt = 0;
    for (uint32_t p = lane; p < n; p += 32)
    {
        const double px = x[p], py = y[p];
        cnt += (px >= x0 && px <= x1 && py >= y0 && py <= y1) ? 1u : 0u;
    }
    for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if (lane == 0) pop[warp] = cnt;
}

// order[rank] = window, over the non-empty windows; *count = their number.  Windows are few (hundreds to a few thousand).
__global__ void k_tile_order(const uint32_t* __restrict__ pop, uint32_t nwin, uint32_t* __restrict__ order, uint32_t* __restrict__ count)
{
    for (uint32_t w = blockIdx.x * blockDim.x + threadIdx.x; w < nwin; w += gridDim.x * blockDim.x)
    {
        const uint32_t pw = pop[w];
        if (pw == 0) continue;
        uint32_t r = 0;
        for (uint32_t v = 0; v < nwin; ++v) { const uint32_t pv = pop[v]; r += (pv > pw || (pv == pw && v < w)) ? 1u : 0u; }
        order[r] = w;
        atomicAdd(count, 1u);
    }
}

// One CTA: the members of window w in point order.
__global__ void k_tile_members(const double* __restrict__ x, const double* __restrict__ y, uint32_t n,
                               const double* __restrict__ xr, const double* __restrict__ yr, uint32_t nyw, uint32_t w, double patch,
                               double* __restrict__ ox, double* __restrict__ oy, uint32_t* __restrict__ oidx)
{
    __shared__ uint32_t s_base;
    __shared__ uint32_t s_warp[32];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const double x0 = xr[w / nyw], y0 = yr[w % nyw], x1 = x0 + patch, y1 = y0 + patch;
    if (threadIdx.x == 0) s_base = 0;
    __syncthreads();
    for (uint32_t p0 = 0; p0 < n; p0 += blockDim.x)
    {
        const uint32_t p = p0 + threadIdx.x;
        double px = 0.0, py = 0.0;
        bool in = false;
        if (p < n) { px = x[p]; py = y[p]; in = px >= x0 && px <= x1 && py >= y0 && py <= y1; }
        const uint32_t bm = __ballot_sync(0xffffffffu, in);
        if (lane == 0) s_warp[warp] = __popc(bm);
        __syncthreads();
        uint32_t off = s_base;
        for (uint32_t v = 0; v < warp; ++v) off += s_warp[v];
        if (in)
        {
            const uint32_t at = off + __popc(bm & ((1u << lane) - 1u));
            ox[at] = px; oy[at] = py; oidx[at] = p;
        }
        __syncthreads();
        if (threadIdx.x == 0) { uint32_t t = 0; for (uint32_t v = 0; v < nw; ++v) t += s_warp[v]; s_base += t; }
        __syncthreads();
    }
}

// np.arange(start, stop, step): ceil((stop - start) / step) values start + i * step
std::vector<double> arange(double start, double stop, double step)
{
    std::vector<double> v;
    const double len = std::ceil((stop - start) / step);
    const size_t cnt = len > 0 ? (size_t) len : 0;
    v.reserve(cnt);
    for (size_t i = 0; i < cnt; ++i) v.push_back(start + (double) i * step);
    return v;
}
}  // namespace

int area_tiled(Ctx* c, const double* hx, const double* hy, uint32_t n, double patch, double step, uint64_t kmax, double amax,
               adk_result* out, uint32_t* hull, adk_tiled_stats* st)
{
    std::memset(out, 0, sizeof(*out)); out->entries_alloc = -1; out->entries_required = -1;
    if (st) std::memset(st, 0, sizeof(*st));
    if (!hx || !hy || !(patch > 0.0) || !(step > 0.0)) { c->err = "adk_solve_area_tiled: bad argument"; return ADK_ERR_ARG; }
    if (n == 0) return ADK_OK;
    // DataFrame.drop_duplicates(): the first of equal rows is kept; the windows see the remaining rows in their original order
    std::vector<uint32_t> idx(n);
    std::iota(idx.begin(), idx.end(), 0u);
    std::stable_sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return hx[a] != hx[b] ? hx[a] < hx[b] : hy[a] < hy[b]; });
    std::vector<uint32_t> keep;
    for (uint32_t t = 0; t < n; ++t) if (t == 0 || hx[idx[t]] != hx[idx[t - 1]] || hy[idx[t]] != hy[idx[t - 1]]) keep.push_back(idx[t]);
    std::sort(keep.begin(), keep.end());
    const uint32_t m = (uint32_t) keep.size();
    std::vector<double> xs(m), ys(m);
    for (uint32_t t = 0; t < m; ++t) { xs[t] = hx[keep[t]]; ys[t] = hy[keep[t]]; }
    const double xmin = *std::min_element(xs.begin(), xs.end()), xmax = *std::max_element(xs.begin(), xs.end());
    const double ymin = *std::min_element(ys.begin(), ys.end()), ymax = *std::max_element(ys.begin(), ys.end());
    const std::vector<double> xr = (xmax - xmin) < patch ? std::vector<double>{xmin - 0.01} : arange(xmin - 0.01, xmax, step);
    const std::vector<double> yr = (ymax - ymin) < patch ? std::vector<double>{ymin - 0.01} : arange(ymin - 0.01, ymax, step);
    const uint32_t nxw = (uint32_t) xr.size(), nyw = (uint32_t) yr.size(), nwin = nxw * nyw;
    if (nwin == 0) return ADK_OK;

    // device buffers: the slide, the grid, populations / order, one window instance
    ADK_TRY(ensure(c, c->tile_buf, sizeof(double) * (2 * (size_t) m + nxw + nyw) + sizeof(uint32_t) * (2 * (size_t) nwin + 8)
                                   + (sizeof(double) * 2 + sizeof(uint32_t)) * (size_t) m + 256));
    double* d_x = c->tile_buf.as<double>();
    double* d_y = d_x + m;
    double* d_xr = d_y + m;
    double* d_yr = d_xr + nxw;
    double* d_wx = d_yr + nyw;
    double* d_wy = d_wx + m;
    uint32_t* d_pop = reinterpret_cast<uint32_t*>(d_wy + m);
    uint32_t* d_order = d_pop + nwin;
    uint32_t* d_cnt = d_order + nwin;
    uint32_t* d_widx = d_cnt + 8;
    ADK_CUDA(c, cudaMemcpyAsync(d_x, xs.data(), sizeof(double) * m, cudaMemcpyHostToDevice, c->stream));
    ADK_CUDA(c, cudaMemcpyAsync(d_y, ys.data(), sizeof(double) * m, cudaMemcpyHostToDevice, c->stream));
    ADK_CUDA(c, cudaMemcpyAsync(d_xr, xr.data(), sizeof(double) * nxw, cudaMemcpyHostToDevice, c->stream));
    ADK_CUDA(c, cudaMemcpyAsync(d_yr, yr.data(), sizeof(double) * nyw, cudaMemcpyHostToDevice, c->stream));
    ADK_CUDA(c, cudaMemsetAsync(d_cnt, 0, sizeof(uint32_t) * 8, c->stream));
    k_tile_population<<<(nwin * 32 + 255) / 256, 256, 0, c->stream>>>(d_x, d_y, m, d_xr, nxw, d_yr, nyw, patch, d_pop);
    k_tile_order<<<std::min<uint32_t>((nwin + 127) / 128, 256), 128, 0, c->stream>>>(d_pop, nwin, d_order, d_cnt);
    c->launches += 2;
    ADK_CUDA(c, cudaGetLastError());
    std::vector<uint32_t> pop(nwin), order(nwin);
    uint32_t nonempty = 0;
    ADK_CUDA(c, cudaMemcpyAsync(pop.data(), d_pop, sizeof(uint32_t) * nwin, cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaMemcpyAsync(order.data(), d_order, sizeof(uint32_t) * nwin, cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaMemcpyAsync(&nonempty, d_cnt, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    ADK_CUDA(c, cudaStreamSynchronize(c->stream));
    if (st) st->patches = nonempty;

    // the walk (quality_of_results_comparison.py:82-98)
    bool have = false;
    adk_result best; std::memset(&best, 0, sizeof(best));
    std::vector<uint32_t> best_hull, whull(m + 4), widx(m);
    uint64_t solved = 0, cells_n = 0, cells_e = 0;
    double kernel_ms = 0.0, dp_ms = 0.0;
    for (uint32_t r = 0; r < nonempty; ++r)
    {
        const uint32_t w = order[r], pw = pop[w];
        if (have && best.count > pw) break;
        ++solved;
        if (pw < 3 || kmax < 3) continue;                                     // Eppstein.cpp:137-140: std::nullopt
        k_tile_members<<<1, 256, 0, c->stream>>>(d_x, d_y, m, d_xr, d_yr, nyw, w, patch, d_wx, d_wy, d_widx);
        ++c->launches;
        ADK_CUDA(c, cudaGetLastError());
        ADK_TRY(upload_points(c, d_wx, d_wy, pw, true, true));
        adk_partial part; adk_stats ps;
        ADK_TRY(area_partial(c, kmax, amax, 0, 1, &part, &ps));
        cells_n += ps.cells_nominal; cells_e += ps.cells_executed; kernel_ms += ps.kernel_ms; dp_ms += ps.dp_ms;
        if (!part.found || (have && !(best.count < part.count))) continue;   // strict improvement in the count only
        adk_result res; std::memset(&res, 0, sizeof(res));
        ADK_TRY(area_finish(c, kmax, amax, &part, ADK_FLAG_RECONSTRUCT, &res, whull.data()));
        ADK_CUDA(c, cudaMemcpyAsync(widx.data(), d_widx, sizeof(uint32_t) * pw, cudaMemcpyDeviceToHost, c->stream));
        ADK_CUDA(c, cudaStreamSynchronize(c->stream));
        best = res; have = true;
        best_hull.resize(res.hull_len);
        for (uint32_t t = 0; t < res.hull_len; ++t) best_hull[t] = keep[widx[whull[t]]];      // back to indices of the caller's points
        if (res.diam_valid) { best.diam[0] = keep[widx[res.diam[0]]]; best.diam[1] = keep[widx[res.diam[1]]]; }
    }
    if (st) st->solved = solved;
    if (have)
    {
        *out = best;
        if (hull) for (uint32_t t = 0; t < best.hull_len && t < n; ++t) hull[t] = best_hull[t];
    }
    out->entries_alloc = -1; out->entries_required = -1;
    out->cells_nominal = cells_n; out->cells_executed = cells_e; out->kernel_ms = kernel_ms; out->dp_ms = dp_ms;
    return ADK_OK;
}

}  // namespace adk
