// adk_mt.cpp — implementation of the namespace-MT mirror (adk_mt.hpp) over the C ABI (include/adk.h).
// Marshalling only: AoS long double points -> SoA double, one C call, result widened back.
#include "adk_mt.hpp"

#include <algorithm>
#include <cmath>
#include <stdexcept>
#include <string>

#include "../../include/adk.h"

namespace MT
{
namespace
{
struct ThreadSolver
{
    adk_ctx* ctx = nullptr;
    int device = -1;
    B200::LastSolveStats last{0, 0, 0, 0, 0.0, 0.0};
    ~ThreadSolver() { if (ctx) adk_destroy(ctx); }
    adk_ctx* get()
    {
        if (!ctx)
        {
            const int rc = adk_create(&ctx, device);
            if (rc != ADK_OK)
                throw std::runtime_error(std::string("adk: no usable CUDA device (") + adk_last_error(nullptr) +
                                         "); this build has no CPU fallback");
        }
        return ctx;
    }
};

ThreadSolver& solver()
{
    static thread_local ThreadSolver s;
    return s;
}

struct Soa
{
    std::vector<double> x, y;
    explicit Soa(const std::vector<CM::Point2>& pts) : x(pts.size()), y(pts.size())
    {
        for (size_t i = 0; i < pts.size(); ++i)
        {
            if (pts[i].myIndex != i) throw std::invalid_argument("adk: points[i].myIndex must equal i");
            x[i] = (double) pts[i].myX;
            y[i] = (double) pts[i].myY;
        }
    }
};

void check(int rc, adk_ctx* ctx, const char* what)
{
    if (rc != ADK_OK) throw std::runtime_error(std::string(what) + " failed: " + adk_last_error(ctx));
}

uint64_t narrow_count(size_t k) { return k == (size_t) -1 ? ADK_KMAX_INF : (uint64_t) k; }

std::optional<ConvexArea> widen(const adk_result& r, const std::vector<uint32_t>& hull, std::optional<BenchmarkInfo>& info, ThreadSolver& s)
{
    s.last = B200::LastSolveStats{r.cells_nominal, r.cells_executed, r.states, r.pairs, r.kernel_ms, r.dp_ms};
    if (info)
    {
        info->myAllocatedEntriesCount = r.entries_alloc;
        info->myRequiredEntriesCount = r.entries_required;
    }
    if (!r.found) return std::nullopt;
    ConvexArea out;
    out.myHullArea = r.area;
    out.myPointsCount = r.count;
    if (r.diam_valid) out.myDiameterOpt = Diameter{r.diam[0], r.diam[1]};
    out.myHullIndices.assign(hull.begin(), hull.begin() + r.hull_len);
    return out;
}

std::optional<ConvexArea> run_area_diameter(const std::vector<CM::Point2>& pts, std::optional<BenchmarkInfo>& info, size_t k,
                                            long double a, long double d, bool reconstruct)
{
    if (std::min(pts.size(), k) < 3)
    {   // AntipodalOptimized.cpp:461-464 returns before touching the counters
        return std::nullopt;
    }
    ThreadSolver& s = solver();
    Soa soa(pts);
    adk_result r;
    std::vector<uint32_t> hull(pts.size() + 4);
    const uint32_t flags = (reconstruct ? ADK_FLAG_RECONSTRUCT : 0u) | (info ? ADK_FLAG_COUNTERS : 0u);
    check(adk_solve_area_diameter(s.get(), soa.x.data(), soa.y.data(), (uint32_t) pts.size(), narrow_count(k), (double) a, (double) d,
                                  flags, &r, hull.data()), s.ctx, "adk_solve_area_diameter");
    return widen(r, hull, info, s);
}
}  // namespace

// ------------------------------------------------------------------------------------------------- scalar layer (host copies)
namespace CM
{
bool Point2::operator==(const Point2& o) const { return o.myX == myX && o.myY == myY && o.myIndex == myIndex; }

long double Dot(const Point2& a, const Point2& b) { return a.myX * b.myX + a.myY * b.myY; }

long double Distance2(const Point2& a, const Point2& b)
{
    const long double dx = b.myX - a.myX, dy = b.myY - a.myY;
    return dx * dx + dy * dy;
}

bool IsCloseToZero(long double v, long double eps) { return (v < 0 ? -v : v) < eps; }

long double SignedArea(const Point2& a, const Point2& b, const Point2& c)
{
    return 0.5l * ((b.myX - a.myX) * (c.myY - a.myY) - (c.myX - a.myX) * (b.myY - a.myY));
}

bool AreCollinear(const Point2& a, const Point2& b, const Point2& c) { return IsCloseToZero(SignedArea(a, b, c)); }

ORIENTATION Orientation(const Point2& a, const Point2& b, const Point2& c)
{
    const long double s = SignedArea(a, b, c);
    if (IsCloseToZero(s)) return ORIENTATION::COLLINEAR;
    return s > 0 ? ORIENTATION::LEFT_TURN : ORIENTATION::RIGHT_TURN;
}

long double Angle(const Point2& ref, const Point2& p)
{
    const long double a = ::atan2l(p.myY - ref.myY, p.myX - ref.myX);
    return a < 0 ? a + 2.f * M_PI : a;
}

long double ProjectedDistance2(const Point2& s, const Point2& e, const Point2& p)
{
    if (p.myIndex == s.myIndex) return 0;
    const long double d2 = Distance2(s, e);
    if (p.myIndex == e.myIndex) return d2;
    const long double dot = (e.myX - s.myX) * (p.myX - s.myX) + (e.myY - s.myY) * (p.myY - s.myY);
    return std::clamp((dot * dot) / d2, 0.l, d2);
}

long double PointLinePseudoDistance(const Point2& s, const Point2& e, const Point2& p)
{
    const long double v = (p.myX - s.myX) * (-e.myY + s.myY) + (p.myY - s.myY) * (e.myX - s.myX);
    return v < 0 ? -v : v;
}
}  // namespace CM

bool Diameter::operator==(const Diameter& o) const
{
    return (myFirstIndex == o.myFirstIndex && mySecondIndex == o.mySecondIndex) ||
           (mySecondIndex == o.myFirstIndex && myFirstIndex == o.mySecondIndex);
}

bool ArePointsClockwise(const CM::Point2& ref, const CM::Point2& a, const CM::Point2& b)
{
    const long double fa = CM::Angle(ref, a), fb = CM::Angle(ref, b);
    const long double d = fa - fb;
    if ((d < 0 ? -d : d) < 0.000001l && CM::AreCollinear(ref, a, b)) return CM::Distance2(ref, a) < CM::Distance2(ref, b);
    return fa > fb;
}

void GetHullPoints(const std::vector<size_t>& idx, const std::vector<CM::Point2>& pts, std::vector<CM::Point2>& out)
{
    out.clear();
    out.reserve(idx.size());
    for (size_t i : idx) out.push_back(pts[i]);
}

// ------------------------------------------------------------------------------------------------- precompute on the GPU
void CountPointsBelowAllSegments(const std::vector<CM::Point2>& pts, const std::vector<std::vector<CM::Point2>>&, PointsInTriangleCache& out)
{
    // The clockwise lists are not needed: the tables come from their definitions on the device (k_count_tables).
    ThreadSolver& s = solver();
    Soa soa(pts);
    const size_t n = pts.size();
    std::vector<int32_t> below(n * n), coll(n * n), rb(n);
    check(adk_count_tables(s.get(), soa.x.data(), soa.y.data(), (uint32_t) n, below.data(), coll.data(), rb.data()), s.ctx, "adk_count_tables");
    out.myPointsBelowSegmentCount.assign(n, std::vector<int>(n));
    out.myCollinearPointsCount.assign(n, std::vector<int>(n));
    out.myPointsRightBelowCount.assign(rb.begin(), rb.end());
    for (size_t i = 0; i < n; ++i)
        for (size_t j = 0; j < n; ++j)
        {
            out.myPointsBelowSegmentCount[i][j] = below[i * n + j];
            out.myCollinearPointsCount[i][j] = coll[i * n + j];
        }
}

void SortPointsClockWiseAroundPoint(const CM::Point2& ref, std::vector<CM::Point2>& io)
{
    // point 0 = the centre, points 1.. = the list; row 0 of the device's clockwise order is the answer
    ThreadSolver& s = solver();
    const size_t m = io.size();
    if (m < 2) return;
    std::vector<double> x(m + 1), y(m + 1);
    x[0] = (double) ref.myX; y[0] = (double) ref.myY;
    for (size_t i = 0; i < m; ++i) { x[i + 1] = (double) io[i].myX; y[i + 1] = (double) io[i].myY; }
    std::vector<int32_t> order((m + 1) * m);
    check(adk_cw_order(s.get(), x.data(), y.data(), (uint32_t) (m + 1), order.data()), s.ctx, "adk_cw_order");
    std::vector<CM::Point2> sorted(m);
    for (size_t p = 0; p < m; ++p) sorted[p] = io[(size_t) order[p] - 1];
    io.swap(sorted);
}

std::optional<Diameter> ComputeDiameter(const std::vector<CM::Point2>& hull)
{
    // Standalone helper of the public surface; the solvers compute their diameter on the device.
    if (hull.size() < 2) return std::nullopt;
    size_t ba = 0, bb = 0; long double best = 0;
    for (size_t a = 0; a < hull.size(); ++a)
        for (size_t b = a + 1; b < hull.size(); ++b)
        {
            const long double d = CM::Distance2(hull[a], hull[b]);
            if (d > best) { best = d; ba = a; bb = b; }
        }
    return Diameter{hull[ba].myIndex, hull[bb].myIndex};
}

// ------------------------------------------------------------------------------------------------- path A
std::optional<ConvexArea> EppsteinAlgorithmWithBenchmarkInfo(const std::vector<CM::Point2>& pts, std::optional<BenchmarkInfo>& info,
                                                            size_t k, long double a, bool reconstruct, bool)
{
    if (pts.size() < 3 || k < 3) return std::nullopt;   // Eppstein.cpp:137-140
    ThreadSolver& s = solver();
    Soa soa(pts);
    adk_result r;
    std::vector<uint32_t> hull(pts.size() + 4);
    const uint32_t flags = (reconstruct ? ADK_FLAG_RECONSTRUCT : 0u) | (info ? ADK_FLAG_COUNTERS : 0u);
    check(adk_solve_area(s.get(), soa.x.data(), soa.y.data(), (uint32_t) pts.size(), narrow_count(k), (double) a, flags, &r, hull.data()),
          s.ctx, "adk_solve_area");
    return widen(r, hull, info, s);
}

std::optional<ConvexArea> EppsteinAlgorithm(const std::vector<CM::Point2>& pts, size_t k, long double a, bool reconstruct, bool opt)
{
    std::optional<BenchmarkInfo> none;
    return EppsteinAlgorithmWithBenchmarkInfo(pts, none, k, a, reconstruct, opt);
}

// ------------------------------------------------------------------------------------------------- path AD
std::optional<ConvexArea> AntipodalOptimizedAlgorithmWithBenchmarkInfo(const std::vector<CM::Point2>& pts, std::optional<BenchmarkInfo>& info,
                                                                      size_t k, long double a, long double d, bool reconstruct, bool)
{
    return run_area_diameter(pts, info, k, a, d, reconstruct);
}

std::optional<ConvexArea> AntipodalOptimizedAlgorithm(const std::vector<CM::Point2>& pts, size_t k, long double a, long double d,
                                                      bool reconstruct, bool opt)
{
    std::optional<BenchmarkInfo> none;
    return AntipodalOptimizedAlgorithmWithBenchmarkInfo(pts, none, k, a, d, reconstruct, opt);
}

// The 4-index "basic" variant (Antipodal.cpp:369-464) returns the same (count, area, hull cycle, diameter) as the
// optimized one on every input the reference tests (src/test/main.cpp:105-117); it is served by the same kernels.
std::optional<ConvexArea> AntipodalAlgorithmWithBenchmarkInfo(const std::vector<CM::Point2>& pts, std::optional<BenchmarkInfo>& info,
                                                             size_t k, long double a, long double d, bool reconstruct, bool)
{
    return run_area_diameter(pts, info, k, a, d, reconstruct);
}

std::optional<ConvexArea> AntipodalAlgorithm(const std::vector<CM::Point2>& pts, size_t k, long double a, long double d,
                                             bool reconstruct, bool opt)
{
    std::optional<BenchmarkInfo> none;
    return AntipodalAlgorithmWithBenchmarkInfo(pts, none, k, a, d, reconstruct, opt);
}

namespace B200
{
void SetDevice(int ordinal)
{
    ThreadSolver& s = solver();
    if (s.ctx && s.device != ordinal) { adk_destroy(s.ctx); s.ctx = nullptr; }
    s.device = ordinal;
}
LastSolveStats GetLastSolveStats() { return solver().last; }
}  // namespace B200
}  // namespace MT
