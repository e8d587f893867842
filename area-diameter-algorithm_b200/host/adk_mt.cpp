// adk_mt.cpp — implementation of the namespace-MT mirror (adk_mt.hpp) over the C ABI (include/adk.h).
// Marshalling only: AoS long double points -> SoA double, one C call, result widened back.
#include "adk_mt.hpp"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <stdexcept>
#include <string>

#include "../../include/adk.h"

namespace MT
{
namespace
{
// ADK_DEVICES: "all", a count ("8") or a list ("0,1,2,3"); read once.
std::vector<int> env_devices()
{
    std::vector<int> out;
    const char* e = std::getenv("ADK_DEVICES");
    if (!e || !*e) return out;
    const std::string v(e);
    if (v == "all") { out.push_back(-1); return out; }       // resolved by adk_multi_create
    size_t at = 0;
    while (at < v.size())
    {
        size_t end = v.find(',', at);
        if (end == std::string::npos) end = v.size();
        if (end > at) out.push_back(std::atoi(v.substr(at, end - at).c_str()));
        at = end + 1;
    }
    if (out.size() == 1 && v.find(',') == std::string::npos && out[0] > 1)
    {   // a bare count
        const int cnt = out[0];
        out.clear();
        for (int i = 0; i < cnt; ++i) out.push_back(i);
    }
    return out;
}

struct ThreadSolver
{
    adk_ctx* ctx = nullptr;
    adk_multi* multi = nullptr;
    int device = -1;
    std::vector<int> devices = env_devices();
    B200::LastSolveStats last{0, 0, 0, 0, 0.0, 0.0};
    ~ThreadSolver() { if (ctx) adk_destroy(ctx); if (multi) adk_multi_destroy(multi); }
    bool several() const { return devices.size() > 1 || (devices.size() == 1 && devices[0] == -1); }
    adk_multi* get_multi()
    {
        if (!multi)
        {
            const bool all = devices.size() == 1 && devices[0] == -1;
            const int rc = adk_multi_create(&multi, all ? nullptr : devices.data(), all ? 0 : (int) devices.size());
            if (rc != ADK_OK)
                throw std::runtime_error(std::string("adk: cannot open the configured devices (") + adk_multi_last_error(nullptr) +
                                         "); this build has no CPU fallback");
        }
        return multi;
    }
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
    if (s.several())
    {
        adk_multi* m = s.get_multi();
        if (adk_multi_solve_area_diameter(m, soa.x.data(), soa.y.data(), (uint32_t) pts.size(), narrow_count(k), (double) a, (double) d,
                                          flags, &r, hull.data()) != ADK_OK)
            throw std::runtime_error(std::string("adk_multi_solve_area_diameter failed: ") + adk_multi_last_error(m));
        return widen(r, hull, info, s);
    }
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
    // Standalone helper of the public surface (Utils.cpp:318-338): the farthest antipodal pair, first one met wins on ties, as
    // in the reference; the solvers compute their diameter on the device with the same walk.
    if (hull.size() < 2) return std::nullopt;
    size_t ba = 0, bb = 0; long double best = 0;
    if (hull.size() == 2) { ba = 0; bb = 0; }      // the reference's walk needs three vertices; it reports the first vertex twice
    else
        ForAllAntipodalPairs(hull, [&](const size_t a, const size_t b) {
            const long double d = CM::Distance2(hull[a], hull[b]);
            if (d > best) { best = d; ba = a; bb = b; }
        });
    return Diameter{hull[ba].myIndex, hull[bb].myIndex};
}

int PointsInTriangle(const CM::Point2& p0, const CM::Point2& p1, const CM::Point2& p2, const PointsInTriangleCache& cache)
{
    // Inclusion-exclusion over the "points below the segment" counts (Utils.cpp:251-293): order the corners by (x, y), then the
    // middle one lies either below or above the long side.
    if (CM::AreCollinear(p0, p1, p2)) return 0;
    const CM::Point2* v[3] = {&p0, &p1, &p2};
    std::sort(v, v + 3, [](const CM::Point2* a, const CM::Point2* b) { return CM::SortPointsHorizontally(*a, *b); });
    const size_t l = v[0]->myIndex, m = v[1]->myIndex, r = v[2]->myIndex;
    const auto& B = cache.myPointsBelowSegmentCount;
    const auto& C = cache.myCollinearPointsCount;
    if (v[0]->myX == v[1]->myX) return B[m][r] - B[l][r] - C[l][r];                 // the two left corners share their x
    if (v[1]->myX == v[2]->myX) return B[l][r] - B[l][m] - C[l][m];                 // the two right corners share their x
    if (CM::Orientation(*v[2], *v[0], *v[1]) == CM::ORIENTATION::LEFT_TURN)         // the middle corner is below the long side
        return B[l][r] - B[l][m] - B[m][r] - C[l][m] - C[m][r] - cache.myPointsRightBelowCount[m] - 1;
    return B[l][m] + B[m][r] + cache.myPointsRightBelowCount[m] - B[l][r] - C[l][r];
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
    if (s.several())
    {
        adk_multi* m = s.get_multi();
        if (adk_multi_solve_area(m, soa.x.data(), soa.y.data(), (uint32_t) pts.size(), narrow_count(k), (double) a, flags, &r, hull.data()) != ADK_OK)
            throw std::runtime_error(std::string("adk_multi_solve_area failed: ") + adk_multi_last_error(m));
        return widen(r, hull, info, s);
    }
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
    if (s.multi) { adk_multi_destroy(s.multi); s.multi = nullptr; }
    s.device = ordinal;
    s.devices.clear();
}
void SetDevices(const std::vector<int>& ordinals)
{
    ThreadSolver& s = solver();
    if (s.multi) { adk_multi_destroy(s.multi); s.multi = nullptr; }
    s.devices = ordinals;
    if (ordinals.size() == 1) SetDevice(ordinals[0]);
}
LastSolveStats GetLastSolveStats() { return solver().last; }
}  // namespace B200
}  // namespace MT
