// TEST INFRASTRUCTURE — not part of the product.
//
// Thin extern "C" face over the UNMODIFIED reference solver so that tests/ and
// bench.py's cpu_baseline / --impl reference legs can call it through ctypes.
// Compiled by oracle/build_ref.sh together with the reference's own
// src/common/{CustomMath,Utils,Eppstein,Antipodal,AntipodalOptimized}.cpp
// (read in place from /root/reference; nothing is copied) into
// oracle/_ref/libadkref*.so.  Only tests/, __graft_entry__.smoke() and
// bench.py may load that library.
#include <cstdint>
#include <cstring>
#include <cfloat>
#include <chrono>

#include "Eppstein.h"
#include "Antipodal.h"
#include "AntipodalOptimized.h"
#include "Utils.h"

extern "C" {

struct adkref_result
{
    int32_t found;
    uint32_t count;
    double area;
    uint32_t hull_len;
    uint32_t diam_valid;
    uint32_t diam[2];
    int64_t entries_alloc;
    int64_t entries_required;
    double solve_ms;
};

static void locToPoints(const double* x, const double* y, uint32_t n, std::vector<MT::CM::Point2>& out)
{
    out.reserve(n);
    for (uint32_t i = 0; i < n; ++i) out.emplace_back((long double) x[i], (long double) y[i], (size_t) i);
}

// LDBL_MANT_DIG of the build: 64 for the native x87 build, 53 for the binary64 variant.
int adkref_mant_dig(void) { return LDBL_MANT_DIG; }

// algo: 0 = EppsteinAlgorithm, 1 = AntipodalAlgorithm, 2 = AntipodalOptimizedAlgorithm (enum Algorithm, TestingUtils.h:11-17)
int adkref_solve(int algo, const double* x, const double* y, uint32_t n, uint64_t kmax, double amax, double dmax,
                 int reconstruct, int optimise, adkref_result* out, uint32_t* hull, uint32_t hull_cap)
{
    std::vector<MT::CM::Point2> points;
    locToPoints(x, y, n, points);
    std::optional<MT::BenchmarkInfo> info = MT::BenchmarkInfo{};
    std::optional<MT::ConvexArea> res;
    const size_t k = kmax == UINT64_MAX ? (size_t) -1 : (size_t) kmax;
    const long double a = amax, d = dmax;
    const auto t0 = std::chrono::steady_clock::now();
    switch (algo)
    {
        case 0: res = MT::EppsteinAlgorithmWithBenchmarkInfo(points, info, k, a, reconstruct != 0, optimise != 0); break;
        case 1: res = MT::AntipodalAlgorithmWithBenchmarkInfo(points, info, k, a, d, reconstruct != 0, optimise != 0); break;
        case 2: res = MT::AntipodalOptimizedAlgorithmWithBenchmarkInfo(points, info, k, a, d, reconstruct != 0, optimise != 0); break;
        default: return 1;
    }
    const auto t1 = std::chrono::steady_clock::now();
    std::memset(out, 0, sizeof(*out));
    out->solve_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    out->entries_alloc = info->myAllocatedEntriesCount;
    out->entries_required = info->myRequiredEntriesCount;
    if (!res) return 0;
    out->found = 1;
    out->count = (uint32_t) res->myPointsCount;
    out->area = (double) res->myHullArea;
    out->hull_len = (uint32_t) res->myHullIndices.size();
    for (uint32_t i = 0; i < out->hull_len && i < hull_cap; ++i) hull[i] = (uint32_t) res->myHullIndices[i];
    if (res->myDiameterOpt)
    {
        out->diam_valid = 1;
        out->diam[0] = (uint32_t) res->myDiameterOpt->myFirstIndex;
        out->diam[1] = (uint32_t) res->myDiameterOpt->mySecondIndex;
    }
    return 0;
}

// The reference's clockwise neighbour lists (Utils.cpp:311-316): order[i*(n-1)+p] = index of the p-th point around i.
int adkref_cw_order(const double* x, const double* y, uint32_t n, int32_t* order)
{
    std::vector<MT::CM::Point2> points;
    locToPoints(x, y, n, points);
    for (uint32_t i = 0; i < n; ++i)
    {
        std::vector<MT::CM::Point2> others;
        for (uint32_t j = 0; j < n; ++j) if (j != i) others.push_back(points[j]);
        MT::SortPointsClockWiseAroundPoint(points[i], others);
        for (uint32_t p = 0; p + 1 < n; ++p) order[(size_t) i * (n - 1) + p] = (int32_t) others[p].myIndex;
    }
    return 0;
}

// The reference's three count tables (Utils.cpp:74-143), row-major n x n, plus PointsInTriangle on request.
int adkref_count_tables(const double* x, const double* y, uint32_t n, int32_t* below, int32_t* coll, int32_t* right_below)
{
    std::vector<MT::CM::Point2> points;
    locToPoints(x, y, n, points);
    std::vector<std::vector<MT::CM::Point2>> cw(n);
    for (uint32_t i = 0; i < n; ++i)
    {
        for (uint32_t j = 0; j < n; ++j) if (j != i) cw[i].push_back(points[j]);
        MT::SortPointsClockWiseAroundPoint(points[i], cw[i]);
    }
    MT::PointsInTriangleCache cache;
    MT::CountPointsBelowAllSegments(points, cw, cache);
    for (uint32_t i = 0; i < n; ++i)
    {
        right_below[i] = cache.myPointsRightBelowCount[i];
        for (uint32_t j = 0; j < n; ++j)
        {
            below[(size_t) i * n + j] = cache.myPointsBelowSegmentCount[i][j];
            coll[(size_t) i * n + j] = cache.myCollinearPointsCount[i][j];
        }
    }
    return 0;
}

// PointsInTriangle (Utils.cpp:251-293) for a list of triples; tri = 3*m indices.
int adkref_points_in_triangle(const double* x, const double* y, uint32_t n, const uint32_t* tri, uint32_t m, int32_t* out)
{
    std::vector<MT::CM::Point2> points;
    locToPoints(x, y, n, points);
    std::vector<std::vector<MT::CM::Point2>> cw(n);
    for (uint32_t i = 0; i < n; ++i)
    {
        for (uint32_t j = 0; j < n; ++j) if (j != i) cw[i].push_back(points[j]);
        MT::SortPointsClockWiseAroundPoint(points[i], cw[i]);
    }
    MT::PointsInTriangleCache cache;
    MT::CountPointsBelowAllSegments(points, cw, cache);
    for (uint32_t t = 0; t < m; ++t)
        out[t] = MT::PointsInTriangle(points[tri[3 * t]], points[tri[3 * t + 1]], points[tri[3 * t + 2]], cache);
    return 0;
}

// ForAllAntipodalPairs (Utils.h:57-109) over the polygon hull[0..m) of the point set: the (first, second) POSITIONS it reports,
// in call order; returns their number (pairs beyond cap are counted, not stored).
int adkref_antipodal_pairs(const double* x, const double* y, uint32_t n, const uint32_t* hull, uint32_t m, uint32_t* out_pairs, uint32_t cap)
{
    std::vector<MT::CM::Point2> points, poly;
    locToPoints(x, y, n, points);
    for (uint32_t t = 0; t < m; ++t) poly.push_back(points[hull[t]]);
    uint32_t cnt = 0;
    MT::ForAllAntipodalPairs(poly, [&](const size_t a, const size_t b) {
        if (cnt < cap) { out_pairs[2 * cnt] = (uint32_t) a; out_pairs[2 * cnt + 1] = (uint32_t) b; }
        ++cnt;
    });
    return (int) cnt;
}

// ComputeDiameter (Utils.cpp:318-338) of the same polygon: the two point indices.
int adkref_compute_diameter(const double* x, const double* y, uint32_t n, const uint32_t* hull, uint32_t m, uint32_t* out_pair)
{
    std::vector<MT::CM::Point2> points, poly;
    locToPoints(x, y, n, points);
    for (uint32_t t = 0; t < m; ++t) poly.push_back(points[hull[t]]);
    const auto d = MT::ComputeDiameter(poly);
    if (!d) return 0;
    out_pair[0] = (uint32_t) d->myFirstIndex; out_pair[1] = (uint32_t) d->mySecondIndex;
    return 1;
}

}
