// A caller of the public HELPERS of the reference's src/common/Utils.h, written against the reference's declarations and
// linked against this build's libadk_host.so: MT::ForAllAntipodalPairs (Utils.h:57-109), MT::ComputeDiameter (Utils.cpp:318-338)
// and MT::PointsInTriangle over a caller-supplied cache (Utils.h:45-49).  None of them needs a GPU.
// stdin:  n, then n x y | m, then m hull positions | T, then T triples | the three count tables (n*n, n*n, n)
// stdout: "pairs" k then k pairs | "diameter" a b | "pit" T counts
#include <cstdio>
#include <vector>

#include "Utils.h"

int main()
{
    size_t n = 0, m = 0, T = 0;
    if (scanf("%zu", &n) != 1) return 1;
    std::vector<MT::CM::Point2> pts;
    for (size_t i = 0; i < n; ++i) { double x, y; if (scanf("%lf %lf", &x, &y) != 2) return 1; pts.emplace_back((long double) x, (long double) y, i); }
    if (scanf("%zu", &m) != 1) return 1;
    std::vector<size_t> hull(m);
    for (size_t i = 0; i < m; ++i) if (scanf("%zu", &hull[i]) != 1) return 1;
    if (scanf("%zu", &T) != 1) return 1;
    std::vector<size_t> tri(3 * T);
    for (size_t i = 0; i < 3 * T; ++i) if (scanf("%zu", &tri[i]) != 1) return 1;
    MT::PointsInTriangleCache cache;
    cache.myPointsBelowSegmentCount.assign(n, std::vector<int>(n));
    cache.myCollinearPointsCount.assign(n, std::vector<int>(n));
    cache.myPointsRightBelowCount.assign(n, 0);
    for (size_t i = 0; i < n; ++i) for (size_t j = 0; j < n; ++j) if (scanf("%d", &cache.myPointsBelowSegmentCount[i][j]) != 1) return 1;
    for (size_t i = 0; i < n; ++i) for (size_t j = 0; j < n; ++j) if (scanf("%d", &cache.myCollinearPointsCount[i][j]) != 1) return 1;
    for (size_t i = 0; i < n; ++i) if (scanf("%d", &cache.myPointsRightBelowCount[i]) != 1) return 1;

    std::vector<MT::CM::Point2> poly;
    MT::GetHullPoints(hull, pts, poly);
    std::vector<std::pair<size_t, size_t>> pairs;
    if (m > 2) MT::ForAllAntipodalPairs(poly, [&](const size_t a, const size_t b) { pairs.emplace_back(a, b); });
    printf("pairs %zu\n", pairs.size());
    for (const auto& p : pairs) printf("%zu %zu\n", p.first, p.second);
    const auto d = MT::ComputeDiameter(poly);
    if (d) printf("diameter %zu %zu\n", d->myFirstIndex, d->mySecondIndex); else printf("diameter none\n");
    printf("pit %zu\n", T);
    for (size_t t = 0; t < T; ++t) printf("%d\n", MT::PointsInTriangle(pts[tri[3 * t]], pts[tri[3 * t + 1]], pts[tri[3 * t + 2]], cache));
    return 0;
}
