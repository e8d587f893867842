// adk_mt.hpp — host-side C++17 mirror of the reference's public solver interface (namespace MT), implemented on
// top of the C ABI in include/adk.h.  A caller written against the reference's src/common headers
// (src/bench/main.cpp, src/test/main.cpp, src/main/main.cpp, src/py-binding/main.cpp) compiles and links against
// this library unchanged: same namespaces, type names, member names, default arguments and nullopt behaviour.
//
// Interface being mirrored (reference repo paths):
//   CM::Point2, ORIENTATION, scalar predicates ........ src/common/CustomMath.h:11-58
//   Diameter, ConvexArea, PointsInTriangleCache,
//   BenchmarkInfo, ComputeDiameter, GetHullPoints ..... src/common/Utils.h:14-55,111-113
//   PointsInTriangle, ForAllAntipodalPairs ............ src/common/Utils.h:45-49,57-109
//   EppsteinAlgorithm[WithBenchmarkInfo] .............. src/common/Eppstein.h:11-24
//   AntipodalAlgorithm[WithBenchmarkInfo] ............. src/common/Antipodal.h:10-25
//   AntipodalOptimizedAlgorithm[WithBenchmarkInfo] .... src/common/AntipodalOptimized.h:10-25
//
// The solver bodies run on the GPU; only marshalling (Point2 AoS long double -> SoA double) happens here.
// Errors from the device layer (no GPU, out of memory) surface as std::runtime_error: there is no CPU fallback.
#ifndef ADK_MT_HPP
#define ADK_MT_HPP

#include <cstddef>
#include <cstdint>
#include <functional>
#include <limits>
#include <optional>
#include <vector>

namespace MT
{
namespace CM
{
struct Point2
{
    Point2() = default;
    Point2(long double anX, long double anY, size_t anIndex) : myX(anX), myY(anY), myIndex(anIndex) {}
    long double myX, myY;
    size_t myIndex;
    bool operator==(const Point2&) const;
};

constexpr auto SortPointsHorizontally = [](const Point2& a, const Point2& b) {
    return (a.myX < b.myX) || (a.myX == b.myX && a.myY < b.myY);
};
constexpr auto SortPointsVertically = [](const Point2& a, const Point2& b) { return a.myY < b.myY; };

enum class ORIENTATION { RIGHT_TURN = -1, COLLINEAR = 0, LEFT_TURN = 1 };

template <typename T> int Sign(const T& aValue) { return (T(0) < aValue) - (aValue < T(0)); }

long double Dot(const Point2& aFirstPoint, const Point2& aSecondPoint);
long double Distance2(const Point2& aFirstPoint, const Point2& aSecondPoint);
bool IsCloseToZero(long double aValue, long double anEpsilon = 0.0000000001l);
bool AreCollinear(const Point2& aFirstPoint, const Point2& aSecondPoint, const Point2& aThirdPoint);
long double SignedArea(const Point2& aFirstPoint, const Point2& aSecondPoint, const Point2& aThirdPoint);
ORIENTATION Orientation(const Point2& aFirstPoint, const Point2& aSecondPoint, const Point2& aThirdPoint);
long double Angle(const Point2& aReferencePoint, const Point2& aPoint);
long double ProjectedDistance2(const Point2& aStartPoint, const Point2& anEndPoint, const Point2& aPoint);
long double PointLinePseudoDistance(const Point2& aStartPoint, const Point2& anEndPoint, const Point2& aPoint);
}  // namespace CM

constexpr auto INVALID_INDEX = (size_t) -1;

struct Diameter
{
    bool operator==(const Diameter&) const;   // unordered pair equality
    size_t myFirstIndex{INVALID_INDEX}, mySecondIndex{INVALID_INDEX};
};

struct ConvexArea
{
    long double myHullArea{std::numeric_limits<long double>::infinity()};
    size_t myPointsCount{0};
    std::optional<Diameter> myDiameterOpt;
    std::vector<size_t> myHullIndices{};
};

struct PointsInTriangleCache
{
    std::vector<std::vector<int>> myPointsBelowSegmentCount;
    std::vector<std::vector<int>> myCollinearPointsCount;
    std::vector<int> myPointsRightBelowCount;
};

struct BenchmarkInfo
{
    int64_t myAllocatedEntriesCount{-1};
    int64_t myRequiredEntriesCount{-1};
};

// Geometry precompute entry points kept for source compatibility; both run on the GPU.
void CountPointsBelowAllSegments(const std::vector<CM::Point2>& somePoints,
                                 const std::vector<std::vector<CM::Point2>>& someClockWiseSortedPoints,
                                 PointsInTriangleCache& anOutCache);
void SortPointsClockWiseAroundPoint(const CM::Point2& aReferencePoint, std::vector<CM::Point2>& someOutClockWiseSortedPoints);
bool ArePointsClockwise(const CM::Point2& aReferencePoint, const CM::Point2& aFirstPoint, const CM::Point2& aSecondPoint);
std::optional<Diameter> ComputeDiameter(const std::vector<CM::Point2>& somePoints);

// O(1) number of input points strictly inside a triangle, from the three count tables (Utils.h:45-49, Utils.cpp:251-293).
int PointsInTriangle(const CM::Point2& aFirstPoint, const CM::Point2& aSecondPoint, const CM::Point2& aThirdPoint,
                     const PointsInTriangleCache& aCache);

// Rotating calipers over the vertices of a convex polygon given in order (Utils.h:57-109): aFunction(p, q) is called for every
// antipodal vertex pair, in the reference's order (edge p-1 -> p against the vertices farthest from it; an edge parallel to the
// caliper reports both of its ends).  aMaxLength restricts the walk to the first aMaxLength vertices.  Unlike the reference the
// walk is bounded (4 n + 8 steps): on an exactly degenerate polygon it ends instead of circling.
template <typename Container>
void ForAllAntipodalPairs(const Container& somePoints, const std::function<void(const size_t, const size_t)>& aFunction,
                          size_t aMaxLength = (size_t) -1)
{
    const size_t n = aMaxLength == (size_t) -1 ? somePoints.size() : aMaxLength;
    if (n < 3) return;
    const auto succ = [n](size_t v) { return v + 1 == n ? (size_t) 0 : v + 1; };
    // height gained over the edge p -> succ(p) when the far vertex moves from q to succ(q)
    const auto gain = [&](size_t p, size_t q)
    {
        return CM::SignedArea(somePoints[p], somePoints[succ(p)], somePoints[succ(q)]) -
               CM::SignedArea(somePoints[p], somePoints[succ(p)], somePoints[q]);
    };
    const auto rises = [&](size_t p, size_t q)
    {
        return CM::SignedArea(somePoints[p], somePoints[succ(p)], somePoints[succ(q)]) >
               CM::SignedArea(somePoints[p], somePoints[succ(p)], somePoints[q]);
    };
    size_t budget = 4 * n + 8;
    size_t p = n - 1, q = 0;
    while (rises(p, q) && budget) { q = succ(q); --budget; }
    const size_t qFirst = q;
    while (q != 0 && budget)
    {
        --budget;
        p = succ(p);
        aFunction(p, q);
        while (rises(p, q) && budget)
        {
            --budget;
            q = succ(q);
            if (p == qFirst && q == 0) break;          // back at the first pair
            aFunction(p, q);
        }
        if (CM::IsCloseToZero(gain(p, q)))
        {   // the edge succ(q) is parallel to the caliper: its other end is antipodal to p as well
            if (p == qFirst && q == n - 1) break;
            aFunction(p, succ(q));
        }
    }
}

void GetHullPoints(const std::vector<size_t>& someIndices, const std::vector<CM::Point2>& somePoints,
                   std::vector<CM::Point2>& someOutHullPoints);

// ---- path A ---------------------------------------------------------------------------------------------------
std::optional<ConvexArea> EppsteinAlgorithm(const std::vector<CM::Point2>& somePoints,
                                            size_t aMaxPointsCount = (size_t) -1,
                                            long double aMaxArea = std::numeric_limits<long double>::max(),
                                            bool aShouldReconstructHull = false,
                                            bool aShouldEnableOptimizations = false);

std::optional<ConvexArea> EppsteinAlgorithmWithBenchmarkInfo(const std::vector<CM::Point2>& somePoints,
                                                            std::optional<BenchmarkInfo>& anOutBenchmarkInfoOpt,
                                                            size_t aMaxPointsCount = (size_t) -1,
                                                            long double aMaxArea = std::numeric_limits<long double>::max(),
                                                            bool aShouldReconstructHull = false,
                                                            bool aShouldEnableOptimizations = false);

// ---- path AD (basic and optimized share the GPU kernels) --------------------------------------------
std::optional<ConvexArea> AntipodalAlgorithm(const std::vector<CM::Point2>& somePoints,
                                             size_t aMaxPointsCount = (size_t) -1,
                                             long double aMaxArea = std::numeric_limits<long double>::max(),
                                             long double aMaxDiameter = std::numeric_limits<long double>::max(),
                                             bool aShouldReconstructHull = false,
                                             bool aShouldEnableOptimizations = false);

std::optional<ConvexArea> AntipodalAlgorithmWithBenchmarkInfo(const std::vector<CM::Point2>& somePoints,
                                                             std::optional<BenchmarkInfo>& anOutBenchmarkInfoOpt,
                                                             size_t aMaxPointsCount = (size_t) -1,
                                                             long double aMaxArea = std::numeric_limits<long double>::max(),
                                                             long double aMaxDiameter = std::numeric_limits<long double>::max(),
                                                             bool aShouldReconstructHull = false,
                                                             bool aShouldEnableOptimizations = false);

std::optional<ConvexArea> AntipodalOptimizedAlgorithm(const std::vector<CM::Point2>& somePoints,
                                                      size_t aMaxPointsCount = (size_t) -1,
                                                      long double aMaxArea = std::numeric_limits<long double>::max(),
                                                      long double aMaxDiameter = std::numeric_limits<long double>::max(),
                                                      bool aShouldReconstructHull = false,
                                                      bool aShouldEnableOptimizations = false);

std::optional<ConvexArea> AntipodalOptimizedAlgorithmWithBenchmarkInfo(const std::vector<CM::Point2>& somePoints,
                                                                      std::optional<BenchmarkInfo>& anOutBenchmarkInfoOpt,
                                                                      size_t aMaxPointsCount = (size_t) -1,
                                                                      long double aMaxArea = std::numeric_limits<long double>::max(),
                                                                      long double aMaxDiameter = std::numeric_limits<long double>::max(),
                                                                      bool aShouldReconstructHull = false,
                                                                      bool aShouldEnableOptimizations = false);

// ---- additions of this build (not in the reference) -----------------------------------------------------------------
namespace B200
{
// Selects the CUDA device used by the calling thread's solver context (default: the current device).
void SetDevice(int aDeviceOrdinal);
// Several devices of one box for the calling thread's solves: anchors / candidate pairs are sharded over them inside the
// library (adk_multi_*, one NCCL all-gather + one broadcast per solve).  The environment variable ADK_DEVICES ("0,1,2,3",
// or "all") sets the same default for every thread.  An empty list / one device = single-GPU path.
void SetDevices(const std::vector<int>& someDeviceOrdinals);
// Device-side statistics of the calling thread's most recent solve.
struct LastSolveStats { uint64_t cellsNominal, cellsExecuted, states, pairs; double kernelMs, dpMs; };
LastSolveStats GetLastSolveStats();
}  // namespace B200
}  // namespace MT

#endif
