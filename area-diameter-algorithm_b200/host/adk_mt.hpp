// adk_mt.hpp — host-side C++17 mirror of the reference's public solver interface (namespace MT), implemented on
// top of the C ABI in include/adk.h.  A caller written against the reference's src/common headers
// (src/bench/main.cpp, src/test/main.cpp, src/main/main.cpp, src/py-binding/main.cpp) compiles and links against
// this library unchanged: same namespaces, type names, member names, default arguments and nullopt behaviour.
//
// Interface being mirrored (reference repo paths):
//   CM::Point2, ORIENTATION, scalar predicates ........ src/common/CustomMath.h:11-58
//   Diameter, ConvexArea, PointsInTriangleCache,
//   BenchmarkInfo, ComputeDiameter, GetHullPoints ..... src/common/Utils.h:14-55,111-113
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
// Device-side statistics of the calling thread's most recent solve.
struct LastSolveStats { uint64_t cellsNominal, cellsExecuted, states, pairs; double kernelMs, dpMs; };
LastSolveStats GetLastSolveStats();
}  // namespace B200
}  // namespace MT

#endif
