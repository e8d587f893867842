/* Force-included when the reference's UNMODIFIED callers (src/test/main.cpp, src/py-binding/main.cpp, ...) are
 * compiled against this repo's libadk_host.so.  Same libstdc++ shims as oracle/ref_compat.h, plus a replacement for
 * the reference's CMake-configured Constants.h (the committed copy carries the author's /Users/... paths): defining
 * its include guard here makes the reference's copy empty. */
#ifndef ADK_DROPIN_COMPAT_H
#define ADK_DROPIN_COMPAT_H
#include <cmath>
#include <math.h>
#include <array>
#include <optional>
#include <algorithm>
#include <unordered_map>
#include <functional>
#include <tuple>
#include <numeric>
#include <cstdlib>
#include <cstdint>
#include <limits>
#include <vector>
#include <string>
#include <string_view>
#include <stdexcept>
namespace std { using ::fabsl; using ::atan2l; using ::sqrtl; }
#define MASTER_THESIS_CONSTANTS_H_IN_H
#ifndef ADK_DROPIN_DATA
#define ADK_DROPIN_DATA "/root/repo/tests/dropin/_data"
#endif
namespace MT { namespace Constants {
constexpr std::string_view PATH_TO_TESTS = ADK_DROPIN_DATA "/tests";
constexpr std::string_view PATH_TO_EXPERIMENTS = ADK_DROPIN_DATA "/experiments";
constexpr std::string_view PATH_TO_BENCHMARK_RUNS = ADK_DROPIN_DATA "/benchmark_runs.json";
constexpr std::string_view PATH_TO_BENCHMARK_RESULTS = ADK_DROPIN_DATA "/benchmark_results.json";
constexpr std::array<long double, 5> SYNTHETIC_BENCHMARK_DIAMETERS = {2, 3, 4, 5, 6};
constexpr size_t SYNTHETIC_BENCHMARK_UNIFORM_ITERATIONS = 100;
constexpr size_t SYNTHETIC_BENCHMARK_GAUSSIAN_ITERATIONS = 100;
constexpr size_t DENSITIES_COUNT = 11;
constexpr size_t STDDEVS_COUNT = 13;
constexpr bool ENABLE_OPTIMIZATIONS_WITH_SYNTHETIC_DATA = false;
} }
#endif
