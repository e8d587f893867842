/* TEST INFRASTRUCTURE — not part of the product.
 *
 * Force-included (-include) when the UNMODIFIED reference sources under
 * /root/reference/src/common are compiled with g++/libstdc++ into oracle/_ref/.
 * The reference was written against clang/libc++ and leans on transitive
 * includes and on std::fabsl / std::atan2l / std::sqrtl, which libstdc++ does
 * not put in namespace std.  Nothing here changes the reference's arithmetic.
 *
 * With -DADK_REF_F64 (used together with g++ -mlong-double-64) the libm
 * "l" functions are mapped onto the double ones, which yields the reference
 * algorithm evaluated in IEEE binary64 instead of x87 80-bit: the
 * tie-sensitivity detector described in DESIGN.md.
 */
#ifndef ADK_ORACLE_REF_COMPAT_H
#define ADK_ORACLE_REF_COMPAT_H
#ifdef __cplusplus
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
#include <stdexcept>
#ifdef ADK_REF_F64
/* casts matter: std::atan2(long double, long double) would still call glibc's 80-bit atan2l */
#define fabsl(a) fabs((double) (a))
#define atan2l(a, b) atan2((double) (a), (double) (b))
#define sqrtl(a) sqrt((double) (a))
#else
namespace std { using ::fabsl; using ::atan2l; using ::sqrtl; }
#endif
#endif
#endif
