// adk_device.cuh — device-side scalar predicates and the O(1) points-in-triangle query.
//
// FP64 restatement of the reference's scalar layer.  Every product / sum is an explicit round-to-nearest
// intrinsic so that nvcc can never contract a*b-c*d into an FMA: sign decisions must be the ones the
// reference takes in binary64.  Reference: src/common/CustomMath.cpp, Utils.cpp:251-293.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace adk
{
constexpr double kEps = 1e-10;       // CustomMath.h:46-47  (applied to the HALF cross product)
constexpr double kAngleTol = 1e-6;   // Utils.cpp:301
constexpr double kTwoPi = 6.283185307179586476925286766559;
constexpr double kPi = 3.141592653589793238462643383279;

// CustomMath.cpp:32-36
__device__ __forceinline__ double signed_area(double ax, double ay, double bx, double by, double cx, double cy)
{
    const double l = __dmul_rn(__dsub_rn(bx, ax), __dsub_rn(cy, ay));
    const double r = __dmul_rn(__dsub_rn(cx, ax), __dsub_rn(by, ay));
    return __dmul_rn(0.5, __dsub_rn(l, r));
}
__device__ __forceinline__ bool near_zero(double v) { return fabs(v) < kEps; }   // CustomMath.cpp:22-25
// CustomMath.cpp:43-54 : -1 right turn, 0 collinear, +1 left turn
__device__ __forceinline__ int orientation(double ax, double ay, double bx, double by, double cx, double cy)
{
    const double s = signed_area(ax, ay, bx, by, cx, cy);
    if (near_zero(s)) return 0;
    return s > 0.0 ? 1 : -1;
}
// CustomMath.cpp:15-20
__device__ __forceinline__ double dist2(double ax, double ay, double bx, double by)
{
    const double dx = __dsub_rn(bx, ax), dy = __dsub_rn(by, ay);
    return __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
}
// CustomMath.cpp:56-65
__device__ __forceinline__ double angle_of(double rx, double ry, double px, double py)
{
    const double a = atan2(__dsub_rn(py, ry), __dsub_rn(px, rx));
    return a < 0.0 ? __dadd_rn(a, kTwoPi) : a;
}
// CustomMath.cpp:89-93
__device__ __forceinline__ double pseudo_dist(double sx, double sy, double ex, double ey, double px, double py)
{
    const double l = __dmul_rn(__dsub_rn(px, sx), __dadd_rn(-ey, sy));
    const double r = __dmul_rn(__dsub_rn(py, sy), __dsub_rn(ex, sx));
    return fabs(__dadd_rn(l, r));
}
// (x, then y) order of CustomMath.h:21-23
__device__ __forceinline__ bool xy_less(double ax, double ay, double bx, double by)
{
    return ax < bx || (ax == bx && ay < by);
}

// The three count tables of Utils.h:28-33, n*n row-major int32 (+ n).
struct Tables
{
    const int32_t* below;
    const int32_t* coll;
    const int32_t* rb;
    uint32_t n;
    __device__ __forceinline__ int B(int a, int b) const { return __ldg(below + (size_t) a * n + b); }
    __device__ __forceinline__ int C(int a, int b) const { return __ldg(coll + (size_t) a * n + b); }
    __device__ __forceinline__ int R(int a) const { return __ldg(rb + a); }
};

// Utils.cpp:251-293 — number of input points strictly inside triangle (a,b,c), by inclusion-exclusion.
__device__ __forceinline__ int points_in_triangle(const Tables& T, int a, double ax, double ay, int b, double bx, double by,
                                                  int c, double cx, double cy)
{
    if (near_zero(signed_area(ax, ay, bx, by, cx, cy))) return 0;
    int l = a, m = b, r = c;
    double lx = ax, ly = ay, mx = bx, my = by, rx = cx, ry = cy;
#define ADK_SWAP(p, px, py, q, qx, qy) { int t = p; p = q; q = t; double tx = px; px = qx; qx = tx; double ty = py; py = qy; qy = ty; }
    if (xy_less(mx, my, lx, ly)) ADK_SWAP(l, lx, ly, m, mx, my)
    if (xy_less(rx, ry, mx, my)) ADK_SWAP(m, mx, my, r, rx, ry)
    if (xy_less(mx, my, lx, ly)) ADK_SWAP(l, lx, ly, m, mx, my)
#undef ADK_SWAP
    if (lx == mx) return T.B(m, r) - T.B(l, r) - T.C(l, r);
    if (mx == rx) return T.B(l, r) - T.B(l, m) - T.C(l, m);
    if (orientation(rx, ry, lx, ly, mx, my) == 1)
        return T.B(l, r) - T.B(l, m) - T.B(m, r) - T.C(l, m) - T.C(m, r) - T.R(m) - 1;
    return T.B(l, m) + T.B(m, r) + T.R(m) - T.B(l, r) - T.C(l, r);
}

// order-preserving image of a non-negative double (areas) for 64-bit atomicMin
__device__ __forceinline__ unsigned long long area_bits(double a) { return (unsigned long long) __double_as_longlong(a); }
__device__ __forceinline__ double bits_area(unsigned long long b) { return __longlong_as_double((long long) b); }
constexpr unsigned long long kInfBits = 0x7ff0000000000000ull;

}  // namespace adk
