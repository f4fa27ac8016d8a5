// rt_math.cuh — FP64 vector helpers for the device code.
//
// Same definitions and evaluation order as the reference's Apple-simd calls
// (see oracle/vecmath.hpp for the restated forms): no FMA contraction — this
// translation unit MUST be compiled with -fmad=false so that the primitive
// tests and ray generation round exactly like the reference's x86-64 Swift
// (SURVEY.md §9.6).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rtg {

struct V3 {
    double x, y, z;
};

__device__ __forceinline__ V3 mk(double x, double y, double z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 ld3(const double* p) { return mk(p[0], p[1], p[2]); }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
__device__ __forceinline__ V3 operator*(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ V3 operator*(V3 a, double s) { return mk(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 operator*(double s, V3 a) { return mk(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ V3 operator/(V3 a, double s);
__device__ __forceinline__ double dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
    return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
// FP64 sqrt and 1/x expand to ~70 SASS instructions; `normalize` has some twenty call sites on the render path.  Out of
// line (RT_NOINLINE_MATH, one copy each) the executed code of the kernels with recursion shrinks by a third, which is what
// their instruction-cache stalls are made of.  Same operations, same results.
#ifndef RT_NOINLINE_MATH
#define RT_NOINLINE_MATH 0
#endif
#if RT_NOINLINE_MATH && !defined(RT_EMU)
#define RT_MATH_FN static __device__ __noinline__
#else
#define RT_MATH_FN __device__ __forceinline__
#endif
// FP64 a / b: ~25 inlined instructions plus a slow-path call at every site (59 sites in the fast kernel with the ray queue,
// 136 in the general kernel with area lights); behind RT_NOINLINE_MATH one copy serves them all.  IEEE division either way.
RT_MATH_FN double ddiv(double a, double b) { return a / b; }
RT_MATH_FN double dsqrt(double a) { return sqrt(a); }
RT_MATH_FN double dpow(double a, double b) { return pow(a, b); }
RT_MATH_FN double rsqrtExact(double s) { return 1.0 / sqrt(s); }
__device__ __forceinline__ V3 operator/(V3 a, double s) { return mk(ddiv(a.x, s), ddiv(a.y, s), ddiv(a.z, s)); }
__device__ __forceinline__ V3 normalize(V3 a) { return a * rsqrtExact(dot(a, a)); }
RT_MATH_FN double length(V3 a) { return sqrt(dot(a, a)); }
// Swift.max/min on Doubles as used by the reference (a > b ? a : b), NaN-propagation irrelevant here
__device__ __forceinline__ double dmax(double a, double b) { return a > b ? a : b; }
__device__ __forceinline__ double dmin(double a, double b) { return a < b ? a : b; }
__device__ __forceinline__ bool finite3(V3 a) { return isfinite(a.x) && isfinite(a.y) && isfinite(a.z); }

// 3x4 affine (rows 0..2 of a column-major 4x4), m[c*3 + r]:
// simd_mul(double4x4, double4) = ((c0*x + c1*y) + c2*z) + c3*w
__device__ __forceinline__ V3 xfmPoint(const double* m, V3 p) {
    return mk(((m[0] * p.x + m[3] * p.y) + m[6] * p.z) + m[9] * 1.0,
              ((m[1] * p.x + m[4] * p.y) + m[7] * p.z) + m[10] * 1.0,
              ((m[2] * p.x + m[5] * p.y) + m[8] * p.z) + m[11] * 1.0);
}
__device__ __forceinline__ V3 xfmVector(const double* m, V3 d) {
    return mk(((m[0] * d.x + m[3] * d.y) + m[6] * d.z) + m[9] * 0.0,
              ((m[1] * d.x + m[4] * d.y) + m[7] * d.z) + m[10] * 0.0,
              ((m[2] * d.x + m[5] * d.y) + m[8] * d.z) + m[11] * 0.0);
}
__device__ __forceinline__ V3 mul3(const double* m, V3 v) {  // column-major 3x3
    return mk((m[0] * v.x + m[3] * v.y) + m[6] * v.z, (m[1] * v.x + m[4] * v.y) + m[7] * v.z,
              (m[2] * v.x + m[5] * v.y) + m[8] * v.z);
}

// PCG32 exactly as Object+Extension.swift:556-589 (non-standard seeding)
struct PCG32 {
    uint64_t state, inc;
    __device__ __forceinline__ uint32_t next() {
        uint64_t old = state;
        state = old * 6364136223846793005ULL + inc;
        uint32_t xorshifted = (uint32_t)(((old >> 18) ^ old) >> 27);
        uint32_t rot = (uint32_t)(old >> 59);
        return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31u));
    }
    __device__ __forceinline__ void seed(uint64_t s) {
        state = 0;
        inc = (s << 1) | 1;
        next();
        state += 0x9E3779B97F4A7C15ULL;
        next();
    }
    __device__ __forceinline__ double nextFloat() { return (double)next() * 2.3283064365386963e-10; }
    // (A^k, 1 + A + ... + A^(k-1)) mod 2^64: with them k draws are skipped by `state = A^k state + G inc`
    __device__ __forceinline__ static void skipFactors(uint64_t k, uint64_t& ak, uint64_t& gk) {
        uint64_t curMult = 6364136223846793005ULL, curPlus = 1, accMult = 1, accPlus = 0;
        while (k > 0) {
            if (k & 1) { accMult *= curMult; accPlus = accPlus * curMult + curPlus; }
            curPlus = (curMult + 1) * curPlus;
            curMult *= curMult;
            k >>= 1;
        }
        ak = accMult; gk = accPlus;
    }
    __device__ __forceinline__ void skip(uint64_t ak, uint64_t gk) { state = ak * state + gk * inc; }
    // skip `delta` draws in O(log delta): the LCG state after k steps is A^k*s + (A^k-1)/(A-1)*inc
    __device__ __forceinline__ void advance(uint64_t delta) {
        uint64_t curMult = 6364136223846793005ULL, curPlus = inc, accMult = 1, accPlus = 0;
        while (delta > 0) {
            if (delta & 1) { accMult *= curMult; accPlus = accPlus * curMult + curPlus; }
            curPlus = (curMult + 1) * curPlus;
            curMult *= curMult;
            delta >>= 1;
        }
        state = accMult * state + accPlus;
    }
};

}  // namespace rtg
