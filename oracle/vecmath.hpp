// oracle/vecmath.hpp — TEST INFRASTRUCTURE (CPU oracle), not product code.
//
// FP64 vector helpers restating the Apple `simd` operations the reference's
// render path uses (Models/RTContext.swift:13-21: Vec3 = SIMD3<Double>,
// Mat4 = simd_double4x4).  The simd headers are not in /root/reference, so the
// definitions below are the standard ones with a fixed evaluation order and NO
// fused multiply-add (compile with -ffp-contract=off):
//   dot       = (x*x' + y*y') + z*z'
//   cross     = (y*z' - z*y', z*x' - x*z', x*y' - y*x')
//   normalize = v * (1 / sqrt(dot(v,v)))          (simd_precise_normalize form)
//   length    = sqrt(dot(v,v))
//   M4 * v4   = ((c0*x + c1*y) + c2*z) + c3*w
// Parity against Apple's exact rounding is unpinned (ULP level).
#pragma once
#include <cmath>
#include <cstdint>
#include <limits>

namespace orc {

static constexpr double kInf = std::numeric_limits<double>::infinity();

struct Vec3 {
    double x, y, z;
    Vec3() : x(0), y(0), z(0) {}
    Vec3(double a, double b, double c) : x(a), y(b), z(c) {}
    explicit Vec3(double a) : x(a), y(a), z(a) {}
    double operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};

inline Vec3 operator+(Vec3 a, Vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline Vec3 operator-(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vec3 operator-(Vec3 a) { return {-a.x, -a.y, -a.z}; }
inline Vec3 operator*(Vec3 a, Vec3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline Vec3 operator*(Vec3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
inline Vec3 operator*(double s, Vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline Vec3 operator/(Vec3 a, double s) { return {a.x / s, a.y / s, a.z / s}; }
inline Vec3 operator/(Vec3 a, Vec3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline Vec3 operator/(double s, Vec3 a) { return {s / a.x, s / a.y, s / a.z}; }
inline Vec3& operator+=(Vec3& a, Vec3 b) { a = a + b; return a; }
inline Vec3& operator-=(Vec3& a, Vec3 b) { a = a - b; return a; }
inline Vec3& operator*=(Vec3& a, Vec3 b) { a = a * b; return a; }
inline bool operator==(Vec3 a, Vec3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
inline bool operator!=(Vec3 a, Vec3 b) { return !(a == b); }

inline double dmin(double a, double b) { return a < b ? a : b; }
inline double dmax(double a, double b) { return a > b ? a : b; }
inline Vec3 vmin(Vec3 a, Vec3 b) { return {dmin(a.x, b.x), dmin(a.y, b.y), dmin(a.z, b.z)}; }
inline Vec3 vmax(Vec3 a, Vec3 b) { return {dmax(a.x, b.x), dmax(a.y, b.y), dmax(a.z, b.z)}; }
inline Vec3 vabs(Vec3 a) { return {std::fabs(a.x), std::fabs(a.y), std::fabs(a.z)}; }

inline double dot(Vec3 a, Vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline Vec3 cross(Vec3 a, Vec3 b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
inline double length_squared(Vec3 a) { return dot(a, a); }
inline double length(Vec3 a) { return std::sqrt(dot(a, a)); }
inline Vec3 normalize(Vec3 a) { return a * (1.0 / std::sqrt(dot(a, a))); }

// Column-major 3x3 / 4x4 (simd_double3x3 / simd_double4x4): c[col] is a column.
struct Mat3 {
    Vec3 c[3];
};
struct Vec4 {
    double x, y, z, w;
};
struct Mat4 {
    double m[16];  // m[col*4 + row]
    double at(int r, int col) const { return m[col * 4 + r]; }
};

inline Mat4 mat4Identity() {
    Mat4 M{};
    for (int i = 0; i < 16; i++) M.m[i] = (i % 5 == 0) ? 1.0 : 0.0;
    return M;
}

inline Vec4 mul(const Mat4& M, Vec4 v) {
    Vec4 r;
    r.x = ((M.m[0] * v.x + M.m[4] * v.y) + M.m[8] * v.z) + M.m[12] * v.w;
    r.y = ((M.m[1] * v.x + M.m[5] * v.y) + M.m[9] * v.z) + M.m[13] * v.w;
    r.z = ((M.m[2] * v.x + M.m[6] * v.y) + M.m[10] * v.z) + M.m[14] * v.w;
    r.w = ((M.m[3] * v.x + M.m[7] * v.y) + M.m[11] * v.z) + M.m[15] * v.w;
    return r;
}
inline Vec3 mul(const Mat3& M, Vec3 v) {
    return {(M.c[0].x * v.x + M.c[1].x * v.y) + M.c[2].x * v.z,
            (M.c[0].y * v.x + M.c[1].y * v.y) + M.c[2].y * v.z,
            (M.c[0].z * v.x + M.c[1].z * v.y) + M.c[2].z * v.z};
}
inline Mat3 upper3(const Mat4& M) {
    Mat3 r;
    r.c[0] = {M.m[0], M.m[1], M.m[2]};
    r.c[1] = {M.m[4], M.m[5], M.m[6]};
    r.c[2] = {M.m[8], M.m[9], M.m[10]};
    return r;
}
inline double determinant(const Mat3& A) { return dot(A.c[0], cross(A.c[1], A.c[2])); }
inline Mat3 transpose(const Mat3& A) {
    Mat3 r;
    r.c[0] = {A.c[0].x, A.c[1].x, A.c[2].x};
    r.c[1] = {A.c[0].y, A.c[1].y, A.c[2].y};
    r.c[2] = {A.c[0].z, A.c[1].z, A.c[2].z};
    return r;
}
// Inverse by adjugate / determinant (simd_inverse's algorithm is not in the
// tree; parity unpinned at ULP level).  inverse(A) rows are cross products of
// A's columns divided by det.
inline Mat3 inverse(const Mat3& A) {
    Vec3 r0 = cross(A.c[1], A.c[2]), r1 = cross(A.c[2], A.c[0]), r2 = cross(A.c[0], A.c[1]);
    double inv = 1.0 / dot(A.c[0], r0);
    r0 = r0 * inv; r1 = r1 * inv; r2 = r2 * inv;
    Mat3 rows;  // rows of the inverse as columns -> transpose to get the inverse
    rows.c[0] = r0; rows.c[1] = r1; rows.c[2] = r2;
    return transpose(rows);
}
// 4x4 inverse: Gauss-Jordan with partial pivoting (general matrices;
// stands in for simd_double4x4.inverse, RTContext.swift:451).
inline Mat4 inverse(const Mat4& M) {
    double a[4][8];
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) { a[r][c] = M.at(r, c); a[r][c + 4] = (r == c) ? 1.0 : 0.0; }
    for (int col = 0; col < 4; col++) {
        int piv = col;
        for (int r = col + 1; r < 4; r++)
            if (std::fabs(a[r][col]) > std::fabs(a[piv][col])) piv = r;
        if (piv != col)
            for (int c = 0; c < 8; c++) { double t = a[col][c]; a[col][c] = a[piv][c]; a[piv][c] = t; }
        double d = 1.0 / a[col][col];
        for (int c = 0; c < 8; c++) a[col][c] *= d;
        for (int r = 0; r < 4; r++) {
            if (r == col) continue;
            double f = a[r][col];
            if (f == 0.0) continue;
            for (int c = 0; c < 8; c++) a[r][c] -= f * a[col][c];
        }
    }
    Mat4 R{};
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) R.m[c * 4 + r] = a[r][c + 4];
    return R;
}

}  // namespace orc
