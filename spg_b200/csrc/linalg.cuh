// Tiny fixed-size linear algebra over double or Dual scalars.
//
// Restates the handful of Eigen fixed-size primitives the reference's hot path touches
// (external/eigen/Eigen/src/Core/Dot.h:92-102 `normalized` with its z > 0 guard,
//  .../LU/Determinant.h:21-24,47-51 3x3 determinant, .../LU/InverseImpl.h:131-142 cofactor inverse,
//  cross / dot / norm) so that the energy bodies read like the reference lambdas.
#pragma once
#include "dual.cuh"

namespace spgx
{
template <class T>
struct V3 {
    T x, y, z;
    SPGX_HD V3() {}
    SPGX_HD V3(const T &a, const T &b, const T &c) : x(a), y(b), z(c) {}
    SPGX_HD T &operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
    SPGX_HD const T &operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
using V3d = V3<double>;

template <class T>
SPGX_HD V3<T> operator+(const V3<T> &a, const V3<T> &b)
{
    return {a.x + b.x, a.y + b.y, a.z + b.z};
}
template <class T>
SPGX_HD V3<T> operator-(const V3<T> &a, const V3<T> &b)
{
    return {a.x - b.x, a.y - b.y, a.z - b.z};
}
template <class T>
SPGX_HD V3<T> operator-(const V3<T> &a)
{
    return {-a.x, -a.y, -a.z};
}
template <class T, class S>
SPGX_HD V3<T> operator*(const V3<T> &a, const S &s)
{
    return {a.x * s, a.y * s, a.z * s};
}
template <class T>
SPGX_HD T dot(const V3<T> &a, const V3<T> &b)
{
    return a.x * b.x + a.y * b.y + a.z * b.z;
}
template <class T>
SPGX_HD V3<T> cross(const V3<T> &a, const V3<T> &b)
{
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
template <class T>
SPGX_HD T squaredNorm(const V3<T> &a)
{
    return dot(a, a);
}
template <class T>
SPGX_HD T norm(const V3<T> &a)
{
    return dsqrt(squaredNorm(a));
}
// Eigen's normalized(): x / sqrt(z) when z > 0, x unchanged otherwise (Core/Dot.h:92-102)
template <class T>
SPGX_HD V3<T> normalized(const V3<T> &a)
{
    const T z = squaredNorm(a);
    if (value_of(z) > 0.0) {
        const T r = drsqrt(z);
        return {a.x * r, a.y * r, a.z * r};
    }
    return a;
}

// ---- plain-double 3x3 helpers used by the VBD vertex solve (vbd.cpp:139-145) ------------------
// symmetric 3x3 stored as {xx, xy, xz, yy, yz, zz}
struct Sym3 {
    double xx, xy, xz, yy, yz, zz;
};
SPGX_HD double det(const Sym3 &m)
{
    // LU/Determinant.h:21-24: sum over cyclic (a,b,c) of m(0,a) * (m(1,b) m(2,c) - m(1,c) m(2,b))
    return m.xx * (m.yy * m.zz - m.yz * m.yz) + m.xy * (m.yz * m.xz - m.xy * m.zz) + m.xz * (m.xy * m.yz - m.yy * m.xz);
}
// adj(m) * f / det  ==  m^-1 f for a symmetric m (cofactor inverse, LU/InverseImpl.h:131-142)
SPGX_HD V3d solveSym3(const Sym3 &m, double d, const V3d &f)
{
    const double c00 = m.yy * m.zz - m.yz * m.yz;
    const double c01 = m.xz * m.yz - m.xy * m.zz;
    const double c02 = m.xy * m.yz - m.xz * m.yy;
    const double c11 = m.xx * m.zz - m.xz * m.xz;
    const double c12 = m.xy * m.xz - m.xx * m.yz;
    const double c22 = m.xx * m.yy - m.xy * m.xy;
    const double id = 1.0 / d;
    return {(c00 * f.x + c01 * f.y + c02 * f.z) * id, (c01 * f.x + c11 * f.y + c12 * f.z) * id,
            (c02 * f.x + c12 * f.y + c22 * f.z) * id};
}

// 3x3 determinant over a generic scalar, column vectors a, b, c
template <class T>
SPGX_HD T det3(const V3<T> &a, const V3<T> &b, const V3<T> &c)
{
    return dot(a, cross(b, c));
}

// In-place solve of the small SPD system A y = b (n <= 4) by Gaussian elimination without
// pivoting; A is row-major n x n.  Used for the C x C block of the XPBD projection (energy.h:139-145).
template <int N>
SPGX_HD void solveSmall(double (&A)[N][N], double (&b)[N])
{
#pragma unroll
    for (int k = 0; k < N; ++k) {
        const double ip = 1.0 / A[k][k];
#pragma unroll
        for (int r = k + 1; r < N; ++r) {
            const double f = A[r][k] * ip;
#pragma unroll
            for (int c = k + 1; c < N; ++c) A[r][c] -= f * A[k][c];
            b[r] -= f * b[k];
        }
    }
#pragma unroll
    for (int k = N - 1; k >= 0; --k) {
        double s = b[k];
#pragma unroll
        for (int c = k + 1; c < N; ++c) s -= A[k][c] * b[c];
        b[k] = s / A[k][k];
    }
}
}  // namespace spgx
