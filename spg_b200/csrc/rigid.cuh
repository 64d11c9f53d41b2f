// Rigid bodies (reference: spg/include/spg/sim/simObject/rigidBodyGroup.h:14-111,
// spg/source/sim/simObject/rigidBodyGroup.cpp:6-243 and the four rigid-body energies spg/source/sim/energy/rigidBody*.cpp).
//
// A body carries 6 DOFs: the centre of mass and an orientation kept BOTH as a rotation matrix (the state that is composed
// incrementally, rigidBodyGroup.cpp:64-78, :96-116) and as the axis-angle vector derived from it.  The energies are written
// the way the reference writes them: a constraint over (com, dtheta) per body with the orientation linearised around the
// current rotation, globalPoint = com + (I + skew(dtheta)) R local, dtheta = 0 (rigidBodySpringEnergy.cpp:16-33) -- one
// template body per kind, instantiated with doubles, first-order duals over the 6 S stencil DOFs (XPBD) and second-order duals
// seeded on ONE body's 6 DOFs (the 6 x 6 VBD block the reference cuts out of a full 6 S x 6 S autodiff pass,
// energy.h:240-258).
#pragma once
#include <cmath>
#include <cstdint>

#include "dual.cuh"
#include "linalg.cuh"

namespace spgx
{
enum RbKind : int {
    RB_SPRING = 0,                 // rigidBodySpringEnergy.cpp: C = |p0 - p1| - L0
    RB_SPRING_SQUARED = 1,         // rigidBodySpringSquaredEnergy.cpp: C = |p0 - p1|^2 - L0^2
    RB_SPRING_ANCHOR = 2,          // rigidBodySpringAnchorEnergy.cpp: C = |p - a| (XPBD), E = k/2 |p - a|^2 through the point map (VBD)
    RB_ORIENTATION_ALIGNMENT = 3,  // rigidBodyOrientationAlignmentEnergy.cpp: C = sum_c |R0 col c x R1 col c|^2
    RB_KIND_COUNT = 4
};

// 3 x 3 matrices are row-major double[9]
SPGX_HD void mat3Mul(const double *A, const double *B, double *C)
{
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}
SPGX_HD void mat3MulT(const double *A, const double *B, double *C)  // C = A B^T
{
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) C[3 * i + j] = A[3 * i] * B[3 * j] + A[3 * i + 1] * B[3 * j + 1] + A[3 * i + 2] * B[3 * j + 2];
}
SPGX_HD V3d mat3Vec(const double *A, const V3d &v)
{
    return V3d(A[0] * v.x + A[1] * v.y + A[2] * v.z, A[3] * v.x + A[4] * v.y + A[5] * v.z, A[6] * v.x + A[7] * v.y + A[8] * v.z);
}
// Eigen::AngleAxis(angle, axis).toRotationMatrix() (Eigen/src/Geometry/AngleAxis.h:219-242)
SPGX_HD void angleAxisToMatrix(double angle, const V3d &axis, double *R)
{
    const double s = ::sin(angle), c = ::cos(angle);
    const V3d sinAxis(s * axis.x, s * axis.y, s * axis.z);
    const V3d cos1Axis((1.0 - c) * axis.x, (1.0 - c) * axis.y, (1.0 - c) * axis.z);
    double tmp = cos1Axis.x * axis.y;
    R[1] = tmp - sinAxis.z;
    R[3] = tmp + sinAxis.z;
    tmp = cos1Axis.x * axis.z;
    R[2] = tmp + sinAxis.y;
    R[6] = tmp - sinAxis.y;
    tmp = cos1Axis.y * axis.z;
    R[5] = tmp - sinAxis.x;
    R[7] = tmp + sinAxis.x;
    R[0] = cos1Axis.x * axis.x + c;
    R[4] = cos1Axis.y * axis.y + c;
    R[8] = cos1Axis.z * axis.z + c;
}
// axisAngleToRotMatrix (rigidBodyGroup.cpp:6-10)
SPGX_HD void thetaToMatrix(const V3d &theta, double *R)
{
    const double angle = ::sqrt(dot(theta, theta));
    if (angle != 0) {
        angleAxisToMatrix(angle, V3d(theta.x / angle, theta.y / angle, theta.z / angle), R);
    } else {
        R[0] = R[4] = R[8] = 1.0;
        R[1] = R[2] = R[3] = R[5] = R[6] = R[7] = 0.0;
    }
}
// rotMatrixToAxisAngle (rigidBodyGroup.cpp:12-16): Eigen::AngleAxis(R) goes through a quaternion
// (Eigen/src/Geometry/Quaternion.h:862-894 for the matrix, AngleAxis.h:171-189 for the angle and the axis)
SPGX_HD void matrixToAngleAxis(const double *m, double &angle, V3d &axis)
{
    double q[3], w;
    double t = m[0] + m[4] + m[8];
    if (t > 0) {
        t = ::sqrt(t + 1.0);
        w = 0.5 * t;
        t = 0.5 / t;
        q[0] = (m[7] - m[5]) * t;
        q[1] = (m[2] - m[6]) * t;
        q[2] = (m[3] - m[1]) * t;
    } else {
        int i = 0;
        if (m[4] > m[0]) i = 1;
        if (m[8] > m[4 * i]) i = 2;
        const int j = (i + 1) % 3, k = (j + 1) % 3;
        t = ::sqrt(m[4 * i] - m[4 * j] - m[4 * k] + 1.0);
        q[i] = 0.5 * t;
        t = 0.5 / t;
        w = (m[3 * k + j] - m[3 * j + k]) * t;
        q[j] = (m[3 * j + i] + m[3 * i + j]) * t;
        q[k] = (m[3 * k + i] + m[3 * i + k]) * t;
    }
    double n = ::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
    // (below machine epsilon Eigen recomputes the norm with stableNorm: the same value up to rounding for these magnitudes)
    if (n != 0) {
        angle = 2.0 * ::atan2(n, ::fabs(w));
        if (w < 0) n = -n;
        axis = V3d(q[0] / n, q[1] / n, q[2] / n);
    } else {
        angle = 0;
        axis = V3d(1.0, 0.0, 0.0);
    }
}
SPGX_HD V3d matrixToTheta(const double *R)
{
    double angle;
    V3d axis;
    matrixToAngleAxis(R, angle, axis);
    return V3d(axis.x * angle, axis.y * angle, axis.z * angle);
}
// updateInertia (rigidBodyGroup.cpp:207-211): I = R Ilocal R^T (and the same for the inverse)
SPGX_HD void rotateInertia(const double *R, const double *local, double *out)
{
    double tmp[9];
    mat3Mul(R, local, tmp);
    mat3MulT(tmp, R, out);
}
// incremental composition R <- Rot(|d|, d / |d|) R (rigidBodyGroup.cpp:69-74, :88-93, :107-112)
SPGX_HD void composeRotation(const V3d &d, double *R)
{
    const double n = ::sqrt(dot(d, d));
    if (n != 0) {
        double dR[9], out[9];
        angleAxisToMatrix(n, V3d(d.x / n, d.y / n, d.z / n), dR);
        mat3Mul(dR, R, out);
#pragma unroll
        for (int i = 0; i < 9; ++i) R[i] = out[i];
    }
}
// computeIntegratedOmega (rigidBodyGroup.cpp:132-136): the rotation from thetaOld to R, as angle * axis * invdt
SPGX_HD V3d integratedOmega(const double *R, const V3d &thetaOld, double invdt)
{
    double Rold[9], dR[9];
    thetaToMatrix(thetaOld, Rold);
    mat3MulT(R, Rold, dR);
    double angle;
    V3d axis;
    matrixToAngleAxis(dR, angle, axis);
    return V3d(angle * axis.x * invdt, angle * axis.y * invdt, angle * axis.z * invdt);
}

// com + (I + skew(dtheta)) (R local): the body point under the linearised orientation increment
template <class T>
SPGX_HD V3<T> rbPoint(const V3<T> &com, const V3<T> &dtheta, const double *R, const V3d &local)
{
    const V3d r = mat3Vec(R, local);
    // (I + skew(t)) r = r + t x r
    return V3<T>(com.x + (dtheta.y * r.z - dtheta.z * r.y + r.x), com.y + (dtheta.z * r.x - dtheta.x * r.z + r.y),
                 com.z + (dtheta.x * r.y - dtheta.y * r.x + r.z));
}

template <int KIND>
struct RbEnergyKind;

// p = {local0(3), local1(3), L0, k, 1/k}; model = {local0(3), local1(3), restLength, stiffness}
template <>
struct RbEnergyKind<RB_SPRING> {
    static constexpr int S = 2, NP = 9, NMODEL = 8;
    template <class T>
    SPGX_HD static T constraint(const V3<T> *com, const V3<T> *dth, const double *const *R, const double *p)
    {
        const V3<T> a = rbPoint(com[0], dth[0], R[0], V3d(p[0], p[1], p[2]));
        const V3<T> b = rbPoint(com[1], dth[1], R[1], V3d(p[3], p[4], p[5]));
        return norm(a - b) - p[6];
    }
    SPGX_HD static double K(const double *p) { return p[7]; }
    SPGX_HD static double alpha(const double *p) { return p[8]; }
};
// p = {local0(3), local1(3), L0^2, k, 1/k}
template <>
struct RbEnergyKind<RB_SPRING_SQUARED> {
    static constexpr int S = 2, NP = 9, NMODEL = 8;
    template <class T>
    SPGX_HD static T constraint(const V3<T> *com, const V3<T> *dth, const double *const *R, const double *p)
    {
        const V3<T> a = rbPoint(com[0], dth[0], R[0], V3d(p[0], p[1], p[2]));
        const V3<T> b = rbPoint(com[1], dth[1], R[1], V3d(p[3], p[4], p[5]));
        const V3<T> d = a - b;
        return dot(d, d) - p[6];
    }
    SPGX_HD static double K(const double *p) { return p[7]; }
    SPGX_HD static double alpha(const double *p) { return p[8]; }
};
// p = {local(3), anchor(3), k, 1/k}; model = {local(3), anchor(3), stiffness}
template <>
struct RbEnergyKind<RB_SPRING_ANCHOR> {
    static constexpr int S = 1, NP = 8, NMODEL = 7;
    template <class T>
    SPGX_HD static T constraint(const V3<T> *com, const V3<T> *dth, const double *const *R, const double *p)
    {
        const V3<T> a = rbPoint(com[0], dth[0], R[0], V3d(p[0], p[1], p[2]));
        const V3<T> d(a.x - p[3], a.y - p[4], a.z - p[5]);
        const T c = norm(d);
        if (value_of(c) == 0) return T(0.0);  // rigidBodySpringAnchorEnergy.cpp:25-30: the gradient is undefined at 0
        return c;
    }
    SPGX_HD static double K(const double *p) { return p[6]; }
    SPGX_HD static double alpha(const double *p) { return p[7]; }
};
// p = {k, 1/k}; model = {stiffness}
template <>
struct RbEnergyKind<RB_ORIENTATION_ALIGNMENT> {
    static constexpr int S = 2, NP = 2, NMODEL = 1;
    template <class T>
    SPGX_HD static T constraint(const V3<T> *, const V3<T> *dth, const double *const *R, const double *)
    {
        // rotMatrix_b = (I + skew(dtheta_b)) R_b; C = sum over the three columns of |col_c(0) x col_c(1)|^2
        T c = T(0.0);
#pragma unroll
        for (int col = 0; col < 3; ++col) {
            V3<T> u[2];
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                const V3d r(R[b][col], R[b][3 + col], R[b][6 + col]);
                u[b] = V3<T>(dth[b].y * r.z - dth[b].z * r.y + r.x, dth[b].z * r.x - dth[b].x * r.z + r.y,
                             dth[b].x * r.y - dth[b].y * r.x + r.z);
            }
            const V3<T> x = cross(u[0], u[1]);
            c = c + dot(x, x);
        }
        return c;
    }
    SPGX_HD static double K(const double *p) { return p[0]; }
    SPGX_HD static double alpha(const double *p) { return p[1]; }
};

inline int rbKindStencilSize(int kind)
{
    return kind == RB_SPRING_ANCHOR ? 1 : (kind >= 0 && kind < RB_KIND_COUNT ? 2 : -1);
}
inline int rbKindNumParams(int kind)
{
    static const int np[RB_KIND_COUNT] = {9, 9, 8, 2};
    return kind >= 0 && kind < RB_KIND_COUNT ? np[kind] : -1;
}
inline int rbKindNumModelParams(int kind)
{
    static const int nm[RB_KIND_COUNT] = {8, 8, 7, 1};
    return kind >= 0 && kind < RB_KIND_COUNT ? nm[kind] : -1;
}
}  // namespace spgx
