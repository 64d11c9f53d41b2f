// Energy kinds: the reference's constraint / energy lambdas written once as
// `template <class T> __host__ __device__` bodies over a flat stencil of positions, so that the same
// formula is instantiated with plain doubles, first-order duals (XPBD Jacobian) and second-order
// duals (VBD vertex block).  This is the device-side equivalent of the authoring contract of
// StencilBlockEnergy<S, C> (reference: spg/include/spg/sim/energy/energy.h:75-76, :305-347):
// supply C(x) (and optionally a direct E(x)) and every derivative follows.
//
// Per-stencil parameter record `p` (doubles) is produced on the host by prepare() in
// energy_prepare.h, which restates each energy's addStencil + preparePrecomputations.
#pragma once
#include <cstdint>

#include "linalg.cuh"

namespace spgx
{
// generated from the registry (kinds.def): one enumerator per line of it, K_COUNT = the number of lines
enum Kind : int {
#define SPGX_KIND(NAME, ID, BYTES) NAME = ID,
#include "kinds.def"
#undef SPGX_KIND
    K_COUNT = 0
#define SPGX_KIND(NAME, ID, BYTES) +1
#include "kinds.def"
#undef SPGX_KIND
};
static_assert(K_COUNT <= 32, "kind masks are 32-bit");

template <int KIND>
struct EnergyKind;

// accumulate the symmetric outer product scale * (u v^T + v u^T) into H
SPGX_HD void addSymOuter2(Sym3 &H, const V3d &u, const V3d &v, double scale)
{
    H.xx += scale * (2.0 * u.x * v.x);
    H.yy += scale * (2.0 * u.y * v.y);
    H.zz += scale * (2.0 * u.z * v.z);
    H.xy += scale * (u.x * v.y + u.y * v.x);
    H.xz += scale * (u.x * v.z + u.z * v.x);
    H.yz += scale * (u.y * v.z + u.z * v.y);
}
// accumulate u v^T + v u^T into H
SPGX_HD void addSymOuter2u(Sym3 &H, const V3d &u, const V3d &v)
{
    H.xx += 2.0 * (u.x * v.x);
    H.yy += 2.0 * (u.y * v.y);
    H.zz += 2.0 * (u.z * v.z);
    // two fused multiply-adds straight into the accumulator (a product, an FMA and an add otherwise)
    H.xy = ::fma(u.x, v.y, ::fma(u.y, v.x, H.xy));
    H.xz = ::fma(u.x, v.z, ::fma(u.z, v.x, H.xz));
    H.yz = ::fma(u.y, v.z, ::fma(u.z, v.y, H.yz));
}
// accumulate scale * u u^T into H
SPGX_HD void addOuter(Sym3 &H, const V3d &u, double scale)
{
    const double sx = scale * u.x, sy = scale * u.y, sz = scale * u.z;
    H.xx += sx * u.x;
    H.xy += sx * u.y;
    H.xz += sx * u.z;
    H.yy += sy * u.y;
    H.yz += sy * u.z;
    H.zz += sz * u.z;
}

// deformation gradient columns of a triangle: F = [x1-x0, x2-x0] * Dinv, Dinv row-major 2x2 in p[0..3]
template <class T>
SPGX_HD void membraneF(const V3<T> *x, const double *p, V3<T> &F0, V3<T> &F1)
{
    const V3<T> u = x[1] - x[0];
    const V3<T> v = x[2] - x[0];
    F0 = u * p[0] + v * p[2];
    F1 = u * p[1] + v * p[3];
}
// tetrahedron: F = [x1-x0, x2-x0, x3-x0] * Dinv, Dinv row-major 3x3 in p[0..8]
template <class T>
SPGX_HD void tetF(const V3<T> *x, const double *p, V3<T> &F0, V3<T> &F1, V3<T> &F2)
{
    const V3<T> u = x[1] - x[0];
    const V3<T> v = x[2] - x[0];
    const V3<T> w = x[3] - x[0];
    F0 = u * p[0] + v * p[3] + w * p[6];
    F1 = u * p[1] + v * p[4] + w * p[7];
    F2 = u * p[2] + v * p[5] + w * p[8];
}

// ---- SpringEnergy (springEnergy.cpp:10-30): C = |x - y| - L0.  p = {L0, k, 1/k} ----------------
template <>
struct EnergyKind<K_SPRING> {
    static constexpr int S = 2, NC = 1, NP = 3, NMODEL = 2;
    static constexpr bool kDirectEnergy = false;
    template <class T>
    SPGX_HD static void constraints(const V3<T> *x, const double *p, T *C)
    {
        C[0] = norm(x[0] - x[1]) - p[0];
    }
    SPGX_HD static double K(const double *p, int, int) { return p[1]; }
    SPGX_HD static double alpha(const double *p, int, int) { return p[2]; }
};

// ---- SpringContinuumEnergy (springContinuumEnergy.cpp:8-16): C = |x - y| / L0 - 1.  p = {L0, k L0, 1/(k L0)}
template <>
struct EnergyKind<K_SPRING_CONTINUUM> {
    static constexpr int S = 2, NC = 1, NP = 3, NMODEL = 2;
    static constexpr bool kDirectEnergy = false;
    template <class T>
    SPGX_HD static void constraints(const V3<T> *x, const double *p, T *C)
    {
        C[0] = norm(x[0] - x[1]) / p[0] - 1.0;
    }
    SPGX_HD static double K(const double *p, int, int) { return p[1]; }
    SPGX_HD static double alpha(const double *p, int, int) { return p[2]; }
};

// ---- DiscreteBendingEnergy (discreteBendingEnergy.cpp:9-30): C = theta - theta0.  p = {theta0, K, 1/K}
template <>
struct EnergyKind<K_DISCRETE_BENDING> {
    static constexpr int S = 4, NC = 1, NP = 3, NMODEL = 2;
    static constexpr bool kDirectEnergy = false;
    template <class T>
    SPGX_HD static T theta(const V3<T> *x)
    {
        const V3<T> e0 = x[1] - x[0];
        const V3<T> e3 = x[2] - x[1];
        const V3<T> e4 = x[3] - x[1];
        const V3<T> n1 = normalized(cross(e0, e3));
        const V3<T> n2 = -normalized(cross(e0, e4));
        return datan2(dot(cross(n1, n2), normalized(e0)), dot(n1, n2));
    }
    template <class T>
    SPGX_HD static void constraints(const V3<T> *x, const double *p, T *C)
    {
        C[0] = theta(x) - p[0];
    }
    SPGX_HD static double K(const double *p, int, int) { return p[1]; }
    SPGX_HD static double alpha(const double *p, int, int) { return p[2]; }

    // Hand-derived override of the VBD vertex block (north_star: "hand-derived overrides where ncu shows
    // register or spill pressure"): g += dE/dx_local, H += d2E/dx_local^2 for E = K/2 (theta - theta0)^2, i.e.
    //   g = K d grad(theta),  H = K (grad grad^T + d Hess(theta)),  d = theta - theta0.
    // With e = x1 - x0, (a, b) = (x2, x3) - q (q = x1, or x0 when the owner is x1), N1 = e x a, N2 = b x e:
    //   theta = atan2(-|e| N1.b, N1.N2)          [(N1 x N2).e = -|e|^2 N1.b]
    //   owner x2: grad = -|e| N1/|N1|^2,  Hess = -|e|/|N1|^4 (N1 T^T + T N1^T), T = e x N1   (x3: N2, sign +)
    //   owner x0/x1: grad = -+(phi_a W1 + phi_b W2), W = N/|N|^2, phi_a = a.e/|e|, and
    //     Hess = sym(W1 dphi_a^T) + sym(W2 dphi_b^T) - phi_a (t1 N1^T + N1 t1^T)/|N1|^4 + phi_b (t2 N2^T + N2 t2^T)/|N2|^4
    //     with dphi_a = a/|e| - e (a.e)/|e|^3, t1 = a x N1, t2 = b x N2.
    // Equal to the autodiff block of the constraint above up to rounding (tests/test_gpu_elements.py).
    static constexpr bool kHandDerived = true;
    SPGX_HD static void vertexGradHess(const V3d *x, const double *p, int local, V3d &gAcc, Sym3 &HAcc)
    {
        const V3d e = x[1] - x[0];
        const V3d q = (local == 1) ? x[0] : x[1];
        const V3d a = x[2] - q, b = x[3] - q;
        const V3d N1 = cross(e, a), N2 = cross(b, e);
        const double ee = dot(e, e);
        const double ie = 1.0 / ::sqrt(ee);
        const double le = ee * ie;
        const double d = ::atan2(-le * dot(N1, b), dot(N1, N2)) - p[0];
        const double K = p[1];
        const double Kd = K * d;
        V3d g;  // grad theta
        if (local >= 2) {
            const V3d N = (local == 2) ? N1 : N2;
            const double i = 1.0 / dot(N, N);
            const double c = -le * i;
            g = N * c;
            const V3d T = cross(e, N);
            addSymOuter2(HAcc, N, T, (local == 2 ? Kd : -Kd) * c * i);
        } else {
            const double i1 = 1.0 / dot(N1, N1), i2 = 1.0 / dot(N2, N2);
            const double ae = dot(a, e), be = dot(b, e);
            const double phia = ae * ie, phib = be * ie;
            const double ie3 = ie * ie * ie;
            const V3d W1 = N1 * i1, W2 = N2 * i2;
            const V3d ga = a * ie - e * (ae * ie3), gb = b * ie - e * (be * ie3);
            const V3d sumW = W1 * phia + W2 * phib;
            g = (local == 0) ? -sumW : sumW;
            const V3d t1 = cross(a, N1), t2 = cross(b, N2);
            addSymOuter2(HAcc, W1, ga, 0.5 * Kd);
            addSymOuter2(HAcc, W2, gb, 0.5 * Kd);
            addSymOuter2(HAcc, t1, N1, -Kd * phia * i1 * i1);
            addSymOuter2(HAcc, t2, N2, Kd * phib * i2 * i2);
        }
        gAcc.x += Kd * g.x;
        gAcc.y += Kd * g.y;
        gAcc.z += Kd * g.z;
        addOuter(HAcc, g, K);
    }
};

// ---- QuadraticBendingEnergy (quadraticBendingEnergy.cpp:9-29): C = L x, L = [c0 I, c1 I, c2 I, c3 I].
// p = {c0, c1, c2, c3, K, 1/K}, K_eff = K I3
template <>
struct EnergyKind<K_QUADRATIC_BENDING> {
    static constexpr int S = 4, NC = 3, NP = 6, NMODEL = 2;
    static constexpr bool kDirectEnergy = false;
    template <class T>
    SPGX_HD static void constraints(const V3<T> *x, const double *p, T *C)
    {
        const V3<T> c = x[0] * p[0] + x[1] * p[1] + x[2] * p[2] + x[3] * p[3];
        C[0] = c.x;
        C[1] = c.y;
        C[2] = c.z;
    }
    SPGX_HD static double K(const double *p, int i, int j) { return i == j ? p[4] : 0.0; }
    SPGX_HD static double alpha(const double *p, int i, int j) { return i == j ? p[5] : 0.0; }
    // Hand-derived XPBD projection (optional override of the generic block projection, energy.h:117-153).  C = L x is
    // linear with grad C_a = c_s I per vertex, so A = grad C^T W grad C + alpha~ + 1e-10 I is (sum_s w_s c_s^2 + alpha~ +
    // 1e-10) I -- the off-diagonal sums are sums of exact zeros -- and dlambda = -C / that scalar: ~40 flops instead of a
    // 12-wide dual pass and a 3 x 3 solve (the survey's "AD on a linear map").  lambda: 3 doubles or null.
    static constexpr bool kHandProject = true;
    // The half of the projection that does not read a position -- alpha~ and 1 / (sum_s w_s c_s^2 + alpha~ + 1e-10): the XPBD
    // pass computes it AHEAD of its wait for the previous pass (xpbd.cu), so that the dependent chain behind the position
    // gather is C, one multiplication and the update instead of C, the sum and three divisions.
    SPGX_HD static void projectPrepare(const double *w, const double *p, double dt2, double *pre)
    {
        const double alphaTilde = p[5] / dt2;
        double a = 0.0;
#pragma unroll
        for (int s = 0; s < 4; ++s) a += p[s] * (w[s] * p[s]);
        a += alphaTilde;
        a += 1e-10;
        pre[0] = alphaTilde;
        pre[1] = 1.0 / a;
    }
    SPGX_HD static void projectPrepared(V3d *x, const double *w, const double *p, const double *pre, double *lambda)
    {
        const V3d C = x[0] * p[0] + x[1] * p[1] + x[2] * p[2] + x[3] * p[3];
        const double alphaTilde = pre[0], inva = pre[1];
        V3d rhs(-C.x, -C.y, -C.z);
        if (lambda) {
            rhs.x -= alphaTilde * lambda[0];
            rhs.y -= alphaTilde * lambda[1];
            rhs.z -= alphaTilde * lambda[2];
        }
        const V3d dl(rhs.x * inva, rhs.y * inva, rhs.z * inva);
        if (lambda) {
            lambda[0] += dl.x;
            lambda[1] += dl.y;
            lambda[2] += dl.z;
        }
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            const double ws = w[s] * p[s];
            x[s] = V3d(x[s].x + ws * dl.x, x[s].y + ws * dl.y, x[s].z + ws * dl.z);
        }
    }
    SPGX_HD static void project(V3d *x, const double *w, const double *p, double dt2, double *lambda)
    {
        double pre[2];
        projectPrepare(w, p, dt2, pre);
        projectPrepared(x, w, p, pre, lambda);
    }
};

// ---- MembraneStvkEnergy (membraneStvkEnergy.cpp:9-28): C = (E00, E11, 2 E01), E = (F^T F - I)/2.
// p = {Dinv(4), K00, K01, K22, a00, a01, a22}; K = [[K00,K01,0],[K01,K00,0],[0,0,K22]], same pattern for alpha
template <>
struct EnergyKind<K_MEMBRANE_STVK> {
    static constexpr int S = 3, NC = 3, NP = 10, NMODEL = 2;
    static constexpr bool kDirectEnergy = false;
    template <class T>
    SPGX_HD static void constraints(const V3<T> *x, const double *p, T *C)
    {
        V3<T> F0, F1;
        membraneF(x, p, F0, F1);
        C[0] = 0.5 * (dot(F0, F0) - 1.0);
        C[1] = 0.5 * (dot(F1, F1) - 1.0);
        C[2] = 2.0 * (0.5 * dot(F0, F1));
    }
    SPGX_HD static double pat(const double *q, int i, int j)
    {
        if (i == 2 || j == 2) return (i == j) ? q[2] : 0.0;
        return (i == j) ? q[0] : q[1];
    }
    SPGX_HD static double K(const double *p, int i, int j) { return pat(p + 4, i, j); }
    SPGX_HD static double alpha(const double *p, int i, int j) { return pat(p + 7, i, j); }

    // Hand-derived VBD vertex block. With (ga, gb) the owner's row of the shape-gradient matrix
    // (F = sum_s x_s g_s^T: g_1 = Dinv row 0, g_2 = Dinv row 1, g_0 = -g_1 - g_2), s = K C and
    //   u0 = ga F0, u1 = gb F1, u2 = gb F0 + ga F1           (dC_i = u_i . dx)
    //   grad = sum_i s_i u_i,   Hess = sum_ij K_ij u_i u_j^T + (s0 ga^2 + s1 gb^2 + 2 s2 ga gb) I.
    static constexpr bool kHandDerived = true;
    SPGX_HD static void blockFromF(const V3d &F0, const V3d &F1, double ga, double gb, double K00, double K01, double K22,
                                   V3d &gAcc, Sym3 &HAcc)
    {
        const double C0 = 0.5 * (dot(F0, F0) - 1.0), C1 = 0.5 * (dot(F1, F1) - 1.0), C2 = dot(F0, F1);
        const double s0 = K00 * C0 + K01 * C1, s1 = K01 * C0 + K00 * C1, s2 = K22 * C2;
        const V3d u0 = F0 * ga, u1 = F1 * gb, u2 = F0 * gb + F1 * ga;
        gAcc.x += s0 * u0.x + s1 * u1.x + s2 * u2.x;
        gAcc.y += s0 * u0.y + s1 * u1.y + s2 * u2.y;
        gAcc.z += s0 * u0.z + s1 * u1.z + s2 * u2.z;
        const V3d w0 = u0 * K00 + u1 * K01, w1 = u0 * K01 + u1 * K00, w2 = u2 * K22;
        const double geo = s0 * ga * ga + s1 * gb * gb + 2.0 * s2 * ga * gb;
        HAcc.xx += u0.x * w0.x + u1.x * w1.x + u2.x * w2.x + geo;
        HAcc.yy += u0.y * w0.y + u1.y * w1.y + u2.y * w2.y + geo;
        HAcc.zz += u0.z * w0.z + u1.z * w1.z + u2.z * w2.z + geo;
        HAcc.xy += u0.x * w0.y + u1.x * w1.y + u2.x * w2.y;
        HAcc.xz += u0.x * w0.z + u1.x * w1.z + u2.x * w2.z;
        HAcc.yz += u0.y * w0.z + u1.y * w1.z + u2.y * w2.z;
    }
    SPGX_HD static void vertexGradHess(const V3d *x, const double *p, int local, V3d &gAcc, Sym3 &HAcc)
    {
        V3d F0, F1;
        membraneF(x, p, F0, F1);
        const double ga = (local == 0) ? -(p[0] + p[2]) : (local == 1 ? p[0] : p[2]);
        const double gb = (local == 0) ? -(p[1] + p[3]) : (local == 1 ? p[1] : p[3]);
        blockFromF(F0, F1, ga, gb, p[4], p[5], p[6], gAcc, HAcc);
    }
};

// ---- MembraneChoiEnergy (membraneChoiEnergy.cpp:9-28).  p = {Dinv(4), K0..K3, a0..a3} (diagonal)
template <>
struct EnergyKind<K_MEMBRANE_CHOI> {
    static constexpr int S = 3, NC = 4, NP = 12, NMODEL = 3;
    static constexpr bool kDirectEnergy = false;
    template <class T>
    SPGX_HD static void constraints(const V3<T> *x, const double *p, T *C)
    {
        V3<T> F0, F1;
        membraneF(x, p, F0, F1);
        const double sqrtTwoInv = 1.0 / ::sqrt(2.0);
        C[0] = norm(F0) - 1.0;
        C[1] = norm(F1) - 1.0;
        C[2] = norm((F0 + F1) * sqrtTwoInv) - 1.0;
        C[3] = norm((F0 - F1) * sqrtTwoInv) - 1.0;
    }
    SPGX_HD static double K(const double *p, int i, int j) { return i == j ? p[4 + i] : 0.0; }
    SPGX_HD static double alpha(const double *p, int i, int j) { return i == j ? p[8 + i] : 0.0; }
};

// ---- MembraneBaraffWitkinEnergy (membraneBaraffWitkinEnergy.cpp:9-26).  p = {Dinv(4), K0..K2, a0..a2}
template <>
struct EnergyKind<K_MEMBRANE_BW> {
    static constexpr int S = 3, NC = 3, NP = 10, NMODEL = 3;
    static constexpr bool kDirectEnergy = false;
    template <class T>
    SPGX_HD static void constraints(const V3<T> *x, const double *p, T *C)
    {
        V3<T> F0, F1;
        membraneF(x, p, F0, F1);
        C[0] = norm(F0) - 1.0;
        C[1] = norm(F1) - 1.0;
        C[2] = dot(F0, F1);
    }
    SPGX_HD static double K(const double *p, int i, int j) { return i == j ? p[4 + i] : 0.0; }
    SPGX_HD static double alpha(const double *p, int i, int j) { return i == j ? p[7 + i] : 0.0; }
};

// Green strain of a tet: E = (F^T F - I)/2, returned as its 6 distinct entries
template <class T>
SPGX_HD void greenStrain(const V3<T> &F0, const V3<T> &F1, const V3<T> &F2, T &e00, T &e11, T &e22, T &e01, T &e02,
                         T &e12)
{
    e00 = 0.5 * (dot(F0, F0) - 1.0);
    e11 = 0.5 * (dot(F1, F1) - 1.0);
    e22 = 0.5 * (dot(F2, F2) - 1.0);
    e01 = 0.5 * dot(F0, F1);
    e02 = 0.5 * dot(F0, F2);
    e12 = 0.5 * dot(F1, F2);
}

// ---- StvkEnergy, tetrahedra (stvkEnergy.cpp:9-39 constraint, :41-65 energy).
// p = {Dinv(9), lambda V, mu V, 1/(lambda V), 1/(mu V)}
template <>
struct EnergyKind<K_TET_STVK> {
    static constexpr int S = 4, NC = 2, NP = 13, NMODEL = 2;
    static constexpr bool kDirectEnergy = true;
    template <class T>
    SPGX_HD static void constraints(const V3<T> *x, const double *p, T *C)
    {
        V3<T> F0, F1, F2;
        tetF(x, p, F0, F1, F2);
        T e00, e11, e22, e01, e02, e12;
        greenStrain(F0, F1, F2, e00, e11, e22, e01, e02, e12);
        C[0] = e00 + e11 + e22;
        const T trEE = e00 * e00 + e11 * e11 + e22 * e22 + 2.0 * (e01 * e01 + e02 * e02 + e12 * e12);
        // zero guard of stvkEnergy.cpp:36-38: a vanishing deviatoric term becomes a passive 0
        if (value_of(trEE) == 0.0)
            C[1] = T(0.0);
        else
            C[1] = dsqrt(2.0 * trEE);
    }
    template <class T>
    SPGX_HD static T energy(const V3<T> *x, const double *p)
    {
        V3<T> F0, F1, F2;
        tetF(x, p, F0, F1, F2);
        T e00, e11, e22, e01, e02, e12;
        greenStrain(F0, F1, F2, e00, e11, e22, e01, e02, e12);
        const T tr = e00 + e11 + e22;
        const T trEE = e00 * e00 + e11 * e11 + e22 * e22 + 2.0 * (e01 * e01 + e02 * e02 + e12 * e12);
        return (0.5 * p[9]) * tr * tr + p[10] * trEE;
    }
    SPGX_HD static double K(const double *p, int i, int j) { return i == j ? p[9 + i] : 0.0; }
    SPGX_HD static double alpha(const double *p, int i, int j) { return i == j ? p[11 + i] : 0.0; }
};

// ---- StableNeoHookeanEnergy (stableNeoHookeanEnergy.cpp:9-31 constraint, :33-61 energy).
// p = {Dinv(9), lambda V, mu V, 1/(lambda V), 1/(mu V), alpha = 1 + mu/lambda}
template <>
struct EnergyKind<K_TET_SNH> {
    static constexpr int S = 4, NC = 2, NP = 14, NMODEL = 2;
    static constexpr bool kDirectEnergy = true;
    template <class T>
    SPGX_HD static void constraints(const V3<T> *x, const double *p, T *C)
    {
        V3<T> F0, F1, F2;
        tetF(x, p, F0, F1, F2);
        C[0] = det3(F0, F1, F2) - p[13];
        C[1] = dsqrt(dot(F0, F0) + dot(F1, F1) + dot(F2, F2));
    }
    template <class T>
    SPGX_HD static T energy(const V3<T> *x, const double *p)
    {
        V3<T> F0, F1, F2;
        tetF(x, p, F0, F1, F2);
        const T hyd = det3(F0, F1, F2) - p[13];
        const T dev = dot(F0, F0) + dot(F1, F1) + dot(F2, F2) - 3.0;
        return (0.5 * p[9]) * hyd * hyd + (0.5 * p[10]) * dev;
    }
    SPGX_HD static double K(const double *p, int i, int j) { return i == j ? p[9 + i] : 0.0; }
    SPGX_HD static double alpha(const double *p, int i, int j) { return i == j ? p[11 + i] : 0.0; }
};

// ---- SpringAnchorEnergy (springAnchorEnergy.cpp:8-26).  p = {ax, ay, az, k, 1/k}
// XPBD: C = |x - a| with the zero guard (:16-18); VBD: E = k/2 |x - a|^2 with the *model* stiffness (:25)
template <>
struct EnergyKind<K_ANCHOR> {
    static constexpr int S = 1, NC = 1, NP = 5, NMODEL = 4;
    static constexpr bool kDirectEnergy = true;
    template <class T>
    SPGX_HD static void constraints(const V3<T> *x, const double *p, T *C)
    {
        const V3<T> d(x[0].x - p[0], x[0].y - p[1], x[0].z - p[2]);
        const T n2 = squaredNorm(d);
        if (value_of(n2) == 0.0)
            C[0] = T(0.0);
        else
            C[0] = dsqrt(n2);
    }
    template <class T>
    SPGX_HD static T energy(const V3<T> *x, const double *p)
    {
        const V3<T> d(x[0].x - p[0], x[0].y - p[1], x[0].z - p[2]);
        return (0.5 * p[3]) * squaredNorm(d);
    }
    SPGX_HD static double K(const double *p, int, int) { return p[3]; }
    SPGX_HD static double alpha(const double *p, int, int) { return p[4]; }
};

// ---- BaraffWitkinBendingEnergy (baraffWitkinBendingEnergy.cpp:9-37): the dihedral angle with the three norms taken
// as PASSIVE values (".val", :25-31: no derivative flows through them).  p = {theta0, k, 1/k}; the effective stiffness
// is the model stiffness (no area factor, :68-69), theta0 is recomputed from the rest pose (:71-74)
template <>
struct EnergyKind<K_BENDING_BW> {
    static constexpr int S = 4, NC = 1, NP = 3, NMODEL = 2;
    static constexpr bool kDirectEnergy = false;
    template <class T>
    SPGX_HD static void constraints(const V3<T> *x, const double *p, T *C)
    {
        const V3<T> e0 = x[1] - x[0];
        const V3<T> e3 = x[2] - x[1];
        const V3<T> e4 = x[3] - x[1];
        const double e0NormInv = 1.0 / value_of(norm(e0));
        const V3<T> n1 = cross(e0, e3);
        const double n1NormInv = 1.0 / value_of(norm(n1));
        const V3<T> n2 = -cross(e0, e4);
        const double n2NormInv = 1.0 / value_of(norm(n2));
        const V3<T> a = n1 * n1NormInv, b = n2 * n2NormInv;
        C[0] = datan2(dot(cross(a, b), e0 * e0NormInv), dot(a, b)) - p[0];
    }
    SPGX_HD static double K(const double *p, int, int) { return p[1]; }
    SPGX_HD static double alpha(const double *p, int, int) { return p[2]; }
};

// ---- SpringSquaredContinuumEnergy (springSquaredContinuumEnergy.cpp:8-17): C = |x - y|^2 / L0^2 - 1.
// p = {L0, k L0, 1/(k L0)}
template <>
struct EnergyKind<K_SPRING_SQUARED_CONTINUUM> {
    static constexpr int S = 2, NC = 1, NP = 3, NMODEL = 2;
    static constexpr bool kDirectEnergy = false;
    template <class T>
    SPGX_HD static void constraints(const V3<T> *x, const double *p, T *C)
    {
        C[0] = squaredNorm(x[0] - x[1]) / (p[0] * p[0]) - 1.0;
    }
    SPGX_HD static double K(const double *p, int, int) { return p[1]; }
    SPGX_HD static double alpha(const double *p, int, int) { return p[2]; }
};

// ---- StableSquaredNeoHookeanEnergy (stableSquaredNeoHookeanEnergy.cpp:9-31): C = (det F - 1, tr(F^T F) - 3), energy
// through the default 1/2 C^T K C.  p = {Dinv(9), lambda V, mu V, 1/(lambda V), 1/(mu V)} with the Lame lambda + 1e-20
template <>
struct EnergyKind<K_TET_SSNH> {
    static constexpr int S = 4, NC = 2, NP = 13, NMODEL = 2;
    static constexpr bool kDirectEnergy = false;
    template <class T>
    SPGX_HD static void constraints(const V3<T> *x, const double *p, T *C)
    {
        V3<T> F0, F1, F2;
        tetF(x, p, F0, F1, F2);
        C[0] = det3(F0, F1, F2) - 1.0;
        C[1] = dot(F0, F0) + dot(F1, F1) + dot(F2, F2) - 3.0;
    }
    SPGX_HD static double K(const double *p, int i, int j) { return i == j ? p[9 + i] : 0.0; }
    SPGX_HD static double alpha(const double *p, int i, int j) { return i == j ? p[11 + i] : 0.0; }
};

// ---- StaticLoadEnergy (staticLoadEnergy.cpp:15-20, staticLoadEnergy.h:11-19): E = -k (d . x), d normalised at
// addStencil.  Energy only: its constraint form throws in the reference (:8-13), so XPBD rejects it (api_step.cu).
// p = {dx, dy, dz, k, 1/k}
template <>
struct EnergyKind<K_STATIC_LOAD> {
    static constexpr int S = 1, NC = 1, NP = 5, NMODEL = 4;
    static constexpr bool kDirectEnergy = true;
    template <class T>
    SPGX_HD static void constraints(const V3<T> *, const double *, T *C)
    {
        C[0] = T(0.0);  // never reached: spgx_xpbd_step refuses groups that hold this kind
    }
    template <class T>
    SPGX_HD static T energy(const V3<T> *x, const double *p)
    {
        return p[3] * -(p[0] * x[0].x + p[1] * x[0].y + p[2] * x[0].z);
    }
    SPGX_HD static double K(const double *p, int, int) { return p[3]; }
    SPGX_HD static double alpha(const double *p, int, int) { return p[4]; }
};

// E(x) of one stencil for any scalar type: the direct energy when the kind has one, otherwise
// 1/2 C^T K_eff C (energy.h:336-347)
template <int KIND, class T>
SPGX_HD T stencilEnergy(const V3<T> *x, const double *p)
{
    using EK = EnergyKind<KIND>;
    if constexpr (EK::kDirectEnergy) {
        return EK::energy(x, p);
    } else {
        T C[EK::NC];
        EK::constraints(x, p, C);
        T e = T(0.0);
#pragma unroll
        for (int i = 0; i < EK::NC; ++i) {
            T row = C[i] * EK::K(p, i, i);
#pragma unroll
            for (int j = 0; j < EK::NC; ++j)
                if (j != i && EK::K(p, i, j) != 0.0) row = row + C[j] * EK::K(p, i, j);
            e = e + C[i] * row;
        }
        return 0.5 * e;
    }
}

// does the kind carry a hand-derived XPBD projection?
template <class EK, class = void>
struct HasHandProject {
    static constexpr bool value = false;
};
template <class EK>
struct HasHandProject<EK, decltype(void(EK::kHandProject))> {
    static constexpr bool value = EK::kHandProject;
};

// does the kind carry a hand-derived vertexGradHess override?
template <class EK, class = void>
struct HasHandDerived {
    static constexpr bool value = false;
};
template <class EK>
struct HasHandDerived<EK, decltype(void(EK::kHandDerived))> {
    static constexpr bool value = EK::kHandDerived;
};

}  // namespace spgx
// kinds defined outside the library sources (tests/native/extra_kind.cuh shows the recipe; INTEGRATION.md "Adding an energy kind")
#ifdef SPGX_EXTRA_KINDS_IMPL
#include SPGX_EXTRA_KINDS_IMPL
#endif
namespace spgx
{
// per-kind sizes, generated from the registry and the EnergyKind<> traits (-1: unknown kind)
SPGX_HD int kindStencilSize(int kind)
{
    switch (kind) {
#define SPGX_KIND(NAME, ID, BYTES) case NAME: return EnergyKind<NAME>::S;
#include "kinds.def"
#undef SPGX_KIND
    default: return -1;
    }
}
SPGX_HD int kindNumParams(int kind)
{
    switch (kind) {
#define SPGX_KIND(NAME, ID, BYTES) case NAME: return EnergyKind<NAME>::NP;
#include "kinds.def"
#undef SPGX_KIND
    default: return -1;
    }
}
SPGX_HD int kindNumModelParams(int kind)
{
    switch (kind) {
#define SPGX_KIND(NAME, ID, BYTES) case NAME: return EnergyKind<NAME>::NMODEL;
#include "kinds.def"
#undef SPGX_KIND
    default: return -1;
    }
}
SPGX_HD int kindNumConstraints(int kind)
{
    switch (kind) {
#define SPGX_KIND(NAME, ID, BYTES) case NAME: return EnergyKind<NAME>::NC;
#include "kinds.def"
#undef SPGX_KIND
    default: return -1;
    }
}
// bytes of the element record the gather model charges per incidence (SURVEY.md 8d / DESIGN.md)
inline double modelRecordBytes(int kind)
{
    switch (kind) {
#define SPGX_KIND(NAME, ID, BYTES) case NAME: return BYTES;
#include "kinds.def"
#undef SPGX_KIND
    default: return 0;
    }
}
}  // namespace spgx
