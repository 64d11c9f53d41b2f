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
#include "linalg.cuh"

namespace spgx
{
enum Kind : int {
    K_SPRING = 0,
    K_SPRING_CONTINUUM = 1,
    K_DISCRETE_BENDING = 2,
    K_QUADRATIC_BENDING = 3,
    K_MEMBRANE_STVK = 4,
    K_MEMBRANE_CHOI = 5,
    K_MEMBRANE_BW = 6,
    K_TET_STVK = 7,
    K_TET_SNH = 8,
    K_ANCHOR = 9,
    K_COUNT = 10
};

template <int KIND>
struct EnergyKind;

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

SPGX_HD int kindStencilSize(int kind)
{
    switch (kind) {
    case K_SPRING:
    case K_SPRING_CONTINUUM: return 2;
    case K_MEMBRANE_STVK:
    case K_MEMBRANE_CHOI:
    case K_MEMBRANE_BW: return 3;
    case K_ANCHOR: return 1;
    default: return 4;
    }
}
}  // namespace spgx
