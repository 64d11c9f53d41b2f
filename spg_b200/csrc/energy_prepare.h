// Host-side restatement of every energy's addStencil + preparePrecomputations: turns the model
// parameters the reference's addStencil takes into the per-stencil record the kernels read
// (layout documented per kind in energies.cuh).  Validation rules and messages follow the reference.
#pragma once
#include <cmath>
#include <stdexcept>
#include <string>

#include "energies.cuh"

namespace spgx
{
namespace prep
{
struct H3 {
    double x, y, z;
};
inline H3 sub(const H3 &a, const H3 &b)
{
    return {a.x - b.x, a.y - b.y, a.z - b.z};
}
inline H3 crossH(const H3 &a, const H3 &b)
{
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
inline double dotH(const H3 &a, const H3 &b)
{
    return a.x * b.x + a.y * b.y + a.z * b.z;
}
inline double normH(const H3 &a)
{
    return std::sqrt(dotH(a, a));
}
inline H3 normalizedH(const H3 &a)
{
    const double z = dotH(a, a);
    if (z > 0) {
        const double s = std::sqrt(z);
        return {a.x / s, a.y / s, a.z / s};
    }
    return a;
}
inline H3 at(const double *x0, int v)
{
    return {x0[3 * v], x0[3 * v + 1], x0[3 * v + 2]};
}

// 3x3 inverse by cofactors, row-major m -> row-major inv (Eigen LU/InverseImpl.h:131-142 order of operations)
inline void inverse3(const double m[9], double inv[9], double &det)
{
    auto M = [&](int r, int c) { return m[3 * r + c]; };
    auto cof = [&](int i, int j) {
        const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
        return M(i1, j1) * M(i2, j2) - M(i1, j2) * M(i2, j1);
    };
    const double c00 = cof(0, 0), c10 = cof(1, 0), c20 = cof(2, 0);
    det = c00 * M(0, 0) + c10 * M(1, 0) + c20 * M(2, 0);
    const double invdet = 1.0 / det;
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) inv[3 * r + c] = cof(c, r) * invdet;
}
inline double det3rm(const double m[9])
{
    auto M = [&](int r, int c) { return m[3 * r + c]; };
    // LU/Determinant.h:21-24,47-51
    auto h = [&](int a, int b, int c) { return M(0, a) * (M(1, b) * M(2, c) - M(1, c) * M(2, b)); };
    return h(0, 1, 2) + h(1, 2, 0) + h(2, 0, 1);
}

// rest-state 2x2 reference matrix of a membrane triangle -> Dinv (row-major) and area
// (membraneStvkEnergy.cpp:45-71; the Choi / Baraff-Witkin versions are identical up to the message)
inline void membraneRest(const double *x0, const int *st, const char *name, double Dinv[4], double &area)
{
    const H3 a = at(x0, st[0]), b = at(x0, st[1]), c = at(x0, st[2]);
    if (a.z != 0.0 || b.z != 0.0 || c.z != 0.0)
        throw std::runtime_error(std::string("Rest positions of triangles for membrane ") + name +
                                 " energy must have coordinate z=0");
    const double ux = b.x - a.x, uy = b.y - a.y, vx = c.x - a.x, vy = c.y - a.y;
    // matrix [[ux, vx], [uy, vy]]; Eigen 2x2 inverse: invdet * [[d, -b], [-c, a]]
    const double det = ux * vy - uy * vx;
    const double invdet = 1.0 / det;
    Dinv[0] = vy * invdet;
    Dinv[1] = -vx * invdet;
    Dinv[2] = -uy * invdet;
    Dinv[3] = ux * invdet;
    area = (ux * vy - vx * uy) * 0.5;
    if (area <= 0)
        throw std::runtime_error(std::string("Rest area of membrane ") + name + " triangle is not positive");
}

inline void tetRest(const double *x0, const int *st, double Dinv[9], double &volume, const char *name = "SNH")
{
    const H3 a = at(x0, st[0]);
    const H3 u = sub(at(x0, st[1]), a), v = sub(at(x0, st[2]), a), w = sub(at(x0, st[3]), a);
    const double m[9] = {u.x, v.x, w.x, u.y, v.y, w.y, u.z, v.z, w.z};
    volume = 1. / 6. * det3rm(m);
    if (volume <= 0) throw std::runtime_error(std::string(name) + " tetrahedron rest volume is not positive");
    double det;
    inverse3(m, Dinv, det);
}

inline double cotTheta(const H3 &a, const H3 &b)
{
    return dotH(a, b) / normH(crossH(a, b));
}
}  // namespace prep

// EnergyKind<K>::prepare(st, model, x0, p), when the specialisation has one
template <class EK, class = void>
struct HasPrepare {
    static constexpr bool value = false;
};
template <class EK>
struct HasPrepare<EK, decltype(void(&EK::prepare))> {
    static constexpr bool value = true;
};
template <int KIND>
inline bool prepareKindByTrait(const int *st, const double *model, const double *x0, double *p)
{
    if constexpr (HasPrepare<EnergyKind<KIND>>::value) {
        EnergyKind<KIND>::prepare(st, model, x0, p);
        return true;
    } else {
        return false;
    }
}
inline bool prepareByTrait(int kind, const int *st, const double *model, const double *x0, double *p)
{
    switch (kind) {
#define SPGX_KIND(NAME, ID, BYTES)                                                                                     \
    case NAME: return prepareKindByTrait<NAME>(st, model, x0, p);
#include "kinds.def"
#undef SPGX_KIND
    default: return false;
    }
}

// model -> record for stencil `st` (vertex ids) with rest positions x0 (3 doubles per vertex)
inline void prepareStencil(int kind, const int *st, const double *model, const double *x0, double *p)
{
    using namespace prep;
    switch (kind) {
    case K_SPRING: {  // springEnergy.cpp:33-46
        if (model[0] <= 0) throw std::runtime_error("Zero/negative rest length invalid");
        p[0] = model[0];
        p[1] = model[1];
        p[2] = 1.0 / model[1];
        break;
    }
    case K_SPRING_CONTINUUM: {  // springContinuumEnergy.cpp:19-33
        if (model[0] <= 0) throw std::runtime_error("Zero/negative rest length invalid");
        p[0] = model[0];
        p[1] = model[1] * model[0];
        p[2] = 1.0 / p[1];
        break;
    }
    case K_DISCRETE_BENDING: {  // discreteBendingEnergy.cpp:33-72
        if (model[0] < -3.12 || model[0] > 3.12) throw std::runtime_error("Invalid restTheta");
        const H3 x0p = at(x0, st[0]), x1p = at(x0, st[1]), x2p = at(x0, st[2]), x3p = at(x0, st[3]);
        const H3 e0 = sub(x1p, x0p), e3 = sub(x2p, x1p), e4 = sub(x3p, x1p);
        const double constantFactor = 2. * 3. * dotH(e0, e0) / (0.5 * (normH(crossH(e0, e3)) + normH(crossH(e0, e4))));
        const double modelCompliance = 1.0 / model[1];
        const double effCompliance = modelCompliance / constantFactor;
        const H3 n1 = normalizedH(crossH(e0, e3));
        const H3 c2 = normalizedH(crossH(e0, e4));
        const H3 n2 = {-c2.x, -c2.y, -c2.z};
        p[0] = std::atan2(dotH(crossH(n1, n2), normalizedH(e0)), dotH(n1, n2));
        p[1] = 1.0 / effCompliance;
        p[2] = effCompliance;
        break;
    }
    case K_QUADRATIC_BENDING: {  // quadraticBendingEnergy.cpp:32-93
        if (model[0] != 0) throw std::runtime_error("Rest theta != 0 not supported for quadratic bending");
        const H3 x0p = at(x0, st[0]), x1p = at(x0, st[1]), x2p = at(x0, st[2]), x3p = at(x0, st[3]);
        const H3 e0 = sub(x1p, x0p), e1 = sub(x2p, x0p), e2 = sub(x3p, x0p), e3 = sub(x2p, x1p), e4 = sub(x3p, x1p);
        const H3 me0 = {-e0.x, -e0.y, -e0.z};
        const double c01 = cotTheta(e0, e1), c02 = cotTheta(e0, e2), c03 = cotTheta(me0, e3), c04 = cotTheta(me0, e4);
        p[0] = c03 + c04;
        p[1] = c01 + c02;
        p[2] = -c01 - c03;
        p[3] = -c02 - c04;
        p[4] = model[1] * (3 / (0.5 * (normH(crossH(e0, e1)) + normH(crossH(e0, e2)))));
        p[5] = 1.0 / p[4];
        break;
    }
    case K_MEMBRANE_STVK: {  // membraneStvkEnergy.cpp:31-71
        const double young = model[0], poisson = model[1];
        const double yc = young / (1 - poisson * poisson);
        const double C[9] = {yc, yc * poisson, 0, yc * poisson, yc, 0, 0, 0, yc * (1 - poisson) / 2};
        double Ci[9], det;
        inverse3(C, Ci, det);
        double area;
        membraneRest(x0, st, "StVK", p, area);
        p[4] = C[0] * area;
        p[5] = C[1] * area;
        p[6] = C[8] * area;
        p[7] = Ci[0] / area;
        p[8] = Ci[1] / area;
        p[9] = Ci[8] / area;
        break;
    }
    case K_MEMBRANE_CHOI: {  // membraneChoiEnergy.cpp:31-71
        double area;
        membraneRest(x0, st, "Choi", p, area);
        const double k[4] = {model[0], model[1], model[2], model[2]};
        for (int i = 0; i < 4; ++i) {
            p[4 + i] = k[i] * area;
            p[8 + i] = (1.0 / k[i]) / area;
        }
        break;
    }
    case K_MEMBRANE_BW: {  // membraneBaraffWitkinEnergy.cpp:29-70
        double area;
        membraneRest(x0, st, "BaraffWitkin", p, area);
        for (int i = 0; i < 3; ++i) {
            p[4 + i] = model[i] * area;
            p[7 + i] = (1.0 / model[i]) / area;
        }
        break;
    }
    case K_TET_STVK: {  // stvkEnergy.cpp:68-110
        const double young = model[0], poisson = model[1];
        const double mu = young / (2 * (1 + poisson));
        const double lambda = young * poisson / ((1 + poisson) * (1 - 2 * poisson)) + 1e-20;
        double volume;
        tetRest(x0, st, p, volume);
        // Eigen 2x2 inverse of diag(lambda, mu): entries mu/det, lambda/det
        const double invdet = 1.0 / (lambda * mu);
        p[9] = lambda * volume;
        p[10] = mu * volume;
        p[11] = (mu * invdet) / volume;
        p[12] = (lambda * invdet) / volume;
        break;
    }
    case K_TET_SNH: {  // stableNeoHookeanEnergy.cpp:64-108
        const double young = model[0], poisson = model[1];
        const double mu = young / (2 * (1 + poisson));
        const double lambda = mu / (1 - 2 * poisson);
        double volume;
        tetRest(x0, st, p, volume);
        const double invdet = 1.0 / (lambda * mu);
        p[9] = lambda * volume;
        p[10] = mu * volume;
        p[11] = (mu * invdet) / volume;
        p[12] = (lambda * invdet) / volume;
        p[13] = 1 + mu / lambda;
        break;
    }
    case K_ANCHOR: {  // springAnchorEnergy.cpp:29-39
        p[0] = model[0];
        p[1] = model[1];
        p[2] = model[2];
        p[3] = model[3];
        p[4] = 1.0 / model[3];
        break;
    }
    case K_BENDING_BW: {  // baraffWitkinBendingEnergy.cpp:40-77: stiffness as given, rest angle from the rest pose
        if (model[0] < -3.12 || model[0] > 3.12) throw std::runtime_error("Invalid restTheta");
        const H3 x0p = at(x0, st[0]), x1p = at(x0, st[1]), x2p = at(x0, st[2]), x3p = at(x0, st[3]);
        const H3 e0 = sub(x1p, x0p), e3 = sub(x2p, x1p), e4 = sub(x3p, x1p);
        const H3 n1 = normalizedH(crossH(e0, e3));
        const H3 c2 = normalizedH(crossH(e0, e4));
        const H3 n2 = {-c2.x, -c2.y, -c2.z};
        p[0] = std::atan2(dotH(crossH(n1, n2), normalizedH(e0)), dotH(n1, n2));
        p[1] = model[1];
        p[2] = 1.0 / model[1];
        break;
    }
    case K_SPRING_SQUARED_CONTINUUM: {  // springSquaredContinuumEnergy.cpp:20-34
        if (model[0] <= 0) throw std::runtime_error("Zero/negative rest length invalid");
        p[0] = model[0];
        p[1] = model[1] * model[0];
        p[2] = 1.0 / p[1];
        break;
    }
    case K_TET_SSNH: {  // stableSquaredNeoHookeanEnergy.cpp:34-78 (Lame parameters, lambda + 1e-20)
        const double young = model[0], poisson = model[1];
        const double mu = young / (2 * (1 + poisson));
        const double lambda = young * poisson / ((1 + poisson) * (1 - 2 * poisson)) + 1e-20;
        double volume;
        tetRest(x0, st, p, volume, "SSNH");
        const double invdet = 1.0 / (lambda * mu);
        p[9] = lambda * volume;
        p[10] = mu * volume;
        p[11] = (mu * invdet) / volume;
        p[12] = (lambda * invdet) / volume;
        break;
    }
    case K_STATIC_LOAD: {  // staticLoadEnergy.h:11-19: the direction is normalised (Eigen normalized(): z > 0 guard)
        const H3 d = normalizedH({model[0], model[1], model[2]});
        p[0] = d.x;
        p[1] = d.y;
        p[2] = d.z;
        p[3] = model[3];
        p[4] = 1.0 / model[3];
        break;
    }
    default:
        // a kind that brings its own prepare() (registered through SPGX_EXTRA_KINDS_DEF / _IMPL)
        if (!prepareByTrait(kind, st, model, x0, p)) throw std::runtime_error("unknown energy kind");
    }
}

}  // namespace spgx
