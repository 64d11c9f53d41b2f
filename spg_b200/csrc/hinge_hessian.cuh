// Full 12 x 12 Hessian of the discrete-bending energy E = K/2 (theta - theta0)^2 (discreteBendingEnergy.cpp:9-30), by
// hand, for the block-sparse assembly of the implicit-Euler family (newton.cu; the reference gets it from
// TinyAD: energy.h:196-237).  The generic path evaluates the energy ten times per hinge with 16-double mixed duals
// (255 registers, spills, 11 ms for the 3.1 M hinges of a 1024^2 cloth); this is one evaluation of ~500 flops.
//
// With e = x1 - x0, L = |e|, a = x2 - x1, b = x3 - x1, N1 = e x a, N2 = b x e, s_k = |N_k|^2 (streams.cuh):
//   theta = atan2(-L N1.b, N1.N2)
//   g2 = -L N1 / s1,  g3 = -L N2 / s2,  g0 = -(phi_a N1 / s1 + phi_b N2 / s2),  phi_a = a.e / L,  phi_b = b.e / L,
//   g1 = -(g0 + g2 + g3)                                                        [theta is translation invariant]
// Second derivatives T(i, j) = d g_i / d x_j:
//   T(2, 2) = -L / s1^2 (N1 t^T + t N1^T),  t = e x N1            T(3, 3) = +L / s2^2 (N2 u^T + u N2^T),  u = e x N2
//   T(0, 0) = sym2(N1, dphi_a / (2 s1) - phi_a (a x N1) / s1^2) + sym2(N2, dphi_b / (2 s2) + phi_b (b x N2) / s2^2),
//             dphi = (a or b) / L - e ((a or b).e) / L^3,  sym2(p, q) = p q^T + q p^T          [the owner blocks of streams.cuh]
//   T(2, 0) = N1 e^T / (L s1) - (L / s1) [a]x + (2 L / s1^2) N1 (N1 x a)^T      [g2 depends on x0 through e only]
//   T(3, 0) = N2 e^T / (L s2) + (L / s2) [b]x - (2 L / s2^2) N2 (N2 x b)^T
//   T(2, 3) = 0                                                                 [g2 does not depend on x3]
// and every other block from "rows sum to zero" and symmetry:
//   T(2, 1) = -T(2, 2) - T(2, 0),  T(3, 1) = -T(3, 3) - T(3, 0),  T(0, 1) = -T(0, 0) - T(2, 0)^T - T(3, 0)^T,
//   T(1, 1) = -T(0, 1)^T - T(2, 1)^T - T(3, 1)^T.
// Element Hessian block (i, j) = K (g_i g_j^T + (theta - theta0) T(i, j)).
#pragma once
#include "energies.cuh"

namespace spgx
{
struct M3 {
    double m[9];  // row-major
};
SPGX_HD M3 m3Outer(const V3d &u, const V3d &v, double s)
{
    return {{s * u.x * v.x, s * u.x * v.y, s * u.x * v.z, s * u.y * v.x, s * u.y * v.y, s * u.y * v.z, s * u.z * v.x,
             s * u.z * v.y, s * u.z * v.z}};
}
SPGX_HD M3 m3Sym2(const V3d &u, const V3d &v, double s)  // s (u v^T + v u^T)
{
    const double xy = s * (u.x * v.y + u.y * v.x), xz = s * (u.x * v.z + u.z * v.x), yz = s * (u.y * v.z + u.z * v.y);
    return {{2.0 * s * u.x * v.x, xy, xz, xy, 2.0 * s * u.y * v.y, yz, xz, yz, 2.0 * s * u.z * v.z}};
}
SPGX_HD M3 m3Skew(const V3d &a, double s)  // s [a]x:  ([a]x d = a x d)
{
    return {{0.0, -s * a.z, s * a.y, s * a.z, 0.0, -s * a.x, -s * a.y, s * a.x, 0.0}};
}
SPGX_HD M3 m3Add(const M3 &a, const M3 &b)
{
    M3 r;
#pragma unroll
    for (int k = 0; k < 9; ++k) r.m[k] = a.m[k] + b.m[k];
    return r;
}
SPGX_HD M3 m3Neg(const M3 &a)
{
    M3 r;
#pragma unroll
    for (int k = 0; k < 9; ++k) r.m[k] = -a.m[k];
    return r;
}
SPGX_HD M3 m3T(const M3 &a)
{
    return {{a.m[0], a.m[3], a.m[6], a.m[1], a.m[4], a.m[7], a.m[2], a.m[5], a.m[8]}};
}

// x: the four stencil positions, p: the record {theta0, K, 1/K}; out: 16 blocks [la][lb][9], row-major 3 x 3 each
SPGX_HD void hingeElementHessian(const V3d *x, const double *p, double *out)
{
    const V3d e = x[1] - x[0], a = x[2] - x[1], b = x[3] - x[1];
    const V3d N1 = cross(e, a), N2 = cross(b, e);
    const double ee = dot(e, e);
    const double L = ::sqrt(ee), iL = 1.0 / L;
    const double i1 = 1.0 / dot(N1, N1), i2 = 1.0 / dot(N2, N2);
    const double theta = ::atan2(-L * dot(N1, b), dot(N1, N2));
    const double K = p[1], d = theta - p[0];
    const double ae = dot(a, e), be = dot(b, e);
    const double phia = ae * iL, phib = be * iL, iL3 = iL * iL * iL;

    V3d g[4];
    g[2] = N1 * (-L * i1);
    g[3] = N2 * (-L * i2);
    g[0] = -(N1 * (phia * i1) + N2 * (phib * i2));
    g[1] = -(g[0] + g[2] + g[3]);

    const M3 T22 = m3Sym2(N1, cross(e, N1), -L * i1 * i1);
    const M3 T33 = m3Sym2(N2, cross(e, N2), L * i2 * i2);
    const V3d dpa = a * iL - e * (ae * iL3), dpb = b * iL - e * (be * iL3);
    const V3d v1 = dpa * (0.5 * i1) - cross(a, N1) * (phia * i1 * i1);
    const V3d v2 = dpb * (0.5 * i2) + cross(b, N2) * (phib * i2 * i2);
    const M3 T00 = m3Add(m3Sym2(N1, v1, 1.0), m3Sym2(N2, v2, 1.0));
    const M3 T20 = m3Add(m3Add(m3Outer(N1, e, iL * i1), m3Skew(a, -L * i1)), m3Outer(N1, cross(N1, a), 2.0 * L * i1 * i1));
    const M3 T30 = m3Add(m3Add(m3Outer(N2, e, iL * i2), m3Skew(b, L * i2)), m3Outer(N2, cross(N2, b), -2.0 * L * i2 * i2));
    const M3 T21 = m3Neg(m3Add(T22, T20));
    const M3 T31 = m3Neg(m3Add(T33, T30));
    const M3 T01 = m3Neg(m3Add(T00, m3Add(m3T(T20), m3T(T30))));
    const M3 T11 = m3Neg(m3Add(m3T(T01), m3Add(m3T(T21), m3T(T31))));
    const M3 Z = {{0, 0, 0, 0, 0, 0, 0, 0, 0}};

    // upper blocks (i <= j) from T(j, i)^T where only the lower one was derived
    const M3 T02 = m3T(T20), T03 = m3T(T30), T12 = m3T(T21), T13 = m3T(T31);
    const M3 *T[4][4] = {{&T00, &T01, &T02, &T03}, {nullptr, &T11, &T12, &T13}, {nullptr, nullptr, &T22, &Z}, {nullptr, nullptr, nullptr, &T33}};
    const double Kd = K * d;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = i; j < 4; ++j) {
            const M3 &t = *T[i][j];
            const double gi[3] = {g[i].x, g[i].y, g[i].z}, gj[3] = {g[j].x, g[j].y, g[j].z};
            double *bij = out + (i * 4 + j) * 9, *bji = out + (j * 4 + i) * 9;
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const double h = K * (gi[r] * gj[c]) + Kd * t.m[3 * r + c];
                    bij[3 * r + c] = h;
                    bji[3 * c + r] = h;
                }
        }
}
}  // namespace spgx
