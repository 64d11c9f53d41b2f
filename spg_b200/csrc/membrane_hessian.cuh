// Full 9 x 9 Hessian of the StVK membrane energy E = 1/2 C^T K C, C = (E00, E11, 2 E01) (membraneStvkEnergy.cpp:9-28), by
// hand, for the block-sparse assembly of the implicit-Euler family (newton.cu; the reference gets it from TinyAD:
// energy.h:196-237).  Generalises the owner block of EnergyKind<K_MEMBRANE_STVK>::blockFromF (energies.cuh) to a pair of
// vertices.  With g_s = (ga_s, gb_s) the rows of the shape-gradient matrix (F = sum_s x_s g_s^T: g_1 = Dinv row 0,
// g_2 = Dinv row 1, g_0 = -g_1 - g_2), s = K C and, per vertex, u0 = ga F0, u1 = gb F1, u2 = gb F0 + ga F1 (dC_i = u_i . dx):
//   block (a, b) = sum_ij K_ij u_i^a (u_j^b)^T + (s0 ga_a ga_b + s1 gb_a gb_b + s2 (ga_a gb_b + gb_a ga_b)) I.
#pragma once
#include "energies.cuh"

namespace spgx
{
// x: the three stencil positions, p: the record {Dinv(4), K00, K01, K22, ...}; out: 9 blocks [la][lb][9], row-major 3 x 3
SPGX_HD void membraneStvkElementHessian(const V3d *x, const double *p, double *out)
{
    const V3d du = x[1] - x[0], dv = x[2] - x[0];
    const V3d F0 = du * p[0] + dv * p[2], F1 = du * p[1] + dv * p[3];
    const double K00 = p[4], K01 = p[5], K22 = p[6];
    const double C0 = 0.5 * (dot(F0, F0) - 1.0), C1 = 0.5 * (dot(F1, F1) - 1.0), C2 = dot(F0, F1);
    const double s0 = K00 * C0 + K01 * C1, s1 = K01 * C0 + K00 * C1, s2 = K22 * C2;
    const double ga[3] = {-(p[0] + p[2]), p[0], p[2]}, gb[3] = {-(p[1] + p[3]), p[1], p[3]};
    V3d u0[3], u1[3], u2[3], w0[3], w1[3], w2[3];
#pragma unroll
    for (int s = 0; s < 3; ++s) {
        u0[s] = F0 * ga[s];
        u1[s] = F1 * gb[s];
        u2[s] = F0 * gb[s] + F1 * ga[s];
        w0[s] = u0[s] * K00 + u1[s] * K01;  // (K u)_0 of vertex s, used as the right factor
        w1[s] = u0[s] * K01 + u1[s] * K00;
        w2[s] = u2[s] * K22;
    }
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = a; b < 3; ++b) {
            const double geo = s0 * ga[a] * ga[b] + s1 * gb[a] * gb[b] + s2 * (ga[a] * gb[b] + gb[a] * ga[b]);
            const double ua0[3] = {u0[a].x, u0[a].y, u0[a].z}, ua1[3] = {u1[a].x, u1[a].y, u1[a].z}, ua2[3] = {u2[a].x, u2[a].y, u2[a].z};
            const double wb0[3] = {w0[b].x, w0[b].y, w0[b].z}, wb1[3] = {w1[b].x, w1[b].y, w1[b].z}, wb2[3] = {w2[b].x, w2[b].y, w2[b].z};
            double *bab = out + (a * 3 + b) * 9, *bba = out + (b * 3 + a) * 9;
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const double h = ua0[r] * wb0[c] + ua1[r] * wb1[c] + ua2[r] * wb2[c] + (r == c ? geo : 0.0);
                    bab[3 * r + c] = h;
                    bba[3 * c + r] = h;
                }
        }
}
}  // namespace spgx
