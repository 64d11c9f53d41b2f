// Owner-centric incidence streams of the VBD colour pass.
//
// The reference evaluates, for every (vertex, incident stencil) pair, the full stencil gradient and Hessian by
// autodiff and keeps one 3-segment / 3x3 block (spg/include/spg/sim/energy/energy.h:240-258).  Here every such
// pair ("incidence") of a streamed energy kind becomes one small record in a per-warp, fully coalesced stream,
// written once on the host at finalize time:
//   * the stencil is CANONICALISED around the owning vertex, so that the device code has no `local`-dependent
//     selects or branches: discrete bending splits into two stream types (owner on the hinge edge / owner opposite
//     to it) and uses the symmetries theta(x1,x0,x2,x3) = theta(x0,x1,x3,x2) = -theta(x0,x1,x2,x3) of the hinge
//     angle of discreteBendingEnergy.cpp:9-30 (the sign moves into theta0);
//   * the other stencil vertices are referenced by their index in the owner's UNIQUE-NEIGHBOUR list (8 bits each),
//     whose positions the kernel gathers once per vertex into shared memory (a cloth vertex has 8-12 unique
//     neighbours but 36-72 neighbour references);
//   * the per-element constants the block needs ride along as NQ doubles.
// Each evaluator adds dE/dx_owner to g and d2E/dx_owner^2 to H; they are the hand-derived overrides north_star asks
// for ("hand-derived overrides where ncu shows register or spill pressure") and are checked against the generic
// dual-number evaluation of the same energy (tests/test_host_elements.py, tests/test_gpu_elements.py).
#pragma once
#include "energies.cuh"

namespace spgx
{
enum StreamType : int {
    ST_BEND_EDGE = 0,      // discrete bending, owner is an end of the hinge edge
    ST_BEND_OPP = 1,       // discrete bending, owner is opposite to the hinge edge
    ST_MEMBRANE_STVK = 2,  // StVK membrane triangle
    ST_TET_SNH = 3,        // stable Neo-Hookean tetrahedron
    ST_TET_STVK = 4,       // StVK tetrahedron
    ST_WEDGE = 5,          // one triangle of the owner's fan: its StVK membrane + the two discrete-bending hinges that share its normal
    ST_COUNT = 6
};

// 1 / x and 1 / sqrt(x) to double precision from the hardware seed (MUFU.RCP64H / RSQ64H) and two Newton steps,
// without the special-case slow path of the IEEE division (denominators here are squared lengths of mesh edges
// and triangle normals: finite, positive, far from the denormal range)
SPGX_HD double fastRcp(double x)
{
#if defined(__CUDA_ARCH__)
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double e = fma(-x, r, 1.0);
    r = fma(r, e, r);
    e = fma(-x, r, 1.0);
    r = fma(r, e, r);
    return r;
#else
    return 1.0 / x;
#endif
}
SPGX_HD double fastRsqrt(double x)
{
#if defined(__CUDA_ARCH__)
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double hx = 0.5 * x;
    double e = fma(-hx * y, y, 0.5);  // (1 - x y^2) / 2
    y = fma(y, e, y);
    e = fma(-hx * y, y, 0.5);
    y = fma(y, e, y);
    return y;
#else
    return 1.0 / ::sqrt(x);
#endif
}

// atan2(y, x) for finite arguments that are not both zero, to ~2 ulp, without the special-case branches, the IEEE
// division and the 20-term polynomial of the library routine (about 120 issued instructions per call in the colour
// pass, profiles/r1b_*): with mn = min(|x|, |y|), mx = max(|x|, |y|) the ratio t = mn / mx in [0, 1] is folded to
// |t'| <= tan(pi/8) by atan(t) = pi/4 + atan((mn - mx) / (mn + mx)) using the SAME single division, and atan(t') =
// t' + t' z q(z), z = t'^2, with a degree-11 q (interpolated at Chebyshev nodes in 60-digit arithmetic; truncation
// error 5e-20).  The polynomial is split into even and odd halves so that two FMA chains run interleaved.
// On the device the polynomial coefficients are operands from the constant bank (DFMA R, R, c[3][..], R).  As literals
// ptxas materialises every one of them with two UMOV into a uniform register pair: 100 extra issue slots per warp-slice
// of the cloth colour pass, 5 % of its instructions (ncu source page, profiles/r2e_*).
#if defined(__CUDACC__)
static __constant__ double kAtanCoef[16] = {
    -0.033387823128967978, 0.015113763707864538,  -0.052173125594086317, 0.044995006697646829,
    -0.066662052597397748, 0.058768340023197307,  -0.090909080590304585, 0.076922810999214885,
    -0.14285714285329708,  0.11111111085322088,   -0.33333333333333324,  0.19999999999997014,
    0.41421356237309503,   0.78539816339744831,   1.5707963267948966,    3.1415926535897932};
#endif
#if defined(__CUDA_ARCH__)
#define SPGX_ATAN_C(i, literal) kAtanCoef[i]
#else
#define SPGX_ATAN_C(i, literal) (literal)
#endif
SPGX_HD double fastAtan2(double y, double x)
{
    const double ax = ::fabs(x), ay = ::fabs(y);
    const double mx = ::fmax(ax, ay), mn = ::fmin(ax, ay);
    const bool fold = mn > SPGX_ATAN_C(12, 0.41421356237309503) * mx;
    const double num = fold ? mn - mx : mn;
    const double den = fold ? mn + mx : mx;
    const double t = num * fastRcp(den);
    const double z = t * t, z2 = z * z;
    // q(z) = c1 + c2 z + ... + c12 z^11 = qe(z^2) + z qo(z^2)
    double qe = SPGX_ATAN_C(0, -0.033387823128967978);          // c11
    double qo = SPGX_ATAN_C(1, 0.015113763707864538);           // c12
    qe = ::fma(qe, z2, SPGX_ATAN_C(2, -0.052173125594086317));  // c9
    qo = ::fma(qo, z2, SPGX_ATAN_C(3, 0.044995006697646829));   // c10
    qe = ::fma(qe, z2, SPGX_ATAN_C(4, -0.066662052597397748));  // c7
    qo = ::fma(qo, z2, SPGX_ATAN_C(5, 0.058768340023197307));   // c8
    qe = ::fma(qe, z2, SPGX_ATAN_C(6, -0.090909080590304585));  // c5
    qo = ::fma(qo, z2, SPGX_ATAN_C(7, 0.076922810999214885));   // c6
    qe = ::fma(qe, z2, SPGX_ATAN_C(8, -0.14285714285329708));   // c3
    qo = ::fma(qo, z2, SPGX_ATAN_C(9, 0.11111111085322088));    // c4
    qe = ::fma(qe, z2, SPGX_ATAN_C(10, -0.33333333333333324));  // c1
    qo = ::fma(qo, z2, SPGX_ATAN_C(11, 0.19999999999997014));   // c2
    const double q = ::fma(qo, z, qe);
    double a = ::fma(t * z, q, t);
    a += fold ? SPGX_ATAN_C(13, 0.78539816339744831) : 0.0;
    if (ay > ax) a = SPGX_ATAN_C(14, 1.5707963267948966) - a;
    if (x < 0.0) a = SPGX_ATAN_C(15, 3.1415926535897932) - a;
    return ::copysign(a, y);
}

template <int ST>
struct Stream;

// ---- discrete bending, E = K/2 (theta - theta0)^2 (discreteBendingEnergy.cpp:9-30) ---------------------------------
// With e = x1 - x0, a = x2 - x1, b = x3 - x1, N1 = e x a, N2 = b x e:
//   theta = atan2(-|e| N1.b, N1.N2)                 [(n1 x n2).e^ = -|e| N1.b / (|N1||N2|),  n1.n2 = N1.N2 / (|N1||N2|)]
//   g = K d grad(theta),  H = K (grad grad^T + d Hess(theta)),  d = theta - theta0.
//
// Owner opposite to the edge (canonical position x2; neighbours {x0, x1, x3}; q = {theta0, K}):
//   grad = -|e| N1 / |N1|^2,   Hess = -|e| / |N1|^4 (N1 T^T + T N1^T),  T = e x N1.
template <>
struct Stream<ST_BEND_OPP> {
    static constexpr int NI = 3, NQ = 2;
    SPGX_HD static void eval(const V3d &xo, const V3d *xn, const double *q, V3d &gAcc, Sym3 &HAcc)
    {
        const V3d e = xn[1] - xn[0];
        const V3d a = xo - xn[1], b = xn[2] - xn[1];
        const V3d N1 = cross(e, a), N2 = cross(b, e);
        const double ee = dot(e, e);
        const double le = ee * fastRsqrt(ee);
        const double d = fastAtan2(-le * dot(N1, b), dot(N1, N2)) - q[0];
        const double K = q[1], Kd = K * d;
        const double i = fastRcp(dot(N1, N1));
        const double c = -le * i;  // grad(theta) = c N1
        const V3d T = cross(e, N1);
        // K grad grad^T + K d Hess = sym2(N1, w),  w = (K d c i) T + (K c^2 / 2) N1,  sym2(u, v) = u v^T + v u^T
        const double Kc = K * c;
        const V3d w = T * (Kd * c * i) + N1 * (0.5 * Kc * c);
        addSymOuter2u(HAcc, N1, w);
        const double Kdc = Kd * c;
        gAcc.x += Kdc * N1.x;
        gAcc.y += Kdc * N1.y;
        gAcc.z += Kdc * N1.z;
    }
};

// Owner at the start of the edge (canonical position x0; neighbours {x1, x2, x3}; q = {theta0, K}):
//   W = N / |N|^2, phi_a = a.e / |e|, phi_b = b.e / |e|,   grad = -(phi_a W1 + phi_b W2),
//   Hess = sym(W1 dphi_a^T) + sym(W2 dphi_b^T) - phi_a (t1 N1^T + N1 t1^T) / |N1|^4 + phi_b (t2 N2^T + N2 t2^T) / |N2|^4,
//   dphi_a = a / |e| - e (a.e) / |e|^3,  t1 = a x N1,  t2 = b x N2.
template <>
struct Stream<ST_BEND_EDGE> {
    static constexpr int NI = 3, NQ = 2;
    SPGX_HD static void eval(const V3d &xo, const V3d *xn, const double *q, V3d &gAcc, Sym3 &HAcc)
    {
        const V3d e = xn[0] - xo;
        const V3d a = xn[1] - xn[0], b = xn[2] - xn[0];
        const V3d N1 = cross(e, a), N2 = cross(b, e);
        const double ee = dot(e, e);
        const double ie = fastRsqrt(ee);
        const double le = ee * ie;
        const double d = fastAtan2(-le * dot(N1, b), dot(N1, N2)) - q[0];
        const double K = q[1], Kd = K * d;
        const double i1 = fastRcp(dot(N1, N1)), i2 = fastRcp(dot(N2, N2));
        const double ae = dot(a, e), be = dot(b, e);
        const double phia = ae * ie, phib = be * ie;
        const double ie3 = ie * ie * ie;
        const V3d ga = a * ie - e * (ae * ie3), gb = b * ie - e * (be * ie3);
        const double w1 = phia * i1, w2 = phib * i2;
        const V3d g = -(N1 * w1 + N2 * w2);
        const V3d t1 = cross(a, N1), t2 = cross(b, N2);
        // K d Hess = sym2(N1, v1) + sym2(N2, v2),  sym2(u, v) = u v^T + v u^T
        const double hKd = 0.5 * Kd;
        const V3d v1 = ga * (hKd * i1) - t1 * (Kd * w1 * i1);
        const V3d v2 = gb * (hKd * i2) + t2 * (Kd * w2 * i2);
        addSymOuter2u(HAcc, N1, v1);
        addSymOuter2u(HAcc, N2, v2);
        gAcc.x += Kd * g.x;
        gAcc.y += Kd * g.y;
        gAcc.z += Kd * g.z;
        addOuter(HAcc, g, K);
    }
};

// ---- StVK membrane (membraneStvkEnergy.cpp:9-28): C = (E00, E11, 2 E01), E = 1/2 C^T K C ---------------------------
// Neighbours {p, q} in cyclic order after the owner; q = {g_p(2), g_q(2), K00, K01, K22} with g_s the rows of the
// shape-gradient matrix (F = sum_s x_s g_s^T, g_o = -g_p - g_q = (ga, gb)).  With s = K C:
//   grad = (s0 ga + s2 gb) F0 + (s1 gb + s2 ga) F1
//   Hess = A F0 F0^T + B F1 F1^T + C (F0 F1^T + F1 F0^T) + (s0 ga^2 + s1 gb^2 + 2 s2 ga gb) I,
//   A = K00 ga^2 + K22 gb^2,  B = K00 gb^2 + K22 ga^2,  C = (K01 + K22) ga gb.
template <>
struct Stream<ST_MEMBRANE_STVK> {
    static constexpr int NI = 2, NQ = 7;
    SPGX_HD static void eval(const V3d &xo, const V3d *xn, const double *q, V3d &gAcc, Sym3 &HAcc)
    {
        const V3d dp = xn[0] - xo, dq = xn[1] - xo;
        const V3d F0 = dp * q[0] + dq * q[2], F1 = dp * q[1] + dq * q[3];
        const double K00 = q[4], K01 = q[5], K22 = q[6];
        const double C0 = 0.5 * (dot(F0, F0) - 1.0), C1 = 0.5 * (dot(F1, F1) - 1.0), C2 = dot(F0, F1);
        const double s0 = K00 * C0 + K01 * C1, s1 = K01 * C0 + K00 * C1, s2 = K22 * C2;
        const double ga = -(q[0] + q[2]), gb = -(q[1] + q[3]);
        const double al0 = s0 * ga + s2 * gb, al1 = s1 * gb + s2 * ga;
        gAcc.x = ::fma(al0, F0.x, ::fma(al1, F1.x, gAcc.x));
        gAcc.y = ::fma(al0, F0.y, ::fma(al1, F1.y, gAcc.y));
        gAcc.z = ::fma(al0, F0.z, ::fma(al1, F1.z, gAcc.z));
        const double gaa = ga * ga, gbb = gb * gb, gab = ga * gb;
        const double geo = s0 * gaa + s1 * gbb + 2.0 * (s2 * gab);
        const double A = K00 * gaa + K22 * gbb, B = K00 * gbb + K22 * gaa, C = (K01 + K22) * gab;
        const V3d P0 = F0 * A + F1 * C, P1 = F1 * B + F0 * C;
        HAcc.xx = ::fma(F0.x, P0.x, ::fma(F1.x, P1.x, HAcc.xx + geo));
        HAcc.yy = ::fma(F0.y, P0.y, ::fma(F1.y, P1.y, HAcc.yy + geo));
        HAcc.zz = ::fma(F0.z, P0.z, ::fma(F1.z, P1.z, HAcc.zz + geo));
        HAcc.xy = ::fma(F0.x, P0.y, ::fma(F1.x, P1.y, HAcc.xy));
        HAcc.xz = ::fma(F0.x, P0.z, ::fma(F1.x, P1.z, HAcc.xz));
        HAcc.yz = ::fma(F0.y, P0.z, ::fma(F1.y, P1.z, HAcc.yz));
    }
};

// ---- tetrahedra: neighbours {p, q, r} = the other three vertices in cyclic stencil order after the owner; the first
// 9 doubles are the rows g_p, g_q, g_r of the shape-gradient matrix (F = sum_s x_s g_s^T, g_o = -(g_p + g_q + g_r)),
// so that F = (x_p - x_o) g_p^T + (x_q - x_o) g_q^T + (x_r - x_o) g_r^T and dF = dx_o g_o^T.
SPGX_HD void tetFfromOwner(const V3d &xo, const V3d *xn, const double *q, V3d &F0, V3d &F1, V3d &F2, V3d &go)
{
    const V3d dp = xn[0] - xo, dq = xn[1] - xo, dr = xn[2] - xo;
    F0 = dp * q[0] + dq * q[3] + dr * q[6];
    F1 = dp * q[1] + dq * q[4] + dr * q[7];
    F2 = dp * q[2] + dq * q[5] + dr * q[8];
    go = V3d(-(q[0] + q[3] + q[6]), -(q[1] + q[4] + q[7]), -(q[2] + q[5] + q[8]));
}

// Stable Neo-Hookean (stableNeoHookeanEnergy.cpp:33-61): E = lambda/2 (J - alpha)^2 + mu/2 (tr(F^T F) - 3).
// J = det F is affine in a single vertex (dF = dx g^T is rank one), so with c = cof(F) g = dJ/dx_o:
//   grad = lambda (J - alpha) c + mu F g,   Hess = lambda c c^T + mu |g|^2 I.      q = {g_p, g_q, g_r, lambda V, mu V, alpha}
template <>
struct Stream<ST_TET_SNH> {
    static constexpr int NI = 3, NQ = 12;
    SPGX_HD static void eval(const V3d &xo, const V3d *xn, const double *q, V3d &gAcc, Sym3 &HAcc)
    {
        V3d F0, F1, F2, go;
        tetFfromOwner(xo, xn, q, F0, F1, F2, go);
        const double lam = q[9], mu = q[10], alpha = q[11];
        // cof(F) columns: F1 x F2, F2 x F0, F0 x F1  (J = F0 . (F1 x F2))
        const V3d c0 = cross(F1, F2), c1 = cross(F2, F0), c2 = cross(F0, F1);
        const double J = dot(F0, c0);
        const V3d c = c0 * go.x + c1 * go.y + c2 * go.z;
        const V3d Fg = F0 * go.x + F1 * go.y + F2 * go.z;
        const double h = lam * (J - alpha);
        gAcc.x += h * c.x + mu * Fg.x;
        gAcc.y += h * c.y + mu * Fg.y;
        gAcc.z += h * c.z + mu * Fg.z;
        addOuter(HAcc, c, lam);
        const double iso = mu * dot(go, go);
        HAcc.xx += iso;
        HAcc.yy += iso;
        HAcc.zz += iso;
    }
};

// StVK tetrahedron (stvkEnergy.cpp:41-65): E = lambda/2 tr(E)^2 + mu tr(E^2), E = (F^T F - I)/2.
// With S = lambda tr(E) I + 2 mu E (second Piola-Kirchhoff), w = F g:
//   grad = F S g,   Hess = (g^T S g) I + (lambda + mu) w w^T + mu |g|^2 F F^T.        q = {g_p, g_q, g_r, lambda V, mu V}
template <>
struct Stream<ST_TET_STVK> {
    static constexpr int NI = 3, NQ = 11;
    SPGX_HD static void eval(const V3d &xo, const V3d *xn, const double *q, V3d &gAcc, Sym3 &HAcc)
    {
        V3d F0, F1, F2, go;
        tetFfromOwner(xo, xn, q, F0, F1, F2, go);
        const double lam = q[9], mu = q[10];
        const double e00 = 0.5 * (dot(F0, F0) - 1.0), e11 = 0.5 * (dot(F1, F1) - 1.0), e22 = 0.5 * (dot(F2, F2) - 1.0);
        const double e01 = 0.5 * dot(F0, F1), e02 = 0.5 * dot(F0, F2), e12 = 0.5 * dot(F1, F2);
        const double ltr = lam * (e00 + e11 + e22), mu2 = 2.0 * mu;
        // S g
        const double sg0 = (ltr + mu2 * e00) * go.x + mu2 * (e01 * go.y + e02 * go.z);
        const double sg1 = (ltr + mu2 * e11) * go.y + mu2 * (e01 * go.x + e12 * go.z);
        const double sg2 = (ltr + mu2 * e22) * go.z + mu2 * (e02 * go.x + e12 * go.y);
        gAcc.x += F0.x * sg0 + F1.x * sg1 + F2.x * sg2;
        gAcc.y += F0.y * sg0 + F1.y * sg1 + F2.y * sg2;
        gAcc.z += F0.z * sg0 + F1.z * sg1 + F2.z * sg2;
        const V3d w = F0 * go.x + F1 * go.y + F2 * go.z;
        addOuter(HAcc, w, lam + mu);
        const double mg = mu * dot(go, go);
        addOuter(HAcc, F0, mg);
        addOuter(HAcc, F1, mg);
        addOuter(HAcc, F2, mg);
        const double iso = go.x * sg0 + go.y * sg1 + go.z * sg2;
        HAcc.xx += iso;
        HAcc.yy += iso;
        HAcc.zz += iso;
    }
};

// ---- wedge: one triangle of the owner's fan with the two hinges that share its normal --------------------------------
// A cloth vertex o with n incident triangles has n membrane incidences, n hinges whose edge starts at o and n hinges whose
// edge lies opposite to o: three incidences per triangle (o, A, B) of the fan --
//   the StVK membrane of (o, A, B),
//   the hinge on the edge o-A, canonical (x0, x1, x2, x3) = (o, A, B, C), C = the vertex beyond o-A,
//   the hinge on the edge A-B seen from o, canonical (x0, x1, x2, x3) = (A, B, o, R), R = the vertex beyond A-B
// -- and as separate incidences each of them fetches o's neighbours, forms the edge vectors and the triangle normal
// N = (A - o) x (B - o) and inverts |N|^2 for itself.  As ONE record (neighbours {A, B, C, R}, 4 x 8 bits;
// q = {g_A(2), g_B(2), K00, K01, K22 | theta0, K of o-A | theta0, K of A-B}) the edge vectors, N, 1 / |N|^2, a = B - A and
// T = a x N are formed once (T is t1 of the first hinge AND e x N1 of the second), the two sym2(N, .) Hessian terms merge
// into one, a third of the neighbour reads and two of three loop steps go away.  Same formulas as the three evaluators
// above (ST_MEMBRANE_STVK, ST_BEND_EDGE, ST_BEND_OPP), to which incidences that do not complete a wedge (mesh boundary,
// non-manifold input) still go; which incidence joins which wedge is decided on the host (api_build.cu: buildSlotRecords).
template <>
struct Stream<ST_WEDGE> {
    static constexpr int NI = 4, NQ = 11;
    SPGX_HD static void eval(const V3d &xo, const V3d *xn, const double *q, V3d &gAcc, Sym3 &HAcc)
    {
        const V3d eA = xn[0] - xo, eB = xn[1] - xo;
        {  // membrane (membraneStvkEnergy.cpp:9-28), dp = eA, dq = eB
            const V3d F0 = eA * q[0] + eB * q[2], F1 = eA * q[1] + eB * q[3];
            const double K00 = q[4], K01 = q[5], K22 = q[6];
            const double C0 = 0.5 * (dot(F0, F0) - 1.0), C1 = 0.5 * (dot(F1, F1) - 1.0), C2 = dot(F0, F1);
            const double s0 = K00 * C0 + K01 * C1, s1 = K01 * C0 + K00 * C1, s2 = K22 * C2;
            const double ga = -(q[0] + q[2]), gb = -(q[1] + q[3]);
            const double al0 = s0 * ga + s2 * gb, al1 = s1 * gb + s2 * ga;
            gAcc.x = ::fma(al0, F0.x, ::fma(al1, F1.x, gAcc.x));
            gAcc.y = ::fma(al0, F0.y, ::fma(al1, F1.y, gAcc.y));
            gAcc.z = ::fma(al0, F0.z, ::fma(al1, F1.z, gAcc.z));
            const double gaa = ga * ga, gbb = gb * gb, gab = ga * gb;
            const double geo = s0 * gaa + s1 * gbb + 2.0 * (s2 * gab);
            const double A = K00 * gaa + K22 * gbb, B = K00 * gbb + K22 * gaa, C = (K01 + K22) * gab;
            const V3d P0 = F0 * A + F1 * C, P1 = F1 * B + F0 * C;
            HAcc.xx = ::fma(F0.x, P0.x, ::fma(F1.x, P1.x, HAcc.xx + geo));
            HAcc.yy = ::fma(F0.y, P0.y, ::fma(F1.y, P1.y, HAcc.yy + geo));
            HAcc.zz = ::fma(F0.z, P0.z, ::fma(F1.z, P1.z, HAcc.zz + geo));
            HAcc.xy = ::fma(F0.x, P0.y, ::fma(F1.x, P1.y, HAcc.xy));
            HAcc.xz = ::fma(F0.x, P0.z, ::fma(F1.x, P1.z, HAcc.xz));
            HAcc.yz = ::fma(F0.y, P0.z, ::fma(F1.y, P1.z, HAcc.yz));
        }
        // shared by the two hinges
        const V3d N = cross(eA, eB);
        const double iN = fastRcp(dot(N, N));
        const V3d a = eB - eA;  // B - A
        const V3d T = cross(a, N);
        V3d vN;        // the Hessian gets sym2(N, vN), the gradient cN * N
        double cN;
        {  // hinge on o-A, owner at the start of the edge (Stream<ST_BEND_EDGE> with e = eA, a, b = C - A, N1 = N, t1 = T)
            const V3d b = xn[2] - xn[0];
            const V3d N2 = cross(b, eA);
            const double ee = dot(eA, eA);
            const double ie = fastRsqrt(ee);
            const double le = ee * ie;
            const double d = fastAtan2(-le * dot(N, b), dot(N, N2)) - q[7];
            const double K = q[8], Kd = K * d;
            const double i2 = fastRcp(dot(N2, N2));
            const double ae = dot(a, eA), be = dot(b, eA);
            const double phia = ae * ie, phib = be * ie;
            const double ie3 = ie * ie * ie;
            const V3d ga = a * ie - eA * (ae * ie3), gb = b * ie - eA * (be * ie3);
            const double w1 = phia * iN, w2 = phib * i2;
            const V3d g = -(N * w1 + N2 * w2);
            const V3d t2 = cross(b, N2);
            const double hKd = 0.5 * Kd;
            vN = ga * (hKd * iN) - T * (Kd * w1 * iN);
            const V3d v2 = gb * (hKd * i2) + t2 * (Kd * w2 * i2);
            addSymOuter2u(HAcc, N2, v2);
            addOuter(HAcc, g, K);
            // Kd g = -(Kd w1) N - (Kd w2) N2: the N part joins the second hinge's
            cN = -(Kd * w1);
            const double c2 = -(Kd * w2);
            gAcc.x = ::fma(c2, N2.x, gAcc.x);
            gAcc.y = ::fma(c2, N2.y, gAcc.y);
            gAcc.z = ::fma(c2, N2.z, gAcc.z);
        }
        {  // hinge on A-B, owner opposite (Stream<ST_BEND_OPP> with e = a, b = R - B, N1 = N, e x N1 = T)
            const V3d b = xn[3] - xn[1];
            const V3d N2 = cross(b, a);
            const double aa = dot(a, a);
            const double la = aa * fastRsqrt(aa);
            const double d = fastAtan2(-la * dot(N, b), dot(N, N2)) - q[9];
            const double K = q[10], Kd = K * d;
            const double c = -la * iN;  // grad(theta) = c N
            const double Kc = K * c;
            vN = vN + T * (Kd * c * iN) + N * (0.5 * Kc * c);
            cN = ::fma(Kd, c, cN);
        }
        addSymOuter2u(HAcc, N, vN);
        gAcc.x = ::fma(cN, N.x, gAcc.x);
        gAcc.y = ::fma(cN, N.y, gAcc.y);
        gAcc.z = ::fma(cN, N.z, gAcc.z);
    }
};

// ---- host side: which stream an incidence of `kind` at stencil position `local` goes to, and its canonical record ----
// returns -1 when the kind has no stream (it then runs through the generic CSR path of the kernel)
SPGX_HD int streamTypeOf(int kind, int local)
{
    switch (kind) {
    case K_DISCRETE_BENDING: return local < 2 ? ST_BEND_EDGE : ST_BEND_OPP;
    case K_MEMBRANE_STVK: return ST_MEMBRANE_STVK;
    case K_TET_SNH: return ST_TET_SNH;
    case K_TET_STVK: return ST_TET_STVK;
    default: return -1;
    }
}
SPGX_HD int streamNI(int st)
{
    return st == ST_MEMBRANE_STVK ? 2 : (st == ST_WEDGE ? 4 : 3);
}
// word of the 8-word slice descriptor (kernels.h: VbdArgs::sliceDesc) that holds the first step of a stream type.  Words 2..6
// are the five single-incidence types; a wedge stream only exists next to cloth streams and takes the word of ST_TET_STVK
// (buildStreams forms no wedges in a group that also holds tetrahedra)
SPGX_HD constexpr int streamDescWord(int st)
{
    return st == ST_WEDGE ? 2 + ST_TET_STVK : 2 + st;
}
SPGX_HD int streamNQ(int st)
{
    switch (st) {
    case ST_BEND_EDGE:
    case ST_BEND_OPP: return 2;
    case ST_MEMBRANE_STVK: return 7;
    case ST_TET_SNH: return 12;
    case ST_TET_STVK: return 11;
    case ST_WEDGE: return 11;
    default: return 0;
    }
}
// stencil: the S vertex ids of the stencil (any id space), p: the kind's record (energies.cuh); nbr: NI ids out
SPGX_HD void makeStreamRecord(int kind, const int32_t *stencil, const double *p, int local, int32_t *nbr, double *q)
{
    switch (kind) {
    case K_DISCRETE_BENDING:
        if (local < 2) {  // owner on the edge: canonical x0; local 1 flips the edge, which negates theta
            nbr[0] = stencil[1 - local], nbr[1] = stencil[2], nbr[2] = stencil[3];
            q[0] = local == 0 ? p[0] : -p[0];
        } else {  // owner opposite: canonical x2; local 3 swaps the opposite vertices, which negates theta
            nbr[0] = stencil[0], nbr[1] = stencil[1], nbr[2] = stencil[local == 2 ? 3 : 2];
            q[0] = local == 2 ? p[0] : -p[0];
        }
        q[1] = p[1];
        break;
    case K_MEMBRANE_STVK: {
        const int ip = (local + 1) % 3, iq = (local + 2) % 3;
        const double gs[3][2] = {{-(p[0] + p[2]), -(p[1] + p[3])}, {p[0], p[1]}, {p[2], p[3]}};
        nbr[0] = stencil[ip], nbr[1] = stencil[iq];
        q[0] = gs[ip][0], q[1] = gs[ip][1], q[2] = gs[iq][0], q[3] = gs[iq][1];
        q[4] = p[4], q[5] = p[5], q[6] = p[6];
        break;
    }
    case K_TET_SNH:
    case K_TET_STVK: {
        // rows of the shape-gradient matrix: g_1..g_3 = rows of Dinv (p[0..8] row-major), g_0 = -(g_1 + g_2 + g_3)
        double gs[4][3];
        for (int c = 0; c < 3; ++c) {
            gs[1][c] = p[c], gs[2][c] = p[3 + c], gs[3][c] = p[6 + c];
            gs[0][c] = -(p[c] + p[3 + c] + p[6 + c]);
        }
        for (int k = 0; k < 3; ++k) {
            const int s = (local + 1 + k) % 4;
            nbr[k] = stencil[s];
            for (int c = 0; c < 3; ++c) q[3 * k + c] = gs[s][c];
        }
        q[9] = p[9], q[10] = p[10];
        if (kind == K_TET_SNH) q[11] = p[13];
        break;
    }
    default: break;
    }
}

// The wedge record of owner o = tri[localT] (streams.cuh: Stream<ST_WEDGE>).  tri / pm: a K_MEMBRANE_STVK stencil and its
// record; he / ph: the K_DISCRETE_BENDING stencil whose hinge edge joins o (position localE < 2) to A, and its record; ho / po:
// the one whose hinge edge is {A, B} with o opposite (position localO >= 2).  A is the triangle vertex after (aNext) or
// before o in the stencil's cyclic order, B the other one.  The hinge angle changes sign when the edge is reversed or the
// opposite vertices are swapped (discreteBendingEnergy.cpp:9-30), so bringing a stencil into the canonical order only
// negates theta0.  nbr: {A, B, C, R} out; returns false when the three stencils do not form a wedge around o.
SPGX_HD bool makeWedgeRecord(const int32_t *tri, const double *pm, int localT, bool aNext, const int32_t *he, const double *ph,
                             int localE, const int32_t *ho, const double *po, int localO, int32_t *nbr, double *q)
{
    const int ia = aNext ? (localT + 1) % 3 : (localT + 2) % 3, ib = aNext ? (localT + 2) % 3 : (localT + 1) % 3;
    const int32_t o = tri[localT], A = tri[ia], B = tri[ib];
    if (localE > 1 || localO < 2 || he[localE] != o || he[1 - localE] != A || ho[localO] != o) return false;
    double th0E = localE == 0 ? ph[0] : -ph[0];  // (o, A, he[2], he[3])
    int32_t C;
    if (he[2] == B)
        C = he[3];
    else if (he[3] == B)
        C = he[2], th0E = -th0E;  // opposite vertices swapped
    else
        return false;
    double th0O = localO == 2 ? po[0] : -po[0];  // (ho[0], ho[1], o, R)
    const int32_t R = ho[localO == 2 ? 3 : 2];
    if (ho[0] == A && ho[1] == B) {
    } else if (ho[0] == B && ho[1] == A)
        th0O = -th0O;  // edge reversed
    else
        return false;
    nbr[0] = A, nbr[1] = B, nbr[2] = C, nbr[3] = R;
    const double gs[3][2] = {{-(pm[0] + pm[2]), -(pm[1] + pm[3])}, {pm[0], pm[1]}, {pm[2], pm[3]}};
    q[0] = gs[ia][0], q[1] = gs[ia][1], q[2] = gs[ib][0], q[3] = gs[ib][1];
    q[4] = pm[4], q[5] = pm[5], q[6] = pm[6];
    q[7] = th0E, q[8] = ph[1], q[9] = th0O, q[10] = po[1];
    return true;
}

// run one canonical evaluator by runtime stream type (tests, host and device)
SPGX_HD void evalStream(int st, const V3d &xo, const V3d *xn, const double *q, V3d &g, Sym3 &H)
{
    switch (st) {
    case ST_BEND_EDGE: Stream<ST_BEND_EDGE>::eval(xo, xn, q, g, H); break;
    case ST_BEND_OPP: Stream<ST_BEND_OPP>::eval(xo, xn, q, g, H); break;
    case ST_MEMBRANE_STVK: Stream<ST_MEMBRANE_STVK>::eval(xo, xn, q, g, H); break;
    case ST_TET_SNH: Stream<ST_TET_SNH>::eval(xo, xn, q, g, H); break;
    case ST_TET_STVK: Stream<ST_TET_STVK>::eval(xo, xn, q, g, H); break;
    case ST_WEDGE: Stream<ST_WEDGE>::eval(xo, xn, q, g, H); break;
    default: break;
    }
}
}  // namespace spgx
