// Forward-mode dual numbers for the device-side energy evaluators.
//
// Replaces TinyAD::Scalar<k, double, with_hessian> (reference: external/TinyAD/include/TinyAD/Scalar.hh,
// ops actually used by the hot path: ctor-active :59-67, chain :176-191, unary minus :193-207,
// sqrt :209-220, + :543-589, - :620-700, * :702-771, / :787-841, atan2 :857-882).  Differences that
// matter on a GPU: K is the number of *seeded* DOFs (3 for a VBD vertex block, 3*S for an XPBD
// constraint Jacobian) rather than always 3*S, the Hessian is stored packed-symmetric
// (K(K+1)/2 doubles, not K*K), and every loop is compile-time unrolled so a Dual lives in registers.
#pragma once
#include <cmath>

#if defined(__CUDACC__)
#define SPGX_HD __host__ __device__ __forceinline__
#else
#define SPGX_HD inline
#endif

namespace spgx
{
template <int K, bool HESS>
struct Dual {
    static constexpr int NG = K;
    static constexpr int NH = HESS ? K * (K + 1) / 2 : 1;
    static constexpr bool kHess = HESS;
    double v;
    double g[K];
    double h[NH];

    SPGX_HD Dual() {}
    // passive constant (Scalar.hh:51-55)
    SPGX_HD Dual(double val) : v(val)
    {
#pragma unroll
        for (int i = 0; i < K; ++i) g[i] = 0.0;
#pragma unroll
        for (int i = 0; i < NH; ++i) h[i] = 0.0;
    }
    // active variable with unit gradient in slot `idx` (Scalar.hh:59-67); idx < 0 or >= K means passive
    SPGX_HD Dual(double val, int idx) : v(val)
    {
#pragma unroll
        for (int i = 0; i < K; ++i) g[i] = (i == idx) ? 1.0 : 0.0;
#pragma unroll
        for (int i = 0; i < NH; ++i) h[i] = 0.0;
    }
    // packed index of the (i, j) Hessian entry, i <= j
    SPGX_HD static constexpr int hidx(int i, int j) { return i * K - (i * (i - 1)) / 2 + (j - i); }
    SPGX_HD double hess(int i, int j) const { return i <= j ? h[hidx(i, j)] : h[hidx(j, i)]; }
};

template <class T>
struct is_dual {
    static constexpr bool value = false;
};
template <int K, bool H>
struct is_dual<Dual<K, H>> {
    static constexpr bool value = true;
};

SPGX_HD double value_of(double a)
{
    return a;
}
template <int K, bool H>
SPGX_HD double value_of(const Dual<K, H> &a)
{
    return a.v;
}

// f(a) given f, f', f'' at a.v (Scalar.hh:176-191)
template <int K, bool H>
SPGX_HD Dual<K, H> chain(const Dual<K, H> &a, double f, double df, double ddf)
{
    Dual<K, H> r;
    r.v = f;
#pragma unroll
    for (int i = 0; i < K; ++i) r.g[i] = df * a.g[i];
    if constexpr (H) {
#pragma unroll
        for (int i = 0; i < K; ++i)
#pragma unroll
            for (int j = i; j < K; ++j) {
                const int q = Dual<K, H>::hidx(i, j);
                r.h[q] = ddf * a.g[i] * a.g[j] + df * a.h[q];
            }
    }
    return r;
}

template <int K, bool H>
SPGX_HD Dual<K, H> operator-(const Dual<K, H> &a)
{
    Dual<K, H> r;
    r.v = -a.v;
#pragma unroll
    for (int i = 0; i < K; ++i) r.g[i] = -a.g[i];
    if constexpr (H) {
#pragma unroll
        for (int i = 0; i < Dual<K, H>::NH; ++i) r.h[i] = -a.h[i];
    }
    return r;
}

template <int K, bool H>
SPGX_HD Dual<K, H> operator+(const Dual<K, H> &a, const Dual<K, H> &b)
{
    Dual<K, H> r;
    r.v = a.v + b.v;
#pragma unroll
    for (int i = 0; i < K; ++i) r.g[i] = a.g[i] + b.g[i];
    if constexpr (H) {
#pragma unroll
        for (int i = 0; i < Dual<K, H>::NH; ++i) r.h[i] = a.h[i] + b.h[i];
    }
    return r;
}
template <int K, bool H>
SPGX_HD Dual<K, H> operator+(const Dual<K, H> &a, double b)
{
    Dual<K, H> r = a;
    r.v += b;
    return r;
}
template <int K, bool H>
SPGX_HD Dual<K, H> operator+(double b, const Dual<K, H> &a)
{
    return a + b;
}

template <int K, bool H>
SPGX_HD Dual<K, H> operator-(const Dual<K, H> &a, const Dual<K, H> &b)
{
    Dual<K, H> r;
    r.v = a.v - b.v;
#pragma unroll
    for (int i = 0; i < K; ++i) r.g[i] = a.g[i] - b.g[i];
    if constexpr (H) {
#pragma unroll
        for (int i = 0; i < Dual<K, H>::NH; ++i) r.h[i] = a.h[i] - b.h[i];
    }
    return r;
}
template <int K, bool H>
SPGX_HD Dual<K, H> operator-(const Dual<K, H> &a, double b)
{
    Dual<K, H> r = a;
    r.v -= b;
    return r;
}
template <int K, bool H>
SPGX_HD Dual<K, H> operator-(double b, const Dual<K, H> &a)
{
    Dual<K, H> r = -a;
    r.v += b;
    return r;
}

template <int K, bool H>
SPGX_HD Dual<K, H> operator*(const Dual<K, H> &a, const Dual<K, H> &b)
{
    Dual<K, H> r;
    r.v = a.v * b.v;
#pragma unroll
    for (int i = 0; i < K; ++i) r.g[i] = b.v * a.g[i] + a.v * b.g[i];
    if constexpr (H) {
#pragma unroll
        for (int i = 0; i < K; ++i)
#pragma unroll
            for (int j = i; j < K; ++j) {
                const int q = Dual<K, H>::hidx(i, j);
                r.h[q] = b.v * a.h[q] + a.v * b.h[q] + a.g[i] * b.g[j] + a.g[j] * b.g[i];
            }
    }
    return r;
}
template <int K, bool H>
SPGX_HD Dual<K, H> operator*(const Dual<K, H> &a, double b)
{
    Dual<K, H> r;
    r.v = a.v * b;
#pragma unroll
    for (int i = 0; i < K; ++i) r.g[i] = a.g[i] * b;
    if constexpr (H) {
#pragma unroll
        for (int i = 0; i < Dual<K, H>::NH; ++i) r.h[i] = a.h[i] * b;
    }
    return r;
}
template <int K, bool H>
SPGX_HD Dual<K, H> operator*(double b, const Dual<K, H> &a)
{
    return a * b;
}

template <int K, bool H>
SPGX_HD Dual<K, H> recip(const Dual<K, H> &b)
{
    const double r = 1.0 / b.v;
    return chain(b, r, -r * r, 2.0 * r * r * r);
}
template <int K, bool H>
SPGX_HD Dual<K, H> operator/(const Dual<K, H> &a, const Dual<K, H> &b)
{
    return a * recip(b);
}
template <int K, bool H>
SPGX_HD Dual<K, H> operator/(const Dual<K, H> &a, double b)
{
    return a * (1.0 / b);
}
template <int K, bool H>
SPGX_HD Dual<K, H> operator/(double a, const Dual<K, H> &b)
{
    return recip(b) * a;
}

template <int K, bool H>
SPGX_HD Dual<K, H> &operator+=(Dual<K, H> &a, const Dual<K, H> &b)
{
    a = a + b;
    return a;
}
template <int K, bool H>
SPGX_HD Dual<K, H> &operator-=(Dual<K, H> &a, const Dual<K, H> &b)
{
    a = a - b;
    return a;
}

SPGX_HD double dsqrt(double a)
{
    return ::sqrt(a);
}
template <int K, bool H>
SPGX_HD Dual<K, H> dsqrt(const Dual<K, H> &a)
{
    const double f = ::sqrt(a.v);
    return chain(a, f, 0.5 / f, -0.25 / (f * a.v));
}
// 1/sqrt(a): one chain instead of sqrt followed by a division
SPGX_HD double drsqrt(double a)
{
    return 1.0 / ::sqrt(a);
}
template <int K, bool H>
SPGX_HD Dual<K, H> drsqrt(const Dual<K, H> &a)
{
    const double f = 1.0 / ::sqrt(a.v);
    const double ia = 1.0 / a.v;
    return chain(a, f, -0.5 * f * ia, 0.75 * f * ia * ia);
}

SPGX_HD double datan2(double y, double x)
{
    return ::atan2(y, x);
}
// Scalar.hh:857-882 written with the second partials of atan2 (same derivative, symmetric form)
template <int K, bool H>
SPGX_HD Dual<K, H> datan2(const Dual<K, H> &y, const Dual<K, H> &x)
{
    Dual<K, H> r;
    r.v = ::atan2(y.v, x.v);
    const double s = 1.0 / (x.v * x.v + y.v * y.v);
    const double uy = x.v * s, ux = -y.v * s;
#pragma unroll
    for (int i = 0; i < K; ++i) r.g[i] = uy * y.g[i] + ux * x.g[i];
    if constexpr (H) {
        const double uxx = 2.0 * x.v * y.v * s * s;  // u_yy = -u_xx
        const double uxy = (y.v * y.v - x.v * x.v) * s * s;
#pragma unroll
        for (int i = 0; i < K; ++i)
#pragma unroll
            for (int j = i; j < K; ++j) {
                const int q = Dual<K, H>::hidx(i, j);
                r.h[q] = uy * y.h[q] + ux * x.h[q] + uxx * (x.g[i] * x.g[j] - y.g[i] * y.g[j]) +
                         uxy * (y.g[i] * x.g[j] + y.g[j] * x.g[i]);
            }
    }
    return r;
}
// ---- mixed second-order dual --------------------------------------------------------------------------------------
// Value, the two 3-gradients with respect to the coordinates of two vertices a and b, and ONLY the mixed second
// derivatives d2/(da_i db_j) (3 x 3): 16 doubles instead of the 28 of Dual<6, true>.  Mixed partials are closed under the
// chain and product rules ((f g)_ab = f_ab g + f_a g_b + f_b g_a + f g_ab, phi(f)_ab = phi'' f_a f_b + phi' f_ab), so an
// off-diagonal 3 x 3 block of an element Hessian needs nothing else.  Seeds: slots 0-2 = a, 3-5 = b, anything else passive.
struct DualMixed {
    double v;
    double ga[3], gb[3];
    double h[9];  // h[3 i + j] = d2 / (da_i db_j)

    SPGX_HD DualMixed() {}
    SPGX_HD DualMixed(double val) : v(val)
    {
#pragma unroll
        for (int i = 0; i < 3; ++i) ga[i] = gb[i] = 0.0;
#pragma unroll
        for (int i = 0; i < 9; ++i) h[i] = 0.0;
    }
    SPGX_HD DualMixed(double val, int idx) : v(val)
    {
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            ga[i] = (i == idx) ? 1.0 : 0.0;
            gb[i] = (i + 3 == idx) ? 1.0 : 0.0;
        }
#pragma unroll
        for (int i = 0; i < 9; ++i) h[i] = 0.0;
    }
};
template <>
struct is_dual<DualMixed> {
    static constexpr bool value = true;
};
SPGX_HD double value_of(const DualMixed &a)
{
    return a.v;
}
SPGX_HD DualMixed chain(const DualMixed &a, double f, double df, double ddf)
{
    DualMixed r;
    r.v = f;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        r.ga[i] = df * a.ga[i];
        r.gb[i] = df * a.gb[i];
    }
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) r.h[3 * i + j] = ddf * a.ga[i] * a.gb[j] + df * a.h[3 * i + j];
    return r;
}
SPGX_HD DualMixed operator-(const DualMixed &a)
{
    return chain(a, -a.v, -1.0, 0.0);
}
SPGX_HD DualMixed operator+(const DualMixed &a, const DualMixed &b)
{
    DualMixed r;
    r.v = a.v + b.v;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        r.ga[i] = a.ga[i] + b.ga[i];
        r.gb[i] = a.gb[i] + b.gb[i];
    }
#pragma unroll
    for (int i = 0; i < 9; ++i) r.h[i] = a.h[i] + b.h[i];
    return r;
}
SPGX_HD DualMixed operator-(const DualMixed &a, const DualMixed &b)
{
    DualMixed r;
    r.v = a.v - b.v;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        r.ga[i] = a.ga[i] - b.ga[i];
        r.gb[i] = a.gb[i] - b.gb[i];
    }
#pragma unroll
    for (int i = 0; i < 9; ++i) r.h[i] = a.h[i] - b.h[i];
    return r;
}
SPGX_HD DualMixed operator+(const DualMixed &a, double b)
{
    DualMixed r = a;
    r.v += b;
    return r;
}
SPGX_HD DualMixed operator+(double b, const DualMixed &a)
{
    return a + b;
}
SPGX_HD DualMixed operator-(const DualMixed &a, double b)
{
    DualMixed r = a;
    r.v -= b;
    return r;
}
SPGX_HD DualMixed operator-(double b, const DualMixed &a)
{
    DualMixed r = -a;
    r.v += b;
    return r;
}
SPGX_HD DualMixed operator*(const DualMixed &a, const DualMixed &b)
{
    DualMixed r;
    r.v = a.v * b.v;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        r.ga[i] = b.v * a.ga[i] + a.v * b.ga[i];
        r.gb[i] = b.v * a.gb[i] + a.v * b.gb[i];
    }
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            r.h[3 * i + j] = b.v * a.h[3 * i + j] + a.v * b.h[3 * i + j] + a.ga[i] * b.gb[j] + b.ga[i] * a.gb[j];
    return r;
}
SPGX_HD DualMixed operator*(const DualMixed &a, double b)
{
    return chain(a, a.v * b, b, 0.0);
}
SPGX_HD DualMixed operator*(double b, const DualMixed &a)
{
    return a * b;
}
SPGX_HD DualMixed recip(const DualMixed &b)
{
    const double r = 1.0 / b.v;
    return chain(b, r, -r * r, 2.0 * r * r * r);
}
SPGX_HD DualMixed operator/(const DualMixed &a, const DualMixed &b)
{
    return a * recip(b);
}
SPGX_HD DualMixed operator/(const DualMixed &a, double b)
{
    return a * (1.0 / b);
}
SPGX_HD DualMixed operator/(double a, const DualMixed &b)
{
    return recip(b) * a;
}
SPGX_HD DualMixed &operator+=(DualMixed &a, const DualMixed &b)
{
    a = a + b;
    return a;
}
SPGX_HD DualMixed &operator-=(DualMixed &a, const DualMixed &b)
{
    a = a - b;
    return a;
}
SPGX_HD DualMixed dsqrt(const DualMixed &a)
{
    const double f = ::sqrt(a.v);
    return chain(a, f, 0.5 / f, -0.25 / (f * a.v));
}
SPGX_HD DualMixed drsqrt(const DualMixed &a)
{
    const double f = 1.0 / ::sqrt(a.v);
    const double ia = 1.0 / a.v;
    return chain(a, f, -0.5 * f * ia, 0.75 * f * ia * ia);
}
SPGX_HD DualMixed datan2(const DualMixed &y, const DualMixed &x)
{
    DualMixed r;
    r.v = ::atan2(y.v, x.v);
    const double s = 1.0 / (x.v * x.v + y.v * y.v);
    const double uy = x.v * s, ux = -y.v * s;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        r.ga[i] = uy * y.ga[i] + ux * x.ga[i];
        r.gb[i] = uy * y.gb[i] + ux * x.gb[i];
    }
    const double uxx = 2.0 * x.v * y.v * s * s;  // u_yy = -u_xx
    const double uxy = (y.v * y.v - x.v * x.v) * s * s;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            r.h[3 * i + j] = uy * y.h[3 * i + j] + ux * x.h[3 * i + j] + uxx * (x.ga[i] * x.gb[j] - y.ga[i] * y.gb[j]) +
                             uxy * (y.ga[i] * x.gb[j] + x.ga[i] * y.gb[j]);
    return r;
}
}  // namespace spgx
