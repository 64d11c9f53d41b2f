// TEST INFRASTRUCTURE -- independent CPU restatement of the solver hot path of alexrodag/spg @ 75122ce.
//
// Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may load the library built from this file
// (oracle/libspgport.so, `make -C oracle port`); the product (spg_b200/, include/) never does.  It shares no code with
// spg_b200/csrc: scalar forward-mode autodiff over the FULL stencil (value, k-gradient, k x k Hessian, k = 3S) exactly
// as the reference does through TinyAD, plain loops, OpenMP over the vertices / stencils of one colour.
// Parity of this restatement is PINNED: tests/test_port_oracle.py checks it against the golden fixtures generated from
// the compiled reference (tests/golden/) and, where oracle/_ref/libspgref.so is present, against the reference itself.
//
// It exports the same `ref_*` C entry points as oracle/ref_driver.cpp (minus the scene builders), so
// tests/oracle_api.py drives either library.  Every function cites the reference lines it follows
// (paths relative to the reference root).
#include <omp.h>

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace port
{
enum Kind {
    K_SPRING = 0, K_SPRING_CONTINUUM, K_DISCRETE_BENDING, K_QUADRATIC_BENDING, K_MEMBRANE_STVK, K_MEMBRANE_CHOI,
    K_MEMBRANE_BW, K_TET_STVK, K_TET_SNH, K_ANCHOR, K_BENDING_BW, K_SPRING_SQUARED_CONTINUUM, K_TET_SSNH, K_STATIC_LOAD,
    K_COUNT
};
constexpr int kStencilSize[K_COUNT] = {2, 2, 4, 4, 3, 3, 3, 4, 4, 1, 4, 2, 4, 1};
constexpr int kNumConstraints[K_COUNT] = {1, 1, 1, 3, 3, 4, 3, 2, 2, 1, 1, 1, 2, 1};
constexpr int kNumModelParams[K_COUNT] = {2, 2, 2, 2, 2, 3, 3, 2, 2, 4, 2, 2, 2, 4};

// ---------------------------------------------------------------------------------------------------------------------
// Forward-mode scalar with dense gradient and Hessian: external/TinyAD/include/TinyAD/Scalar.hh
// (data :1304-1309; active-variable ctor :59-67; chain :176-191; unary minus :193-207; sqrt :209-220; + :543-589;
//  - :620-700; * :702-771; / :787-841; atan2 :857-882; comparisons act on .val :896-1058).
// ---------------------------------------------------------------------------------------------------------------------
template <int K>
struct AD {
    double val = 0;
    double g[K];
    double h[K * K];
    AD() { zero(); }
    AD(double v) : val(v) { zero(); }
    AD(double v, int idx) : val(v)  // Scalar.hh:59-67: d val / d x_idx = 1
    {
        zero();
        g[idx] = 1.0;
    }
    void zero()
    {
        std::fill(g, g + K, 0.0);
        std::fill(h, h + K * K, 0.0);
    }
};

// f(a) with first / second derivative df, ddf at a.val (Scalar.hh:176-191)
template <int K>
AD<K> chain(double f, double df, double ddf, const AD<K> &a)
{
    AD<K> r;
    r.val = f;
    for (int i = 0; i < K; ++i) r.g[i] = df * a.g[i];
    for (int i = 0; i < K; ++i)
        for (int j = 0; j < K; ++j) r.h[i * K + j] = ddf * a.g[i] * a.g[j] + df * a.h[i * K + j];
    return r;
}
template <int K>
AD<K> operator-(const AD<K> &a)
{
    return chain(-a.val, -1.0, 0.0, a);
}
template <int K>
AD<K> sqrt(const AD<K> &a)  // Scalar.hh:209-220
{
    const double f = std::sqrt(a.val);
    return chain(f, 0.5 / f, -0.25 / (f * a.val), a);
}
template <int K>
AD<K> operator+(const AD<K> &a, const AD<K> &b)
{
    AD<K> r;
    r.val = a.val + b.val;
    for (int i = 0; i < K; ++i) r.g[i] = a.g[i] + b.g[i];
    for (int i = 0; i < K * K; ++i) r.h[i] = a.h[i] + b.h[i];
    return r;
}
template <int K>
AD<K> operator-(const AD<K> &a, const AD<K> &b)
{
    AD<K> r;
    r.val = a.val - b.val;
    for (int i = 0; i < K; ++i) r.g[i] = a.g[i] - b.g[i];
    for (int i = 0; i < K * K; ++i) r.h[i] = a.h[i] - b.h[i];
    return r;
}
template <int K>
AD<K> operator*(const AD<K> &a, const AD<K> &b)  // Scalar.hh:702-771
{
    AD<K> r;
    r.val = a.val * b.val;
    for (int i = 0; i < K; ++i) r.g[i] = b.val * a.g[i] + a.val * b.g[i];
    for (int i = 0; i < K; ++i)
        for (int j = 0; j < K; ++j)
            r.h[i * K + j] = b.val * a.h[i * K + j] + a.g[i] * b.g[j] + b.g[i] * a.g[j] + a.val * b.h[i * K + j];
    return r;
}
template <int K>
AD<K> operator/(const AD<K> &a, const AD<K> &b)  // Scalar.hh:787-841
{
    AD<K> r;
    r.val = a.val / b.val;
    for (int i = 0; i < K; ++i) r.g[i] = (b.val * a.g[i] - a.val * b.g[i]) / (b.val * b.val);
    for (int i = 0; i < K; ++i)
        for (int j = 0; j < K; ++j)
            r.h[i * K + j] = (a.h[i * K + j] - r.g[i] * b.g[j] - b.g[i] * r.g[j] - r.val * b.h[i * K + j]) / b.val;
    return r;
}
template <int K>
AD<K> operator+(const AD<K> &a, double b) { AD<K> r = a; r.val += b; return r; }
template <int K>
AD<K> operator-(const AD<K> &a, double b) { AD<K> r = a; r.val -= b; return r; }
template <int K>
AD<K> operator*(double b, const AD<K> &a) { return chain(a.val * b, b, 0.0, a); }
template <int K>
AD<K> operator*(const AD<K> &a, double b) { return chain(a.val * b, b, 0.0, a); }
template <int K>
AD<K> operator/(const AD<K> &a, double b) { return chain(a.val / b, 1.0 / b, 0.0, a); }
template <int K>
AD<K> atan2(const AD<K> &y, const AD<K> &x)  // Scalar.hh:857-882
{
    AD<K> r;
    const double r2 = x.val * x.val + y.val * y.val;
    const double ty = x.val / r2, tx = -y.val / r2;
    const double tyy = -2 * x.val * y.val / (r2 * r2), txx = -tyy, txy = (y.val * y.val - x.val * x.val) / (r2 * r2);
    r.val = std::atan2(y.val, x.val);
    for (int i = 0; i < K; ++i) r.g[i] = ty * y.g[i] + tx * x.g[i];
    for (int i = 0; i < K; ++i)
        for (int j = 0; j < K; ++j)
            r.h[i * K + j] = ty * y.h[i * K + j] + tx * x.h[i * K + j] + tyy * y.g[i] * y.g[j] + txx * x.g[i] * x.g[j] +
                             txy * (x.g[i] * y.g[j] + y.g[i] * x.g[j]);
    return r;
}
inline double valueOf(double a) { return a; }
template <int K>
double valueOf(const AD<K> &a) { return a.val; }

// ---------------------------------------------------------------------------------------------------------------------
// The fixed-size vector primitives the energies use (Eigen: cross, dot, squaredNorm, norm = sqrt(squaredNorm),
// normalized with the z > 0 guard of external/eigen/Eigen/src/Core/Dot.h:92-102).
// ---------------------------------------------------------------------------------------------------------------------
template <class T>
struct V3 {
    T x, y, z;
};
template <class T> V3<T> operator-(const V3<T> &a, const V3<T> &b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
template <class T> V3<T> operator+(const V3<T> &a, const V3<T> &b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
template <class T> V3<T> operator-(const V3<T> &a) { return {-a.x, -a.y, -a.z}; }
template <class T> V3<T> scale(const V3<T> &a, double s) { return {a.x * s, a.y * s, a.z * s}; }
template <class T> T dot(const V3<T> &a, const V3<T> &b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <class T>
V3<T> cross(const V3<T> &a, const V3<T> &b)
{
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
using std::sqrt;
template <class T> T norm(const V3<T> &a) { return sqrt(dot(a, a)); }
template <class T>
V3<T> normalized(const V3<T> &a)
{
    const T z = dot(a, a);
    if (valueOf(z) > 0) {
        const T n = sqrt(z);
        return {a.x / n, a.y / n, a.z / n};
    }
    return a;
}
using Vec3 = V3<double>;

// ---------------------------------------------------------------------------------------------------------------------
// Small dense helpers: n x n inverse by cofactors (n <= 4; Eigen's fixed-size inverse is the closed form too:
// external/eigen/Eigen/src/LU/InverseImpl.h:131-142 for 3 x 3) and the 3 x 3 determinant (LU/Determinant.h:47-51).
// ---------------------------------------------------------------------------------------------------------------------
inline double detN(const double *m, int n, int stride)
{
    if (n == 1) return m[0];
    if (n == 2) return m[0] * m[stride + 1] - m[1] * m[stride];
    double d = 0;
    double minor[9];
    for (int c = 0; c < n; ++c) {
        for (int i = 1; i < n; ++i) {
            int cc = 0;
            for (int j = 0; j < n; ++j)
                if (j != c) minor[(i - 1) * (n - 1) + cc++] = m[i * stride + j];
        }
        d += ((c % 2) ? -1.0 : 1.0) * m[c] * detN(minor, n - 1, n - 1);
    }
    return d;
}
inline void inverseN(const double *m, int n, double *out)
{
    if (n == 1) {
        out[0] = 1.0 / m[0];
        return;
    }
    const double d = detN(m, n, n);
    double minor[9];
    for (int r = 0; r < n; ++r)
        for (int c = 0; c < n; ++c) {
            int k = 0;
            for (int i = 0; i < n; ++i)
                for (int j = 0; j < n; ++j)
                    if (i != r && j != c) minor[k++] = m[i * n + j];
            const double cof = (((r + c) % 2) ? -1.0 : 1.0) * detN(minor, n - 1, n - 1);
            out[c * n + r] = cof / d;  // adjugate = transposed cofactors
        }
}

// ---------------------------------------------------------------------------------------------------------------------
// One energy = one reference class instance: stencils + the per-stencil members the class keeps.
// ---------------------------------------------------------------------------------------------------------------------
struct Energy {
    int kind = 0, S = 0, NC = 0, n = 0;
    bool prepared = false;
    std::vector<int> stencils;                       // S per stencil
    std::vector<double> Kmodel, Keff, Amodel, Aeff;  // NC x NC per stencil, row-major (energy.h: m_modelStiffness, ...)
    std::vector<double> restLength, restTheta, stiffnessFactor, alpha;  // per stencil
    std::vector<double> laplacian;                   // quadratic bending: c0..c3 per stencil (L = [c0 I c1 I c2 I c3 I])
    std::vector<double> DmInv;                       // 2 x 2 (membranes) or 3 x 3 (tets) per stencil, row-major
    std::vector<double> anchor;                      // 3 per stencil
    const int *st(int i) const { return stencils.data() + static_cast<size_t>(i) * S; }
    double *K(std::vector<double> &m, int i) { return m.data() + static_cast<size_t>(i) * NC * NC; }
    const double *K(const std::vector<double> &m, int i) const { return m.data() + static_cast<size_t>(i) * NC * NC; }
};

struct Group {  // spg::ParticleGroup (spg/source/sim/simObject/particleGroup.cpp)
    int n = 0;
    std::vector<double> x, x0, xInitial, v, mass, invMass;  // 3n / n
    std::vector<std::shared_ptr<Energy>> energies;            // shared between copies (particleGroup.h:35, :59)
};

inline void pushDiag(Energy &e, const std::vector<double> &diag)
{
    const int nc = e.NC;
    std::vector<double> Kmat(static_cast<size_t>(nc) * nc, 0.0), A(static_cast<size_t>(nc) * nc, 0.0);
    for (int i = 0; i < nc; ++i) Kmat[i * nc + i] = diag[i];
    inverseN(Kmat.data(), nc, A.data());
    e.Kmodel.insert(e.Kmodel.end(), Kmat.begin(), Kmat.end());
    e.Keff.insert(e.Keff.end(), Kmat.begin(), Kmat.end());
    e.Amodel.insert(e.Amodel.end(), A.begin(), A.end());
    e.Aeff.insert(e.Aeff.end(), A.begin(), A.end());
}
inline void pushMatrix(Energy &e, const std::vector<double> &Kmat)
{
    const int nc = e.NC;
    std::vector<double> A(static_cast<size_t>(nc) * nc, 0.0);
    inverseN(Kmat.data(), nc, A.data());
    e.Kmodel.insert(e.Kmodel.end(), Kmat.begin(), Kmat.end());
    e.Keff.insert(e.Keff.end(), Kmat.begin(), Kmat.end());
    e.Amodel.insert(e.Amodel.end(), A.begin(), A.end());
    e.Aeff.insert(e.Aeff.end(), A.begin(), A.end());
}

// <Class>::addStencil of every kind; p = the model parameters of one stencil in the order oracle/ref_driver.cpp passes
// them to the reference (makeEnergy)
inline void addStencil(Energy &e, const int32_t *st, const double *p)
{
    for (int s = 0; s < e.S; ++s) e.stencils.push_back(st[s]);
    switch (e.kind) {
    case K_SPRING:  // springEnergy.cpp:33-46
        if (p[0] <= 0) throw std::runtime_error("Zero/negative rest length invalid");
        e.restLength.push_back(p[0]);
        pushDiag(e, {p[1]});
        break;
    case K_SPRING_CONTINUUM: {  // springContinuumEnergy.cpp:19-33: effective stiffness = k * L0
        if (p[0] <= 0) throw std::runtime_error("Zero/negative rest length invalid");
        e.restLength.push_back(p[0]);
        pushDiag(e, {p[1]});
        double *Ke = e.K(e.Keff, e.n), *Ae = e.K(e.Aeff, e.n);
        Ke[0] = p[1] * p[0];
        Ae[0] = 1.0 / Ke[0];
        break;
    }
    case K_DISCRETE_BENDING:  // discreteBendingEnergy.cpp:33-46
        if (p[0] < -3.12 || p[0] > 3.12) throw std::runtime_error("Invalid restTheta");
        e.restTheta.push_back(p[0]);
        pushDiag(e, {p[1]});
        break;
    case K_QUADRATIC_BENDING:  // quadraticBendingEnergy.cpp:32-47 (matrices are filled by preparePrecomputations)
        if (p[0] != 0) throw std::runtime_error("Rest theta != 0 not supported for quadratic bending");
        e.stiffnessFactor.push_back(p[1]);
        e.laplacian.insert(e.laplacian.end(), 4, 0.0);
        for (auto *m : {&e.Kmodel, &e.Keff, &e.Amodel, &e.Aeff}) m->insert(m->end(), 9, 0.0);
        break;
    case K_MEMBRANE_STVK: {  // membraneStvkEnergy.cpp:31-43
        const double young = p[0], poisson = p[1];
        const double yc = young / (1 - poisson * poisson);
        pushMatrix(e, {yc, yc * poisson, 0, yc * poisson, yc, 0, 0, 0, yc * (1 - poisson) / 2});
        break;
    }
    case K_MEMBRANE_CHOI:  // membraneChoiEnergy.cpp:31-43
        pushDiag(e, {p[0], p[1], p[2], p[2]});
        break;
    case K_MEMBRANE_BW:  // membraneBaraffWitkinEnergy.cpp:29-41
        pushDiag(e, {p[0], p[1], p[2]});
        break;
    case K_TET_STVK: {  // stvkEnergy.cpp:68-82
        const double young = p[0], poisson = p[1];
        const double mu = young / (2 * (1 + poisson));
        const double lambda = young * poisson / ((1 + poisson) * (1 - 2 * poisson)) + 1e-20;
        pushDiag(e, {lambda, mu});
        break;
    }
    case K_TET_SNH: {  // stableNeoHookeanEnergy.cpp:64-80
        const double young = p[0], poisson = p[1];
        const double mu = young / (2 * (1 + poisson));
        const double lambda = mu / (1 - 2 * poisson);
        pushDiag(e, {lambda, mu});
        e.alpha.push_back(1 + mu / lambda);
        break;
    }
    case K_ANCHOR:  // springAnchorEnergy.cpp:29-39
        pushDiag(e, {p[3]});
        e.anchor.insert(e.anchor.end(), p, p + 3);
        break;
    case K_BENDING_BW:  // baraffWitkinBendingEnergy.cpp:40-53
        if (p[0] < -3.12 || p[0] > 3.12) throw std::runtime_error("Invalid restTheta");
        e.restTheta.push_back(p[0]);
        pushDiag(e, {p[1]});
        break;
    case K_SPRING_SQUARED_CONTINUUM: {  // springSquaredContinuumEnergy.cpp:20-34
        if (p[0] <= 0) throw std::runtime_error("Zero/negative rest length invalid");
        e.restLength.push_back(p[0]);
        pushDiag(e, {p[1]});
        e.K(e.Keff, e.n)[0] = p[1] * p[0];
        e.K(e.Aeff, e.n)[0] = 1.0 / (p[1] * p[0]);
        break;
    }
    case K_TET_SSNH: {  // stableSquaredNeoHookeanEnergy.cpp:34-50
        const double young = p[0], poisson = p[1];
        const double mu = young / (2 * (1 + poisson));
        const double lambda = young * poisson / ((1 + poisson) * (1 - 2 * poisson)) + 1e-20;
        pushDiag(e, {lambda, mu});
        break;
    }
    case K_STATIC_LOAD: {  // staticLoadEnergy.h:11-19: direction stored normalised; `anchor` holds it
        pushDiag(e, {p[3]});
        const Vec3 d = normalized(Vec3{p[0], p[1], p[2]});
        e.anchor.push_back(d.x), e.anchor.push_back(d.y), e.anchor.push_back(d.z);
        break;
    }
    default: throw std::runtime_error("unknown energy kind");
    }
    e.n++;
}

inline Vec3 pos(const std::vector<double> &x, int v) { return {x[3 * v], x[3 * v + 1], x[3 * v + 2]}; }

// <Class>::preparePrecomputations on the rest positions; energy.h:54-61 allows it once
inline void prepare(Energy &e, const Group &g)
{
    if (e.prepared) throw std::runtime_error("Energy was already prepared");
    const int nc = e.NC;
    auto scaleEffective = [&](int i, double measure) {  // compliance / measure, stiffness * measure
        for (int k = 0; k < nc * nc; ++k) {
            e.K(e.Aeff, i)[k] = e.K(e.Amodel, i)[k] / measure;
            e.K(e.Keff, i)[k] = e.K(e.Kmodel, i)[k] * measure;
        }
    };
    for (int i = 0; i < e.n; ++i) {
        const int *s = e.st(i);
        switch (e.kind) {
        case K_DISCRETE_BENDING: {  // discreteBendingEnergy.cpp:48-72
            const Vec3 x0 = pos(g.x0, s[0]), x1 = pos(g.x0, s[1]), x2 = pos(g.x0, s[2]), x3 = pos(g.x0, s[3]);
            const Vec3 e0 = x1 - x0, e3 = x2 - x1, e4 = x3 - x1;
            const double factor = 2. * 3. * dot(e0, e0) / (0.5 * (norm(cross(e0, e3)) + norm(cross(e0, e4))));
            e.K(e.Aeff, i)[0] = e.K(e.Amodel, i)[0] / factor;
            e.K(e.Keff, i)[0] = 1.0 / e.K(e.Aeff, i)[0];
            const Vec3 n1 = normalized(cross(e0, e3)), n2 = -normalized(cross(e0, e4));
            e.restTheta[i] = std::atan2(dot(cross(n1, n2), normalized(e0)), dot(n1, n2));
            break;
        }
        case K_QUADRATIC_BENDING: {  // quadraticBendingEnergy.cpp:49-93
            const Vec3 x0 = pos(g.x0, s[0]), x1 = pos(g.x0, s[1]), x2 = pos(g.x0, s[2]), x3 = pos(g.x0, s[3]);
            const Vec3 e0 = x1 - x0, e1 = x2 - x0, e2 = x3 - x0, e3 = x2 - x1, e4 = x3 - x1;
            auto cot = [](const Vec3 &a, const Vec3 &b) { return dot(a, b) / norm(cross(a, b)); };
            const double c01 = cot(e0, e1), c02 = cot(e0, e2), c03 = cot(-e0, e3), c04 = cot(-e0, e4);
            double *L = e.laplacian.data() + 4 * static_cast<size_t>(i);
            L[0] = c03 + c04, L[1] = c01 + c02, L[2] = -c01 - c03, L[3] = -c02 - c04;
            const double m = e.stiffnessFactor[i] * (3 / (0.5 * (norm(cross(e0, e1)) + norm(cross(e0, e2)))));
            for (int k = 0; k < 9; ++k) e.K(e.Keff, i)[k] = e.K(e.Kmodel, i)[k] = e.K(e.Aeff, i)[k] = e.K(e.Amodel, i)[k] = 0;
            for (int k = 0; k < 3; ++k) {
                e.K(e.Keff, i)[4 * k] = e.K(e.Kmodel, i)[4 * k] = m;
                e.K(e.Aeff, i)[4 * k] = e.K(e.Amodel, i)[4 * k] = 1.0 / m;
            }
            break;
        }
        case K_MEMBRANE_STVK:
        case K_MEMBRANE_CHOI:
        case K_MEMBRANE_BW: {  // membraneStvkEnergy.cpp:45-71 (Choi :45-71, Baraff-Witkin :43-69 are the same code)
            const Vec3 x0 = pos(g.x0, s[0]), x1 = pos(g.x0, s[1]), x2 = pos(g.x0, s[2]);
            const char *name = e.kind == K_MEMBRANE_STVK ? "StVK" : (e.kind == K_MEMBRANE_CHOI ? "Choi" : "BaraffWitkin");
            if (x0.z != 0.0 || x1.z != 0.0 || x2.z != 0.0)
                throw std::runtime_error(std::string("Rest positions of triangles for membrane ") + name +
                                         " energy must have coordinate z=0");
            const double ux = x1.x - x0.x, uy = x1.y - x0.y, vx = x2.x - x0.x, vy = x2.y - x0.y;
            const double Dm[4] = {ux, vx, uy, vy};  // columns u, v
            if (e.DmInv.size() < 4 * static_cast<size_t>(e.n)) e.DmInv.resize(4 * static_cast<size_t>(e.n));
            inverseN(Dm, 2, e.DmInv.data() + 4 * static_cast<size_t>(i));
            const double area = (ux * vy - vx * uy) * 0.5;
            if (area <= 0) throw std::runtime_error(std::string("Rest area of membrane ") + name + " triangle is not positive");
            scaleEffective(i, area);
            break;
        }
        case K_BENDING_BW: {  // baraffWitkinBendingEnergy.cpp:55-77: no area factor, rest angle from the rest pose
            const Vec3 x0 = pos(g.x0, s[0]), x1 = pos(g.x0, s[1]), x2 = pos(g.x0, s[2]), x3 = pos(g.x0, s[3]);
            const Vec3 e0 = x1 - x0, e3 = x2 - x1, e4 = x3 - x1;
            const Vec3 n1 = normalized(cross(e0, e3)), n2 = -normalized(cross(e0, e4));
            e.restTheta[i] = std::atan2(dot(cross(n1, n2), normalized(e0)), dot(n1, n2));
            break;
        }
        case K_TET_SSNH:  // stableSquaredNeoHookeanEnergy.cpp:52-78
        case K_TET_STVK:
        case K_TET_SNH: {  // stvkEnergy.cpp:84-110, stableNeoHookeanEnergy.cpp:82-108
            const Vec3 x0 = pos(g.x0, s[0]), x1 = pos(g.x0, s[1]), x2 = pos(g.x0, s[2]), x3 = pos(g.x0, s[3]);
            const Vec3 u = x1 - x0, v = x2 - x0, w = x3 - x0;
            const double Dm[9] = {u.x, v.x, w.x, u.y, v.y, w.y, u.z, v.z, w.z};
            const double volume = 1. / 6. * detN(Dm, 3, 3);
            if (volume <= 0)
                throw std::runtime_error(std::string(e.kind == K_TET_SSNH ? "SSNH" : "SNH") + " tetrahedron rest volume is not positive");
            if (e.DmInv.size() < 9 * static_cast<size_t>(e.n)) e.DmInv.resize(9 * static_cast<size_t>(e.n));
            inverseN(Dm, 3, e.DmInv.data() + 9 * static_cast<size_t>(i));
            scaleEffective(i, volume);
            break;
        }
        default: break;  // springs, anchors: nothing to precompute
        }
    }
    e.prepared = true;
}

// ---------------------------------------------------------------------------------------------------------------------
// Constraint bodies C(x) of every kind, written once over the scalar type like the reference's generic lambdas.
// xs: the S stencil positions as scalars (DOF s * 3 + c active).  C: NC outputs.
// ---------------------------------------------------------------------------------------------------------------------
template <class T>
void deformationGradient32(const V3<T> *xs, const double *Di, V3<T> &F0, V3<T> &F1)
{
    // F = [u v] * Dm^-1 (3 x 2) -- membraneStvkEnergy.cpp:19-22
    const V3<T> u = xs[1] - xs[0], v = xs[2] - xs[0];
    F0 = scale(u, Di[0]) + scale(v, Di[2]);
    F1 = scale(u, Di[1]) + scale(v, Di[3]);
}
template <class T>
void deformationGradient33(const V3<T> *xs, const double *Di, V3<T> F[3])
{
    // F = [u v w] * Dm^-1 (3 x 3), F[c] = column c -- stvkEnergy.cpp:21-25
    const V3<T> u = xs[1] - xs[0], v = xs[2] - xs[0], w = xs[3] - xs[0];
    for (int c = 0; c < 3; ++c) F[c] = scale(u, Di[c]) + scale(v, Di[3 + c]) + scale(w, Di[6 + c]);
}
template <class T>
T det3(const V3<T> F[3])  // Eigen 3 x 3 determinant of the matrix with columns F[0..2]
{
    return dot(F[0], cross(F[1], F[2]));
}

template <class T>
void constraints(const Energy &e, int i, const V3<T> *xs, T *C)
{
    switch (e.kind) {
    case K_SPRING:  // springEnergy.cpp:10-18
        C[0] = norm(xs[0] - xs[1]) - e.restLength[i];
        break;
    case K_SPRING_CONTINUUM:  // springContinuumEnergy.cpp:8-16
        C[0] = norm(xs[0] - xs[1]) / e.restLength[i] - 1.;
        break;
    case K_DISCRETE_BENDING: {  // discreteBendingEnergy.cpp:9-30
        const V3<T> e0 = xs[1] - xs[0], e3 = xs[2] - xs[1], e4 = xs[3] - xs[1];
        const V3<T> n1 = normalized(cross(e0, e3));
        const V3<T> n2 = -normalized(cross(e0, e4));
        const T theta = atan2(dot(cross(n1, n2), normalized(e0)), dot(n1, n2));
        C[0] = theta - e.restTheta[i];
        break;
    }
    case K_QUADRATIC_BENDING: {  // quadraticBendingEnergy.cpp:9-29: C = L x, L = [c0 I c1 I c2 I c3 I]
        const double *L = e.laplacian.data() + 4 * static_cast<size_t>(i);
        const V3<T> c = scale(xs[0], L[0]) + scale(xs[1], L[1]) + scale(xs[2], L[2]) + scale(xs[3], L[3]);
        C[0] = c.x, C[1] = c.y, C[2] = c.z;
        break;
    }
    case K_MEMBRANE_STVK: {  // membraneStvkEnergy.cpp:9-28: Green strain E = (F^T F - I) / 2, C = (E00, E11, 2 E01)
        V3<T> F0, F1;
        deformationGradient32(xs, e.DmInv.data() + 4 * static_cast<size_t>(i), F0, F1);
        C[0] = (dot(F0, F0) - 1.0) * 0.5;
        C[1] = (dot(F1, F1) - 1.0) * 0.5;
        C[2] = 2.0 * (dot(F0, F1) * 0.5);
        break;
    }
    case K_MEMBRANE_CHOI: {  // membraneChoiEnergy.cpp:9-28
        V3<T> F0, F1;
        deformationGradient32(xs, e.DmInv.data() + 4 * static_cast<size_t>(i), F0, F1);
        const double sqrtTwoInv = 1.0 / std::sqrt(2.0);
        C[0] = norm(F0) - 1.0;
        C[1] = norm(F1) - 1.0;
        C[2] = norm(scale(F0 + F1, sqrtTwoInv)) - 1.0;
        C[3] = norm(scale(F0 - F1, sqrtTwoInv)) - 1.0;
        break;
    }
    case K_MEMBRANE_BW: {  // membraneBaraffWitkinEnergy.cpp:9-26
        V3<T> F0, F1;
        deformationGradient32(xs, e.DmInv.data() + 4 * static_cast<size_t>(i), F0, F1);
        C[0] = norm(F0) - 1.0;
        C[1] = norm(F1) - 1.0;
        C[2] = dot(F0, F1);
        break;
    }
    case K_TET_STVK: {  // stvkEnergy.cpp:9-39: C = (tr E, sqrt(2 tr(E E))), zero guard :36-38
        V3<T> F[3];
        deformationGradient33(xs, e.DmInv.data() + 9 * static_cast<size_t>(i), F);
        T E[3][3];
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) E[a][b] = (dot(F[a], F[b]) - (a == b ? 1.0 : 0.0)) * 0.5;
        C[0] = E[0][0] + E[1][1] + E[2][2];
        T trEE = T(0.0);
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) trEE = trEE + E[a][b] * E[b][a];
        C[1] = sqrt(2.0 * trEE);
        if (valueOf(C[1]) == 0) C[1] = T(0.0);
        break;
    }
    case K_TET_SNH: {  // stableNeoHookeanEnergy.cpp:9-31: C = (det F - alpha, sqrt(tr(F^T F)))
        V3<T> F[3];
        deformationGradient33(xs, e.DmInv.data() + 9 * static_cast<size_t>(i), F);
        C[0] = det3(F) - e.alpha[i];
        C[1] = sqrt(dot(F[0], F[0]) + dot(F[1], F[1]) + dot(F[2], F[2]));
        break;
    }
    case K_BENDING_BW: {  // baraffWitkinBendingEnergy.cpp:9-37: the three norms are passive values (:25-31)
        const V3<T> e0 = xs[1] - xs[0], e3 = xs[2] - xs[1], e4 = xs[3] - xs[1];
        const double e0NormInv = 1 / valueOf(norm(e0));
        const V3<T> n1 = cross(e0, e3);
        const double n1NormInv = 1 / valueOf(norm(n1));
        const V3<T> n2 = -cross(e0, e4);
        const double n2NormInv = 1 / valueOf(norm(n2));
        const V3<T> a = scale(n1, n1NormInv), b = scale(n2, n2NormInv);
        C[0] = atan2(dot(cross(a, b), scale(e0, e0NormInv)), dot(a, b)) - e.restTheta[i];
        break;
    }
    case K_SPRING_SQUARED_CONTINUUM: {  // springSquaredContinuumEnergy.cpp:8-17
        const V3<T> d = xs[0] - xs[1];
        C[0] = dot(d, d) / (e.restLength[i] * e.restLength[i]) - 1.;
        break;
    }
    case K_TET_SSNH: {  // stableSquaredNeoHookeanEnergy.cpp:9-31
        V3<T> F[3];
        deformationGradient33(xs, e.DmInv.data() + 9 * static_cast<size_t>(i), F);
        C[0] = det3(F) - 1.0;
        C[1] = dot(F[0], F[0]) + dot(F[1], F[1]) + dot(F[2], F[2]) - 3.0;
        break;
    }
    case K_STATIC_LOAD:  // staticLoadEnergy.cpp:8-13
        throw std::runtime_error("Constraint-based load energy not implemented");
    case K_ANCHOR: {  // springAnchorEnergy.cpp:8-19 (zero guard :16-18)
        const double *a = e.anchor.data() + 3 * static_cast<size_t>(i);
        const V3<T> d = {xs[0].x - a[0], xs[0].y - a[1], xs[0].z - a[2]};
        C[0] = norm(d);
        if (valueOf(C[0]) == 0) C[0] = T(0.0);
        break;
    }
    default: throw std::runtime_error("unknown energy kind");
    }
}

// dEnergy: the default E = 1/2 C^T K_eff C (energy.h:336-347) or the class's own direct form
template <class T>
T energyOf(const Energy &e, int i, const V3<T> *xs)
{
    switch (e.kind) {
    case K_SPRING: {  // springEnergy.cpp:20-30
        const T diff = norm(xs[0] - xs[1]) - e.restLength[i];
        return 0.5 * e.K(e.Keff, i)[0] * diff * diff;
    }
    case K_TET_STVK: {  // stvkEnergy.cpp:41-65: E = 1/2 lambda (tr E)^2 + mu tr(E^T E)
        V3<T> F[3];
        deformationGradient33(xs, e.DmInv.data() + 9 * static_cast<size_t>(i), F);
        T E[3][3];
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) E[a][b] = (dot(F[a], F[b]) - (a == b ? 1.0 : 0.0)) * 0.5;
        const double lambda = e.K(e.Keff, i)[0], mu = e.K(e.Keff, i)[3];
        const T tr = E[0][0] + E[1][1] + E[2][2];
        T trEtE = T(0.0);
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) trEtE = trEtE + E[a][b] * E[a][b];
        return 0.5 * lambda * tr * tr + mu * trEtE;
    }
    case K_TET_SNH: {  // stableNeoHookeanEnergy.cpp:33-61
        V3<T> F[3];
        deformationGradient33(xs, e.DmInv.data() + 9 * static_cast<size_t>(i), F);
        const double lambda = e.K(e.Keff, i)[0], mu = e.K(e.Keff, i)[3];
        const T hyd = det3(F) - e.alpha[i];
        const T dev = dot(F[0], F[0]) + dot(F[1], F[1]) + dot(F[2], F[2]) - 3.0;
        return 0.5 * lambda * hyd * hyd + 0.5 * mu * dev;
    }
    case K_STATIC_LOAD: {  // staticLoadEnergy.cpp:15-20: E = k * -(d . x)
        const double *d = e.anchor.data() + 3 * static_cast<size_t>(i);
        return e.K(e.Kmodel, i)[0] * -(d[0] * xs[0].x + d[1] * xs[0].y + d[2] * xs[0].z);
    }
    case K_ANCHOR: {  // springAnchorEnergy.cpp:21-26: the MODEL stiffness
        const double *a = e.anchor.data() + 3 * static_cast<size_t>(i);
        const V3<T> d = {xs[0].x - a[0], xs[0].y - a[1], xs[0].z - a[2]};
        return 0.5 * e.K(e.Kmodel, i)[0] * dot(d, d);
    }
    default: {
        T C[4];
        constraints(e, i, xs, C);
        const double *Kmat = e.K(e.Keff, i);
        T acc = T(0.0);
        for (int a = 0; a < e.NC; ++a) {
            T row = T(0.0);
            for (int b = 0; b < e.NC; ++b) row = row + Kmat[a * e.NC + b] * C[b];
            acc = acc + C[a] * row;
        }
        return 0.5 * acc;
    }
    }
}

template <int S>
void seed(const Energy &e, int i, const std::vector<double> &x, V3<AD<3 * S>> *xs)
{
    for (int s = 0; s < S; ++s) {
        const int v = e.st(i)[s];
        xs[s] = {AD<3 * S>(x[3 * v], 3 * s), AD<3 * S>(x[3 * v + 1], 3 * s + 1), AD<3 * S>(x[3 * v + 2], 3 * s + 2)};
    }
}

// value, gradient (k) and Hessian (k x k row-major) of stencil i: energy.h:336-362, with SpringEnergy's hand-written
// overrides (springEnergy.cpp:90-134), which is what VBD reaches through the virtual energyGradient / energyHessian
template <int S>
void energyDerivatives(const Energy &e, int i, const std::vector<double> &x, double *val, double *grad, double *hess)
{
    constexpr int K = 3 * S;
    if (e.kind == K_SPRING) {
        const Vec3 x0 = pos(x, e.st(i)[0]), x1 = pos(x, e.st(i)[1]);
        const double k = e.K(e.Keff, i)[0], L0 = e.restLength[i];
        const Vec3 s = x0 - x1;
        const double sNorm = norm(s);
        if (val) *val = 0.5 * k * (sNorm - L0) * (sNorm - L0);
        const double sv[3] = {s.x, s.y, s.z};
        if (grad)
            for (int c = 0; c < 3; ++c) {
                grad[c] = k * (sNorm - L0) * sv[c] / sNorm;
                grad[3 + c] = -grad[c];
            }
        if (hess)
            for (int a = 0; a < 3; ++a)
                for (int b = 0; b < 3; ++b) {
                    const double sst = (sv[a] / sNorm) * (sv[b] / sNorm);
                    const double h = k * ((1 - L0 / sNorm) * ((a == b ? 1.0 : 0.0) - sst) + sst);
                    hess[a * 6 + b] = hess[(3 + a) * 6 + 3 + b] = h;
                    hess[a * 6 + 3 + b] = hess[(3 + a) * 6 + b] = -h;
                }
        return;
    }
    V3<AD<K>> xs[S];
    seed<S>(e, i, x, xs);
    const AD<K> E = energyOf(e, i, xs);
    if (val) *val = E.val;
    if (grad) std::copy(E.g, E.g + K, grad);
    if (hess) std::copy(E.h, E.h + K * K, hess);
}

inline void energyDerivativesAny(const Energy &e, int i, const std::vector<double> &x, double *val, double *grad, double *hess)
{
    switch (e.S) {
    case 1: energyDerivatives<1>(e, i, x, val, grad, hess); break;
    case 2: energyDerivatives<2>(e, i, x, val, grad, hess); break;
    case 3: energyDerivatives<3>(e, i, x, val, grad, hess); break;
    default: energyDerivatives<4>(e, i, x, val, grad, hess); break;
    }
}

// constraintsAndGradient (energy.h:317-332): C (NC) and grad (k x NC row-major)
template <int S>
void constraintsAndGradient(const Energy &e, int i, const std::vector<double> &x, double *C, double *grad)
{
    constexpr int K = 3 * S;
    V3<AD<K>> xs[S];
    seed<S>(e, i, x, xs);
    AD<K> c[4];
    constraints(e, i, xs, c);
    for (int j = 0; j < e.NC; ++j) {
        C[j] = c[j].val;
        for (int k = 0; k < K; ++k) grad[k * e.NC + j] = c[j].g[k];
    }
}
inline void constraintsAndGradientAny(const Energy &e, int i, const std::vector<double> &x, double *C, double *grad)
{
    switch (e.S) {
    case 1: constraintsAndGradient<1>(e, i, x, C, grad); break;
    case 2: constraintsAndGradient<2>(e, i, x, C, grad); break;
    case 3: constraintsAndGradient<3>(e, i, x, C, grad); break;
    default: constraintsAndGradient<4>(e, i, x, C, grad); break;
    }
}

// projectPosition: the generic block-XPBD step (energy.h:117-153) or SpringEnergy's own (springEnergy.cpp:48-68)
inline void projectPosition(const Energy &e, int i, Group &g, double dt)
{
    if (e.kind == K_SPRING) {
        const int a = e.st(i)[0], b = e.st(i)[1];
        const double w0 = g.invMass[a], w1 = g.invMass[b], wSum = w0 + w1;
        if (wSum < 1.0e-6) return;
        const double alphaTilde = e.K(e.Aeff, i)[0] / (dt * dt);
        Vec3 n = pos(g.x, b) - pos(g.x, a);
        const double d = norm(n);
        if (d < 1.0e-6) return;
        n = {n.x / d, n.y / d, n.z / d};
        const double c = (d - e.restLength[i]) / (wSum + alphaTilde);
        const double corr[3] = {n.x * c, n.y * c, n.z * c};
        for (int k = 0; k < 3; ++k) {
            g.x[3 * a + k] += w0 * corr[k];
            g.x[3 * b + k] -= w1 * corr[k];
        }
        return;
    }
    const int nc = e.NC, K = 3 * e.S;
    double C[4], grad[12 * 4], Wgrad[12 * 4], A[16], Ainv[16], dl[4];
    constraintsAndGradientAny(e, i, g.x, C, grad);
    for (int k = 0; k < K; ++k) {
        const double w = g.invMass[e.st(i)[k / 3]];
        for (int j = 0; j < nc; ++j) Wgrad[k * nc + j] = w * grad[k * nc + j];
    }
    const double *alpha = e.K(e.Aeff, i);
    for (int a = 0; a < nc; ++a)
        for (int b = 0; b < nc; ++b) {
            double s = 0;
            for (int k = 0; k < K; ++k) s += grad[k * nc + a] * Wgrad[k * nc + b];
            A[a * nc + b] = s + alpha[a * nc + b] / (dt * dt) + (a == b ? 1e-10 : 0.0);
        }
    inverseN(A, nc, Ainv);
    for (int a = 0; a < nc; ++a) {
        double s = 0;
        for (int b = 0; b < nc; ++b) s += Ainv[a * nc + b] * C[b];
        dl[a] = -s;
    }
    for (int k = 0; k < K; ++k) {
        double s = 0;
        for (int j = 0; j < nc; ++j) s += Wgrad[k * nc + j] * dl[j];
        g.x[3 * e.st(i)[k / 3] + k % 3] += s;  // particleGroup.cpp:43-46
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Greedy colouring (spg/source/utils/graphColoring.cpp).  Graph::greedyColoring (:21-49) visits nodes in index order and
// gives each the smallest colour no ALREADY COLOURED neighbour holds; node 0 gets colour 0.  The reference refills an
// N-sized availability array per node (quadratic); a stamp array gives the same colours in O(N deg).
// ---------------------------------------------------------------------------------------------------------------------
inline std::vector<int> greedyColoring(const std::vector<std::vector<int>> &adjacency)
{
    const int n = static_cast<int>(adjacency.size());
    std::vector<int> colors(static_cast<size_t>(n), -1), stamp(static_cast<size_t>(n) + 1, -1);
    for (int v = 0; v < n; ++v) {
        for (int u : adjacency[v])
            if (colors[u] != -1) stamp[colors[u]] = v;
        int c = 0;
        while (stamp[c] == v) ++c;
        colors[v] = c;
    }
    return colors;
}
// coloring::colorVertices (:72-84): vertices adjacent iff they share a stencil of any energy
inline std::vector<int> colorVertices(int nVerts, const std::vector<std::pair<const int *, std::pair<int64_t, int>>> &sets)
{
    std::vector<std::vector<int>> adj(static_cast<size_t>(nVerts));
    for (const auto &fs : sets) {
        const int *ent = fs.first;
        const int64_t nEnt = fs.second.first;
        const int S = fs.second.second;
        for (int64_t i = 0; i < nEnt; ++i)
            for (int64_t j = i + 1; j % S != 0; ++j) {
                if (ent[i] >= nVerts || ent[j] >= nVerts)
                    throw std::runtime_error("Trying to add edge to graph with invalid node indices");
                adj[ent[i]].push_back(ent[j]);
                adj[ent[j]].push_back(ent[i]);
            }
    }
    return greedyColoring(adj);
}
// coloring::colorStencils (:51-70): stencils adjacent iff they share a vertex
inline std::vector<int> colorStencils(const int *ent, int64_t nEnt, int S)
{
    std::vector<std::vector<int>> perVertex;
    for (int64_t i = 0; i < nEnt; ++i) {
        if (static_cast<size_t>(ent[i]) >= perVertex.size()) perVertex.resize(static_cast<size_t>(ent[i]) + 1);
        perVertex[ent[i]].push_back(static_cast<int>(i / S));
    }
    std::vector<std::vector<int>> adj(static_cast<size_t>(nEnt / S));
    for (const auto &vs : perVertex)
        for (size_t a = 0; a < vs.size(); ++a)
            for (size_t b = a + 1; b < vs.size(); ++b) {
                adj[vs[a]].push_back(vs[b]);
                adj[vs[b]].push_back(vs[a]);
            }
    return greedyColoring(adj);
}

// ---------------------------------------------------------------------------------------------------------------------
// Solvers.  BaseSolver::addObject copies the group (spg/source/sim/solver/baseSolver.cpp:10-14); energies stay shared.
// ---------------------------------------------------------------------------------------------------------------------
struct Solver {
    int kind = 0;  // 0 VBD, 1 XPBD parallel Gauss-Seidel, 2 XPBD serial, 3 implicit Euler Baraff-Witkin, 4 Newton dx,
                   // 5 Newton dv, 6 symplectic Euler, 7 BDF2, 8 static Newton, 9 quasi-static Newton, 10 Newton robust,
                   // 11 quasi-static Newton robust
    Group g;
    double dt = 0.01, gravity[3] = {0, -9.8, 0};
    int substeps = 1, iterations = 1;
    bool dirty = true;
    std::vector<std::vector<int>> vertexGroups;                         // vbd.cpp:203-229
    std::vector<std::vector<std::vector<std::array<int, 2>>>> incidence;  // [vertex][energy] -> (stencil, local), vbd.cpp:231-268
    std::vector<std::vector<std::vector<int>>> stencilGroups;           // per energy, xpbd.cpp:133-160
    std::vector<double> xOld, xInertial;
    std::vector<int> cgIterations;  // implicit-Euler family: one entry per linear solve of the last step call
    std::vector<double> xMinusOne, vMinusOne;  // BDF2 (bdf2.h:19-20)
    double currentDt = -1.;                    // the adaptive solvers (implicitEulerNewtonRobust.h:21-22)
    int consecutiveSuccessfulSteps = 0;
};

inline void groupBy(const std::vector<int> &colors, std::vector<std::vector<int>> &groups)
{
    groups.clear();
    for (int i = 0; i < static_cast<int>(colors.size()); ++i) {
        if (static_cast<size_t>(colors[i]) >= groups.size()) groups.resize(static_cast<size_t>(colors[i]) + 1);
        groups[colors[i]].push_back(i);
    }
}

inline void vbdPrecompute(Solver &s)
{
    std::vector<std::pair<const int *, std::pair<int64_t, int>>> sets;
    for (const auto &e : s.g.energies) sets.push_back({e->stencils.data(), {static_cast<int64_t>(e->stencils.size()), e->S}});
    if (s.vertexGroups.empty()) groupBy(colorVertices(s.g.n, sets), s.vertexGroups);
    s.incidence.assign(static_cast<size_t>(s.g.n), std::vector<std::vector<std::array<int, 2>>>(s.g.energies.size()));
    for (size_t ei = 0; ei < s.g.energies.size(); ++ei) {
        const Energy &e = *s.g.energies[ei];
        for (int i = 0; i < e.n; ++i)
            for (int l = 0; l < e.S; ++l) s.incidence[e.st(i)[l]][ei].push_back({i, l});
    }
    s.dirty = false;
}

inline double det3x3(const double *H)
{
    return H[0] * (H[4] * H[8] - H[5] * H[7]) - H[1] * (H[3] * H[8] - H[5] * H[6]) + H[2] * (H[3] * H[7] - H[4] * H[6]);
}

// predict (vbd.cpp:49-83 / xpbd.cpp:47-74): xOld <- x; v += dt g; x += v dt
inline void predict(Solver &s, double dt, bool keepInertial)
{
    Group &g = s.g;
    s.xOld = g.x;
    for (int i = 0; i < g.n; ++i)
        for (int c = 0; c < 3; ++c) {
            g.v[3 * i + c] += dt * s.gravity[c];
            g.x[3 * i + c] += g.v[3 * i + c] * dt;  // particleGroup.cpp:57-63
        }
    if (keepInertial) s.xInertial = g.x;
}
// particleGroup.cpp:65-72
inline void finalizeVelocities(Solver &s, double invdt)
{
    for (size_t k = 0; k < s.g.x.size(); ++k) s.g.v[k] = (s.g.x[k] - s.xOld[k]) * invdt;
}

// VBD::step (vbd.cpp:12-195)
inline void vbdStep(Solver &s)
{
    if (s.dirty) vbdPrecompute(s);
    Group &g = s.g;
    const double dt = s.dt / s.substeps, invdt = 1. / dt, invdtSquared = invdt * invdt;
    for (int sub = 0; sub < s.substeps; ++sub) {
        predict(s, dt, true);
        for (int iter = 0; iter < s.iterations; ++iter)
            for (const auto &group : s.vertexGroups) {
                const int count = static_cast<int>(group.size());
#pragma omp parallel for schedule(static)
                for (int q = 0; q < count; ++q) {  // l_vertexDescent, vbd.cpp:99-147
                    const int v = group[q];
                    const double mOverdtSquared = g.mass[v] * invdtSquared;
                    double f[3], H[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
                    for (int c = 0; c < 3; ++c) {
                        f[c] = -mOverdtSquared * (g.x[3 * v + c] - s.xInertial[3 * v + c]);
                        H[4 * c] = mOverdtSquared;
                    }
                    for (size_t ei = 0; ei < g.energies.size(); ++ei) {
                        const Energy &e = *g.energies[ei];
                        for (const auto &ent : s.incidence[v][ei]) {
                            double grad[12], hess[144];
                            energyDerivativesAny(e, ent[0], g.x, nullptr, grad, hess);  // energy.h:240-258
                            const int K = 3 * e.S, o = 3 * ent[1];
                            for (int a = 0; a < 3; ++a) {
                                f[a] -= grad[o + a];
                                for (int b = 0; b < 3; ++b) H[3 * a + b] += hess[(o + a) * K + o + b];
                            }
                        }
                    }
                    if (det3x3(H) > 1e-10) {  // vbd.cpp:139
                        double Hinv[9];
                        inverseN(H, 3, Hinv);
                        for (int a = 0; a < 3; ++a) g.x[3 * v + a] += Hinv[3 * a] * f[0] + Hinv[3 * a + 1] * f[1] + Hinv[3 * a + 2] * f[2];
                    }
                }
            }
        finalizeVelocities(s, invdt);
    }
}

// XPBD::step (xpbd.cpp:16-131).  iterations > 1 is not a reference feature: the pass is simply repeated.
inline void xpbdStep(Solver &s)
{
    Group &g = s.g;
    if (s.kind == 1 && s.dirty) {
        if (s.stencilGroups.size() != g.energies.size()) s.stencilGroups.assign(g.energies.size(), {});
        for (size_t ei = 0; ei < g.energies.size(); ++ei)
            if (s.stencilGroups[ei].empty()) {
                const Energy &e = *g.energies[ei];
                groupBy(colorStencils(e.stencils.data(), static_cast<int64_t>(e.stencils.size()), e.S), s.stencilGroups[ei]);
            }
        s.dirty = false;
    }
    const double dt = s.dt / s.substeps, invdt = 1 / dt;
    for (int sub = 0; sub < s.substeps; ++sub) {
        predict(s, dt, false);
        for (int it = 0; it < s.iterations; ++it)
            for (size_t ei = 0; ei < g.energies.size(); ++ei) {
                const Energy &e = *g.energies[ei];
                if (s.kind == 1) {
                    for (const auto &group : s.stencilGroups[ei]) {
                        const int count = static_cast<int>(group.size());
#pragma omp parallel for schedule(static)
                        for (int q = 0; q < count; ++q) projectPosition(e, group[q], g, dt);
                    }
                } else {
                    for (int i = 0; i < e.n; ++i) projectPosition(e, i, g, dt);
                }
            }
        finalizeVelocities(s, invdt);
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// The implicit-Euler family and symplectic Euler (spg/source/sim/solver/implicitEulerBase.cpp and friends).
// ---------------------------------------------------------------------------------------------------------------------
// A square sparse matrix as sorted COO -> CSR with duplicates summed (what SparseMatrix::setFromTriplets yields)
struct Csr {
    int n = 0;
    std::vector<int64_t> rowptr;
    std::vector<int> col;
    std::vector<double> val;
};
struct Trip {
    int r, c;
    double v;
};
inline Csr fromTriplets(int n, std::vector<Trip> &t)
{
    std::stable_sort(t.begin(), t.end(), [](const Trip &a, const Trip &b) { return a.r != b.r ? a.r < b.r : a.c < b.c; });
    Csr A;
    A.n = n;
    A.rowptr.assign(static_cast<size_t>(n) + 1, 0);
    for (size_t k = 0; k < t.size(); ++k) {
        if (k > 0 && t[k].r == t[k - 1].r && t[k].c == t[k - 1].c) {
            A.val.back() += t[k].v;
        } else {
            A.col.push_back(t[k].c);
            A.val.push_back(t[k].v);
            A.rowptr[t[k].r + 1]++;
        }
    }
    for (int r = 0; r < n; ++r) A.rowptr[r + 1] += A.rowptr[r];
    return A;
}
inline void spmv(const Csr &A, const std::vector<double> &x, std::vector<double> &y)
{
    y.assign(static_cast<size_t>(A.n), 0.0);
#pragma omp parallel for schedule(static)
    for (int r = 0; r < A.n; ++r) {
        double acc = 0;
        for (int64_t k = A.rowptr[r]; k < A.rowptr[r + 1]; ++k) acc += A.val[k] * x[A.col[k]];
        y[r] = acc;
    }
}
inline double dotv(const std::vector<double> &a, const std::vector<double> &b)
{
    double acc = 0;
    for (size_t i = 0; i < a.size(); ++i) acc += a[i] * b[i];
    return acc;
}

// getSystemForce (implicitEulerBase.cpp:102-143): gravity * mass, then accumulateForces (energy.h:196-216) of every
// stencil of every energy: f -= energyGradient
inline void systemForce(const Solver &s, std::vector<double> &f)
{
    const Group &g = s.g;
    f.assign(3 * static_cast<size_t>(g.n), 0.0);
    for (int i = 0; i < g.n; ++i)
        for (int c = 0; c < 3; ++c) f[3 * i + c] += s.gravity[c] * g.mass[i];
    for (const auto &ep : g.energies) {
        const Energy &e = *ep;
        for (int i = 0; i < e.n; ++i) {
            double grad[12];
            energyDerivativesAny(e, i, g.x, nullptr, grad, nullptr);
            for (int l = 0; l < e.S; ++l)
                for (int c = 0; c < 3; ++c) f[3 * e.st(i)[l] + c] += -grad[3 * l + c];
        }
    }
}
// getSystemStiffnessMatrix (:176-221): K = sum of negativeHessianTriplets (energy.h:218-237), i.e. minus the assembled
// energy Hessian
inline Csr systemStiffness(const Solver &s)
{
    const Group &g = s.g;
    std::vector<Trip> t;
    for (const auto &ep : g.energies) {
        const Energy &e = *ep;
        const int K = 3 * e.S;
        for (int i = 0; i < e.n; ++i) {
            double hess[144];
            energyDerivativesAny(e, i, g.x, nullptr, nullptr, hess);
            for (int rs = 0; rs < e.S; ++rs)
                for (int cs = 0; cs < e.S; ++cs)
                    for (int rd = 0; rd < 3; ++rd)
                        for (int cd = 0; cd < 3; ++cd)
                            t.push_back({e.st(i)[rs] * 3 + rd, e.st(i)[cs] * 3 + cd, -hess[(rs * 3 + rd) * K + cs * 3 + cd]});
        }
    }
    return fromTriplets(3 * g.n, t);
}
// getSystemMassMatrix (:145-174) for particles: diagonal
inline Csr systemMass(const Solver &s)
{
    std::vector<Trip> t;
    for (int i = 0; i < s.g.n; ++i)
        for (int c = 0; c < 3; ++c) t.push_back({3 * i + c, 3 * i + c, s.g.mass[i]});
    return fromTriplets(3 * s.g.n, t);
}
// alpha * A + beta * B over the union pattern
inline Csr combine(double alpha, const Csr &A, double beta, const Csr &B)
{
    std::vector<Trip> t;
    for (int r = 0; r < A.n; ++r) {
        for (int64_t k = A.rowptr[r]; k < A.rowptr[r + 1]; ++k) t.push_back({r, A.col[k], alpha * A.val[k]});
        for (int64_t k = B.rowptr[r]; k < B.rowptr[r + 1]; ++k) t.push_back({r, B.col[k], beta * B.val[k]});
    }
    return fromTriplets(A.n, t);
}
// solveLinearSystem (:223-238) = Eigen::ConjugateGradient with the default DiagonalPreconditioner
// (external/eigen/Eigen/src/IterativeLinearSolvers/ConjugateGradient.h:29-87, BasicPreconditioners.h:62-73), zero initial
// guess, tolerance 1e-5 on |r| / |b|, at most 300 iterations (implicitEulerBase.h:22-23).  Returns the iteration count.
inline int conjugateGradient(const Csr &A, const std::vector<double> &b, std::vector<double> &x, double tol = 1e-5,
                             int maxIters = 300)
{
    const size_t n = b.size();
    x.assign(n, 0.0);
    std::vector<double> invdiag(n, 1.0);
    for (int r = 0; r < A.n; ++r)
        for (int64_t k = A.rowptr[r]; k < A.rowptr[r + 1]; ++k)
            if (A.col[k] == r && A.val[k] != 0) invdiag[r] = 1.0 / A.val[k];
    std::vector<double> residual = b, p(n), z(n), tmp(n);
    const double rhsNorm2 = dotv(b, b);
    if (rhsNorm2 == 0) return 0;
    const double threshold = std::max(tol * tol * rhsNorm2, std::numeric_limits<double>::min());
    double residualNorm2 = dotv(residual, residual);
    if (residualNorm2 < threshold) return 0;
    for (size_t i = 0; i < n; ++i) p[i] = invdiag[i] * residual[i];
    double absNew = dotv(residual, p);
    int it = 0;
    while (it < maxIters) {
        spmv(A, p, tmp);
        const double alpha = absNew / dotv(p, tmp);
        for (size_t i = 0; i < n; ++i) x[i] += alpha * p[i];
        for (size_t i = 0; i < n; ++i) residual[i] -= alpha * tmp[i];
        residualNorm2 = dotv(residual, residual);
        if (residualNorm2 < threshold) break;
        for (size_t i = 0; i < n; ++i) z[i] = invdiag[i] * residual[i];
        const double absOld = absNew;
        absNew = dotv(residual, z);
        const double beta = absNew / absOld;
        for (size_t i = 0; i < n; ++i) p[i] = z[i] + beta * p[i];
        ++it;
    }
    return it;
}
// particleGroup.cpp:57-63
inline void integrateVelocities(Group &g, double dt)
{
    for (size_t k = 0; k < g.x.size(); ++k) g.x[k] += g.v[k] * dt;
}
// ImplicitEulerBaraffWitkin::step (implicitEulerBaraffWitkin.cpp:31-54), ImplicitEulerNewtonDx::step
// (implicitEulerNewtonDx.cpp:31-59), ImplicitEulerNewtonDv::step (implicitEulerNewtonDv.cpp:30-62)
inline void implicitStep(Solver &s)
{
    Group &g = s.g;
    const double dt = s.dt / s.substeps;
    const size_t n = g.x.size();
    for (int sub = 0; sub < s.substeps; ++sub) {
        std::vector<double> f, rhs(n), sol;
        if (s.kind == 3) {
            const std::vector<double> v0 = g.v;
            systemForce(s, f);
            const Csr M = systemMass(s), K = systemStiffness(s);
            const Csr lhs = combine(1.0, M, -(dt * dt), K);
            std::vector<double> Kv;
            spmv(K, v0, Kv);
            for (size_t i = 0; i < n; ++i) rhs[i] = dt * (f[i] + dt * Kv[i]);
            s.cgIterations.push_back(conjugateGradient(lhs, rhs, sol));
            for (size_t i = 0; i < n; ++i) g.v[i] = v0[i] + sol[i];
            integrateVelocities(g, dt);
        } else if (s.kind == 4) {
            const std::vector<double> x0 = g.x;
            const Csr M = systemMass(s);
            integrateVelocities(g, dt);
            systemForce(s, f);
            const Csr K = systemStiffness(s);
            const double invdt = 1. / dt;
            const Csr lhs = combine(invdt * invdt, M, -1.0, K);
            s.cgIterations.push_back(conjugateGradient(lhs, f, sol));
            // integrateObjectsVelocitiesFromDx (implicitEulerBase.cpp:76-88): x += dx; v = (x - x0) * invdt
            for (size_t i = 0; i < n; ++i) {
                g.x[i] += sol[i];
                g.v[i] = (g.x[i] - x0[i]) * invdt;
            }
        } else if (s.kind == 7) {  // BDF2::step, bdf2.cpp:31-70
            const std::vector<double> x0 = g.x, v0 = g.v;
            if (s.xMinusOne.size() != n || s.vMinusOne.size() != n) s.xMinusOne = x0, s.vMinusOne = v0;
            systemForce(s, f);
            const Csr M = systemMass(s), K = systemStiffness(s);
            const Csr lhs = combine(1.0, M, -(4.0 / 9.0 * (dt * dt)), K);
            std::vector<double> w(n), Kw, dvm(n), Mdv;
            for (size_t i = 0; i < n; ++i) w[i] = x0[i] - s.xMinusOne[i] + 2 * dt * v0[i];
            spmv(K, w, Kw);
            for (size_t i = 0; i < n; ++i) dvm[i] = v0[i] - s.vMinusOne[i];
            spmv(M, dvm, Mdv);
            for (size_t i = 0; i < n; ++i) rhs[i] = 1.0 / 3.0 * Mdv[i] + 2.0 / 3.0 * dt * (f[i] + 1.0 / 3.0 * Kw[i]);
            s.cgIterations.push_back(conjugateGradient(lhs, rhs, sol));
            for (size_t i = 0; i < n; ++i) {
                g.v[i] = v0[i] + sol[i];
                g.x[i] = 4.0 / 3.0 * x0[i] - 1.0 / 3.0 * s.xMinusOne[i] + 2.0 / 3.0 * g.v[i] * dt;
            }
            s.xMinusOne = x0;
            s.vMinusOne = v0;
        } else if (s.kind == 8) {  // StaticNewton::step, staticNewton.cpp:29-45: K dx = -f
            systemForce(s, f);
            const Csr K = systemStiffness(s);
            for (size_t i = 0; i < n; ++i) rhs[i] = -f[i];
            s.cgIterations.push_back(conjugateGradient(K, rhs, sol));
            for (size_t i = 0; i < n; ++i) g.x[i] += sol[i];
        } else if (s.kind == 9) {  // QuasiStaticNewton::step, quasiStaticNewton.cpp:31-49: (M - dt^2 K) dx = dt^2 f
            systemForce(s, f);
            const Csr M = systemMass(s), K = systemStiffness(s);
            const Csr lhs = combine(1.0, M, -(dt * dt), K);
            for (size_t i = 0; i < n; ++i) rhs[i] = dt * dt * f[i];
            s.cgIterations.push_back(conjugateGradient(lhs, rhs, sol));
            for (size_t i = 0; i < n; ++i) g.x[i] += sol[i];
        } else {
            const std::vector<double> x0 = g.x, v0 = g.v;
            const Csr M = systemMass(s);
            integrateVelocities(g, dt);
            systemForce(s, f);
            const Csr K = systemStiffness(s);
            const Csr lhs = combine(1.0, M, -(dt * dt), K);
            for (size_t i = 0; i < n; ++i) rhs[i] = dt * f[i];
            s.cgIterations.push_back(conjugateGradient(lhs, rhs, sol));
            g.x = x0;
            for (size_t i = 0; i < n; ++i) g.v[i] = v0[i] + sol[i];
            integrateVelocities(g, dt);
        }
    }
}
// ImplicitEulerNewtonRobust::step (implicitEulerNewtonRobust.cpp:31-127) and QuasiStaticNewtonRobust::step
// (quasiStaticNewtonRobust.cpp:33-125): up to 5 Newton iterations per substep, each followed by a halving line search on
// the residual norm; failed substeps are rolled back and shrink the solver's own time step.
inline void robustStep(Solver &s)
{
    Group &g = s.g;
    const bool quasi = s.kind == 11;
    const size_t n = g.x.size();
    const double threshold = quasi ? 1e-2 : 1e-5;
    for (int sub = 0; sub < s.substeps; ++sub) {
        bool successful = true;
        const double maxDt = s.dt / s.substeps;
        if (s.currentDt == -1. || s.currentDt > maxDt) s.currentDt = maxDt;
        const double dt = s.currentDt, invdt = 1 / dt, invdtSquared = invdt * invdt;
        const std::vector<double> x0 = g.x, v0 = g.v;
        if (!quasi) integrateVelocities(g, dt);
        std::vector<double> it = quasi ? x0 : v0;  // xi / vi
        std::vector<double> f, rhs(n), sol;
        auto residual = [&](std::vector<double> *out) {
            double acc = 0;
            for (size_t i = 0; i < n; ++i) {
                const double r = quasi ? f[i] - invdt * (g.mass[i / 3] * g.v[i]) : dt * f[i] - g.mass[i / 3] * (g.v[i] - v0[i]);
                if (out) (*out)[i] = r;
                acc += r * r;
            }
            return std::sqrt(acc);
        };
        double stepResidual = 0;
        int newton = 0;
        do {
            systemForce(s, f);
            const Csr M = systemMass(s), K = systemStiffness(s);
            double initialResidual;
            Csr lhs;
            if (quasi) {
                lhs = combine(invdtSquared, M, -1.0, K);
                for (size_t i = 0; i < n; ++i) g.v[i] = (g.x[i] - x0[i]) * invdt;  // computeSystemIntegratedVelocities
                initialResidual = residual(&rhs);
            } else {
                lhs = combine(1.0, M, -(dt * dt), K);
                double acc = 0;  // RHS uses the accepted iterate vi, not the state's velocities (:66)
                for (size_t i = 0; i < n; ++i) {
                    rhs[i] = dt * f[i] - g.mass[i / 3] * (it[i] - v0[i]);
                    acc += rhs[i] * rhs[i];
                }
                initialResidual = std::sqrt(acc);
            }
            s.cgIterations.push_back(conjugateGradient(lhs, rhs, sol));
            stepResidual = initialResidual;
            int ls = 0;
            double alpha = 1.;
            while (stepResidual >= initialResidual && ls < 5) {
                if (quasi) {
                    for (size_t i = 0; i < n; ++i) {
                        g.x[i] = it[i] + sol[i] * alpha;
                        g.v[i] = (g.x[i] - x0[i]) * invdt;
                    }
                } else {
                    for (size_t i = 0; i < n; ++i) {
                        g.v[i] = it[i] + sol[i] * alpha;
                        g.x[i] = x0[i] + g.v[i] * dt;
                    }
                }
                systemForce(s, f);
                stepResidual = residual(nullptr);
                if (stepResidual < initialResidual) it = quasi ? g.x : g.v;
                alpha *= 0.5;
                ++ls;
            }
            if (stepResidual >= initialResidual && stepResidual > threshold) successful = false;
            ++newton;
        } while (stepResidual > threshold && successful && newton < 5);
        if (stepResidual > threshold) successful = false;
        if (!successful) {
            g.x = x0;
            if (!quasi) g.v = v0;
            s.currentDt *= 0.75;
            s.consecutiveSuccessfulSteps = 0;
        } else {
            s.consecutiveSuccessfulSteps++;
            if (s.consecutiveSuccessfulSteps > 5 && s.currentDt < maxDt) s.currentDt /= 0.75;
        }
    }
}
// SimplecticEuler::step (simplecticEuler.cpp:12-62) with updateExplicitVelocities (energy.h:154-194):
// v += dt g; per stencil v_s += w_s (-grad_s) dt; x += v dt
inline void symplecticStep(Solver &s)
{
    Group &g = s.g;
    const double dt = s.dt / s.substeps;
    for (int sub = 0; sub < s.substeps; ++sub) {
        for (int i = 0; i < g.n; ++i)
            for (int c = 0; c < 3; ++c) g.v[3 * i + c] += dt * s.gravity[c];
        for (const auto &ep : g.energies) {
            const Energy &e = *ep;
            for (int i = 0; i < e.n; ++i) {
                double grad[12];
                energyDerivativesAny(e, i, g.x, nullptr, grad, nullptr);
                for (int l = 0; l < e.S; ++l) {
                    const int v = e.st(i)[l];
                    for (int c = 0; c < 3; ++c) g.v[3 * v + c] += g.invMass[v] * -grad[3 * l + c] * dt;
                }
            }
        }
        integrateVelocities(g, dt);
    }
}

thread_local std::string g_error;
template <class F>
int guarded(F &&f)
{
    try {
        f();
        return 0;
    } catch (const std::exception &e) {
        g_error = e.what();
        return 1;
    } catch (...) {
        g_error = "unknown exception";
        return 2;
    }
}
}  // namespace port

// ---------------------------------------------------------------------------------------------------------------------
// C ABI: the entry points of oracle/ref_driver.cpp that tests/oracle_api.py binds (scene builders excepted).
// ---------------------------------------------------------------------------------------------------------------------
using namespace port;

extern "C" {
const char *ref_last_error() { return g_error.c_str(); }
int ref_omp_max_threads() { return omp_get_max_threads(); }
void ref_omp_set_threads(int n) { omp_set_num_threads(n); }
int ref_kind_stencil_size(int kind) { return (kind >= 0 && kind < K_COUNT) ? kStencilSize[kind] : -1; }
int ref_kind_num_model_params(int kind) { return (kind >= 0 && kind < K_COUNT) ? kNumModelParams[kind] : -1; }

void *ref_group_create(int nVerts, const double *x, const double *x0, const double *mass)
{
    auto *g = new Group;
    g->n = nVerts;
    g->x.assign(x, x + 3 * static_cast<size_t>(nVerts));
    g->x0.assign(x0, x0 + 3 * static_cast<size_t>(nVerts));
    g->xInitial = g->x;
    g->v.assign(3 * static_cast<size_t>(nVerts), 0.0);
    g->mass.assign(mass, mass + nVerts);
    g->invMass.resize(static_cast<size_t>(nVerts));
    for (int i = 0; i < nVerts; ++i) g->invMass[i] = mass[i] == 0 ? 0 : 1 / mass[i];  // particleGroup.cpp:74-82
    return g;
}
void ref_group_destroy(void *h) { delete static_cast<Group *>(h); }

int ref_group_add_energy(void *h, int kind, int nStencils, const int32_t *stencils, const double *params)
{
    return guarded([&] {
        auto *g = static_cast<Group *>(h);
        if (kind < 0 || kind >= K_COUNT) throw std::runtime_error("unknown energy kind");
        auto e = std::make_shared<Energy>();
        e->kind = kind;
        e->S = kStencilSize[kind];
        e->NC = kNumConstraints[kind];
        for (int i = 0; i < nStencils; ++i)
            addStencil(*e, stencils + static_cast<size_t>(i) * e->S, params + static_cast<size_t>(i) * kNumModelParams[kind]);
        prepare(*e, *g);
        g->energies.push_back(e);
    });
}
int ref_update_anchor(void *h, int e, int stencil, const double *anchor3)
{
    return guarded([&] {
        auto &en = *static_cast<Group *>(h)->energies.at(e);
        if (en.kind != K_ANCHOR) throw std::runtime_error("not an anchor energy");  // (static loads share the field)
        std::copy(anchor3, anchor3 + 3, en.anchor.begin() + 3 * static_cast<size_t>(stencil));
    });
}
int ref_group_size(void *h) { return static_cast<Group *>(h)->n; }
void ref_group_get_state(void *h, double *x, double *v)
{
    auto *g = static_cast<Group *>(h);
    if (x) std::copy(g->x.begin(), g->x.end(), x);
    if (v) std::copy(g->v.begin(), g->v.end(), v);
}
void ref_group_set_state(void *h, const double *x, const double *v)
{
    auto *g = static_cast<Group *>(h);
    if (x) g->x.assign(x, x + g->x.size());
    if (v) g->v.assign(v, v + g->v.size());
}
void ref_group_get_masses(void *h, double *m, double *w)
{
    auto *g = static_cast<Group *>(h);
    std::copy(g->mass.begin(), g->mass.end(), m);
    std::copy(g->invMass.begin(), g->invMass.end(), w);
}
int ref_energy_value(void *h, int e, int stencil, double *out)
{
    return guarded([&] {
        auto *g = static_cast<Group *>(h);
        energyDerivativesAny(*g->energies.at(e), stencil, g->x, out, nullptr, nullptr);
    });
}
int ref_energy_grad_hess(void *h, int e, int stencil, double *grad, double *hess)
{
    return guarded([&] {
        auto *g = static_cast<Group *>(h);
        energyDerivativesAny(*g->energies.at(e), stencil, g->x, nullptr, grad, hess);
    });
}
// accumulateVertexForce / accumulateVertexHessian (energy.h:240-258) into zero-initialised outputs
int ref_vertex_force_hessian(void *h, int e, int stencil, int local, double *force3, double *hess9)
{
    return guarded([&] {
        auto *g = static_cast<Group *>(h);
        const Energy &en = *g->energies.at(e);
        double grad[12], hess[144];
        energyDerivativesAny(en, stencil, g->x, nullptr, grad, hess);
        const int K = 3 * en.S, o = 3 * local;
        for (int a = 0; a < 3; ++a) {
            force3[a] = -grad[o + a];
            for (int b = 0; b < 3; ++b) hess9[3 * a + b] = hess[(o + a) * K + o + b];
        }
    });
}
int ref_constraints_and_gradient(void *h, int e, int stencil, double *C, double *grad)
{
    return guarded([&] {
        auto *g = static_cast<Group *>(h);
        constraintsAndGradientAny(*g->energies.at(e), stencil, g->x, C, grad);
    });
}
int ref_project_position(void *h, int e, int stencil, double dt)
{
    return guarded([&] {
        auto *g = static_cast<Group *>(h);
        projectPosition(*g->energies.at(e), stencil, *g, dt);
    });
}
int ref_effective_stiffness(void *h, int e, int stencil, double *K, double *A)
{
    return guarded([&] {
        const Energy &en = *static_cast<Group *>(h)->energies.at(e);
        std::copy(en.K(en.Keff, stencil), en.K(en.Keff, stencil) + en.NC * en.NC, K);
        std::copy(en.K(en.Aeff, stencil), en.K(en.Aeff, stencil) + en.NC * en.NC, A);
    });
}
// same layout as oracle/ref_driver.cpp: bending -> restTheta; quadratic bending -> 3 x 12 laplacian; membranes -> Dm^-1
// (4); tets -> Dm^-1 (9) (+ alpha for SNH)
int ref_rest_data(void *h, int e, int stencil, double *out, int *n)
{
    return guarded([&] {
        const Energy &en = *static_cast<Group *>(h)->energies.at(e);
        *n = 0;
        switch (en.kind) {
        case K_DISCRETE_BENDING:
        case K_BENDING_BW: out[(*n)++] = en.restTheta[stencil]; break;
        case K_STATIC_LOAD:
            for (int c = 0; c < 3; ++c) out[(*n)++] = en.anchor[3 * static_cast<size_t>(stencil) + c];
            break;
        case K_QUADRATIC_BENDING:
            for (int r = 0; r < 3; ++r)
                for (int c = 0; c < 12; ++c) out[(*n)++] = (c % 3 == r) ? en.laplacian[4 * static_cast<size_t>(stencil) + c / 3] : 0.0;
            break;
        case K_MEMBRANE_STVK:
        case K_MEMBRANE_CHOI:
        case K_MEMBRANE_BW:
            for (int k = 0; k < 4; ++k) out[(*n)++] = en.DmInv[4 * static_cast<size_t>(stencil) + k];
            break;
        case K_TET_STVK:
        case K_TET_SNH:
        case K_TET_SSNH:
            for (int k = 0; k < 9; ++k) out[(*n)++] = en.DmInv[9 * static_cast<size_t>(stencil) + k];
            if (en.kind == K_TET_SNH) out[(*n)++] = en.alpha[stencil];
            break;
        default: break;
        }
    });
}

static std::vector<std::pair<const int *, std::pair<int64_t, int>>> setsOf(const Group &g)
{
    std::vector<std::pair<const int *, std::pair<int64_t, int>>> sets;
    for (const auto &e : g.energies) sets.push_back({e->stencils.data(), {static_cast<int64_t>(e->stencils.size()), e->S}});
    return sets;
}
int ref_color_vertices(void *h, int32_t *colors)
{
    return guarded([&] {
        auto *g = static_cast<Group *>(h);
        const auto c = colorVertices(g->n, setsOf(*g));
        std::copy(c.begin(), c.end(), colors);
    });
}
int ref_color_stencils(void *h, int e, int32_t *colors)
{
    return guarded([&] {
        const Energy &en = *static_cast<Group *>(h)->energies.at(e);
        const auto c = colorStencils(en.stencils.data(), static_cast<int64_t>(en.stencils.size()), en.S);
        std::copy(c.begin(), c.end(), colors);
    });
}
int ref_color_vertices_flat(int nVerts, int nSets, const int32_t *const *entries, const int32_t *nEntries,
                            const int32_t *stencilSizes, int32_t *colors)
{
    return guarded([&] {
        std::vector<std::pair<const int *, std::pair<int64_t, int>>> sets;
        for (int s = 0; s < nSets; ++s) sets.push_back({entries[s], {nEntries[s], stencilSizes[s]}});
        const auto c = colorVertices(nVerts, sets);
        std::copy(c.begin(), c.end(), colors);
    });
}
int ref_color_stencils_flat(const int32_t *entries, int nEntries, int stencilSize, int32_t *colors)
{
    return guarded([&] {
        const auto c = colorStencils(entries, nEntries, stencilSize);
        std::copy(c.begin(), c.end(), colors);
    });
}

void *ref_solver_create(int kind, void *group, double dt, int substeps, int iterations, const double *gravity)
{
    auto *s = new Solver;
    s->kind = kind;
    s->g = *static_cast<Group *>(group);  // baseSolver.cpp:10-14: the solver steps its own copy
    s->dt = dt;
    s->substeps = substeps;
    s->iterations = iterations;
    if (gravity) std::copy(gravity, gravity + 3, s->gravity);
    return s;
}
void ref_solver_destroy(void *h) { delete static_cast<Solver *>(h); }
// live topology edits (picking: apps/solverCompare.cpp:717-726, :1446-1455, :1471-1473, :1499-1502) on the solver's copy
// of the group, with requirePrecomputationUpdate (vbd.h:26, xpbd.h:20): colours and incidence are rebuilt on the next step
int ref_solver_add_energy(void *h, int kind, int nStencils, const int32_t *stencils, const double *params)
{
    return guarded([&] {
        auto *s = static_cast<Solver *>(h);
        if (kind < 0 || kind >= K_COUNT) throw std::runtime_error("unknown energy kind");
        auto e = std::make_shared<Energy>();
        e->kind = kind;
        e->S = kStencilSize[kind];
        e->NC = kNumConstraints[kind];
        for (int i = 0; i < nStencils; ++i)
            addStencil(*e, stencils + static_cast<size_t>(i) * e->S, params + static_cast<size_t>(i) * kNumModelParams[kind]);
        prepare(*e, s->g);
        s->g.energies.push_back(e);
        s->vertexGroups.clear();
        s->stencilGroups.clear();
        s->dirty = true;
    });
}
int ref_solver_remove_energy(void *h, int e)
{
    return guarded([&] {
        auto *s = static_cast<Solver *>(h);
        if (e < 0 || e >= static_cast<int>(s->g.energies.size())) throw std::out_of_range("energy index");
        s->g.energies.erase(s->g.energies.begin() + e);
        s->vertexGroups.clear();
        s->stencilGroups.clear();
        s->dirty = true;
    });
}
int ref_solver_update_anchor(void *h, int e, int stencil, const double *anchor3)
{
    return guarded([&] {
        auto &en = *static_cast<Solver *>(h)->g.energies.at(e);
        if (en.kind != K_ANCHOR) throw std::runtime_error("not an anchor energy");
        std::copy(anchor3, anchor3 + 3, en.anchor.begin() + 3 * static_cast<size_t>(stencil));
    });
}
int ref_solver_step(void *h, int nFrames)
{
    return guarded([&] {
        auto *s = static_cast<Solver *>(h);
        s->cgIterations.clear();
        for (int f = 0; f < nFrames; ++f) {
            if (s->kind == 0)
                vbdStep(*s);
            else if ((s->kind >= 3 && s->kind <= 5) || (s->kind >= 7 && s->kind <= 9))
                implicitStep(*s);
            else if (s->kind == 6)
                symplecticStep(*s);
            else if (s->kind == 10 || s->kind == 11)
                robustStep(*s);
            else
                xpbdStep(*s);
        }
    });
}
int ref_solver_cg_iterations(void *h, int32_t *out, int capacity, int *n)
{
    return guarded([&] {
        auto *s = static_cast<Solver *>(h);
        *n = static_cast<int>(s->cgIterations.size());
        for (int i = 0; i < *n && i < capacity; ++i) out[i] = s->cgIterations[i];
    });
}
int ref_system_force(void *h, double *f)
{
    return guarded([&] {
        std::vector<double> fv;
        systemForce(*static_cast<Solver *>(h), fv);
        std::copy(fv.begin(), fv.end(), f);
    });
}
int ref_system_matrix(void *h, int which, int64_t *nnz, int32_t *rows, int32_t *cols, double *vals)
{
    return guarded([&] {
        auto *s = static_cast<Solver *>(h);
        const Csr A = which == 0 ? systemStiffness(*s) : systemMass(*s);
        *nnz = static_cast<int64_t>(A.val.size());
        if (!rows) return;
        for (int r = 0; r < A.n; ++r)
            for (int64_t k = A.rowptr[r]; k < A.rowptr[r + 1]; ++k) rows[k] = r, cols[k] = A.col[k], vals[k] = A.val[k];
    });
}
int ref_linear_solve(void *, int n, int64_t nnz, const int32_t *rows, const int32_t *cols, const double *vals,
                     const double *b, double *x, int *iterations)
{
    return guarded([&] {
        std::vector<Trip> t(static_cast<size_t>(nnz));
        for (int64_t k = 0; k < nnz; ++k) t[k] = {rows[k], cols[k], vals[k]};
        const Csr A = fromTriplets(n, t);
        std::vector<double> bv(b, b + n), xv;
        const int it = conjugateGradient(A, bv, xv);
        if (iterations) *iterations = it;
        std::copy(xv.begin(), xv.end(), x);
    });
}
int ref_solver_reset(void *h)  // ParticleGroup::reset, particleGroup.cpp:97-103
{
    return guarded([&] {
        auto *s = static_cast<Solver *>(h);
        s->g.x = s->g.xInitial;
        std::fill(s->g.v.begin(), s->g.v.end(), 0.0);
        s->xMinusOne.clear();  // BDF2::reset, bdf2.cpp:78-83
        s->vMinusOne.clear();
        s->currentDt = -1.;    // ImplicitEulerNewtonRobust::reset, implicitEulerNewtonRobust.cpp:132-137
        s->consecutiveSuccessfulSteps = 0;
    });
}
void ref_solver_get_state(void *h, double *x, double *v) { ref_group_get_state(&static_cast<Solver *>(h)->g, x, v); }
void ref_solver_set_state(void *h, const double *x, const double *v) { ref_group_set_state(&static_cast<Solver *>(h)->g, x, v); }

int ref_vbd_vertex_colors(void *h, int32_t *colors, int *nColors)
{
    return guarded([&] {
        auto *s = static_cast<Solver *>(h);
        if (s->dirty) vbdPrecompute(*s);
        for (size_t c = 0; c < s->vertexGroups.size(); ++c)
            for (int v : s->vertexGroups[c]) colors[v] = static_cast<int32_t>(c);
        if (nColors) *nColors = static_cast<int>(s->vertexGroups.size());
    });
}
// rowptr[v * nE + e] .. : (stencil, local) pairs of vertex v and energy e in the reference's order (vbd.cpp:231-268)
int ref_vbd_incidence(void *h, int64_t *rowptr, int32_t *entries, int64_t *nEntries)
{
    return guarded([&] {
        auto *s = static_cast<Solver *>(h);
        if (s->dirty) vbdPrecompute(*s);
        const size_t nE = s->g.energies.size();
        int64_t cursor = 0;
        for (int v = 0; v < s->g.n; ++v)
            for (size_t e = 0; e < nE; ++e) {
                if (rowptr) rowptr[v * nE + e] = cursor;
                for (const auto &ent : s->incidence[v][e]) {
                    if (entries) entries[2 * cursor] = ent[0], entries[2 * cursor + 1] = ent[1];
                    ++cursor;
                }
            }
        if (rowptr) rowptr[static_cast<size_t>(s->g.n) * nE] = cursor;
        if (nEntries) *nEntries = cursor;
    });
}
int ref_vbd_inject_colors(void *h, const int32_t *colors, int nVerts)
{
    return guarded([&] {
        auto *s = static_cast<Solver *>(h);
        if (nVerts != s->g.n) throw std::runtime_error("colour array size mismatch");
        groupBy(std::vector<int>(colors, colors + nVerts), s->vertexGroups);
        s->dirty = true;  // incidence is (re)built on the next step; the injected groups are kept
    });
}
int ref_xpbd_stencil_colors(void *h, int e, int32_t *colors, int *nColors)
{
    return guarded([&] {
        auto *s = static_cast<Solver *>(h);
        const Energy &en = *s->g.energies.at(e);
        std::vector<std::vector<int>> groups;
        if (s->stencilGroups.size() == s->g.energies.size() && !s->stencilGroups[e].empty())
            groups = s->stencilGroups[e];
        else
            groupBy(colorStencils(en.stencils.data(), static_cast<int64_t>(en.stencils.size()), en.S), groups);
        for (size_t c = 0; c < groups.size(); ++c)
            for (int i : groups[c]) colors[i] = static_cast<int32_t>(c);
        if (nColors) *nColors = static_cast<int>(groups.size());
    });
}
int ref_xpbd_inject_colors(void *h, int e, const int32_t *colors, int nStencils)
{
    return guarded([&] {
        auto *s = static_cast<Solver *>(h);
        if (s->stencilGroups.size() != s->g.energies.size()) s->stencilGroups.assign(s->g.energies.size(), {});
        if (nStencils != s->g.energies.at(e)->n) throw std::runtime_error("colour array size mismatch");
        groupBy(std::vector<int>(colors, colors + nStencils), s->stencilGroups[e]);
    });
}
}  // extern "C"
