// Rigid-body groups under VBD and XPBD on the device.
//
// Reference: RigidBodyGroup (spg/source/sim/simObject/rigidBodyGroup.cpp:6-243), the rigid-body branches of VBD::step
// (spg/source/sim/solver/vbd.cpp:63-72, :104-147) and XPBD::step (xpbd.cpp:60-70), projectPosition with the 6 x 6 inverse
// mass block (spg/include/spg/sim/energy/energy.h:117-153) and the four energies of spg/source/sim/energy/rigidBody*.cpp.
// These are small systems (chains, articulated bodies: tens to thousands of bodies), so the kernels are the plain form:
// one thread per body of a colour (VBD), one thread per stencil of a colour (XPBD), every launch of a step captured in one
// CUDA graph.  Same control flow, same quirks: the rotation matrix is the state that is composed incrementally and never
// re-orthonormalised, theta is re-derived from it after every update, gravity is applied to every body, a vertex whose
// 6 x 6 system has det <= 1e-10 is skipped.
#include <algorithm>
#include <cstring>
#include <memory>

#include "group.h"
#include "host_build.h"
#include "rigid.cuh"

namespace spgx
{
namespace
{
constexpr int kRbMaxEnergies = 8;

struct RbEnergyDev {
    int kind, n;
    const int32_t *stencils;  // S body ids per stencil (XPBD: ordered by colour)
    const double *records;    // NP doubles per stencil
    const uint32_t *rowptr;   // VBD: nBodies + 1
    const uint32_t *entries;  // VBD: (stencil << 1) | local, ascending per body
};
struct RbState {
    double *x, *theta, *v, *omega, *R, *inertia, *inertiaInv, *xOld, *thetaOld, *xInertial;
    const double *mass, *w, *localI, *localIinv;
    int n;
};

__device__ __forceinline__ V3d ld3(const double *p, int i)
{
    return V3d(p[3 * i], p[3 * i + 1], p[3 * i + 2]);
}
__device__ __forceinline__ void st3(double *p, int i, const V3d &v)
{
    p[3 * i] = v.x, p[3 * i + 1] = v.y, p[3 * i + 2] = v.z;
}

// theta and the world-frame inertia follow the rotation matrix (updateTheta + updateInertia, rigidBodyGroup.cpp:207-237)
__device__ __forceinline__ void refreshDerived(const RbState &s, int i)
{
    st3(s.theta, i, matrixToTheta(s.R + 9 * i));
    rotateInertia(s.R + 9 * i, s.localI + 9 * i, s.inertia + 9 * i);
    rotateInertia(s.R + 9 * i, s.localIinv + 9 * i, s.inertiaInv + 9 * i);
}
// updateElementPositionFromDx (rigidBodyGroup.cpp:64-78)
__device__ __forceinline__ void applyDx(const RbState &s, int i, const double *dx6)
{
    st3(s.x, i, ld3(s.x, i) + V3d(dx6[0], dx6[1], dx6[2]));
    composeRotation(V3d(dx6[3], dx6[4], dx6[5]), s.R + 9 * i);
    refreshDerived(s, i);
}

// vbd.cpp:49-75 / xpbd.cpp:48-74: store the previous state, explicit forces on the velocities (gravity; the gyroscopic term
// omega += dt Iinv (-omega x I omega)), integrateVelocities (rigidBodyGroup.cpp:96-116), inertial position
__global__ void rbPredictKernel(const RbState s, double dt, double gx, double gy, double gz)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s.n) return;
    st3(s.xOld, i, ld3(s.x, i));
    st3(s.thetaOld, i, ld3(s.theta, i));
    const V3d v = ld3(s.v, i) + V3d(dt * gx, dt * gy, dt * gz);
    V3d om = ld3(s.omega, i);
    const V3d Iw = mat3Vec(s.inertia + 9 * i, om);
    const V3d gyro = mat3Vec(s.inertiaInv + 9 * i, -cross(om, Iw));
    om = om + gyro * dt;
    st3(s.v, i, v);
    st3(s.omega, i, om);
    st3(s.x, i, ld3(s.x, i) + v * dt);
    const double on = ::sqrt(dot(om, om));
    if (on != 0) {
        double dR[9], out[9];
        angleAxisToMatrix(on * dt, V3d(om.x / on, om.y / on, om.z / on), dR);
        mat3Mul(dR, s.R + 9 * i, out);
        for (int k = 0; k < 9; ++k) s.R[9 * i + k] = out[k];
    }
    refreshDerived(s, i);
    st3(s.xInertial, i, ld3(s.x, i));
}
// computeIntegratedVelocities (rigidBodyGroup.cpp:118-130)
__global__ void rbFinalizeKernel(const RbState s, double invdt)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s.n) return;
    st3(s.v, i, (ld3(s.x, i) - ld3(s.xOld, i)) * invdt);
    st3(s.omega, i, integratedOmega(s.R + 9 * i, ld3(s.thetaOld, i), invdt));
}
// setPositions (rigidBodyGroup.cpp:43-53): rotation matrices and inertias from theta
__global__ void rbFromThetaKernel(const RbState s)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s.n) return;
    thetaToMatrix(ld3(s.theta, i), s.R + 9 * i);
    rotateInertia(s.R + 9 * i, s.localI + 9 * i, s.inertia + 9 * i);
    rotateInertia(s.R + 9 * i, s.localIinv + 9 * i, s.inertiaInv + 9 * i);
}

// the world-frame inertias of rotation matrices that were restored as they are (a re-finalize keeps the composed matrices)
__global__ void rbInertiaKernel(const RbState s)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s.n) return;
    rotateInertia(s.R + 9 * i, s.localI + 9 * i, s.inertia + 9 * i);
    rotateInertia(s.R + 9 * i, s.localIinv + 9 * i, s.inertiaInv + 9 * i);
}

// C of one stencil over the scalar type T; seedBody >= 0: only that body's 6 DOFs are active (K = 6), otherwise all 6 S
template <int KIND, class T>
__device__ __forceinline__ T rbConstraint(const RbState &s, const int32_t *ids, const double *p, int seedBody)
{
    using EK = RbEnergyKind<KIND>;
    V3<T> com[EK::S], dth[EK::S];
    const double *R[EK::S];
#pragma unroll
    for (int b = 0; b < EK::S; ++b) {
        const int id = ids[b];
        const V3d x = ld3(s.x, id);
        const int base = seedBody < 0 ? 6 * b : (b == seedBody ? 0 : -100);
        com[b] = V3<T>(T(x.x, base), T(x.y, base + 1), T(x.z, base + 2));
        dth[b] = V3<T>(T(0.0, base + 3), T(0.0, base + 4), T(0.0, base + 5));
        R[b] = s.R + 9 * id;
    }
    return EK::template constraint<T>(com, dth, R, p);
}

// the owner body's share of -dE/dq (6) and d2E/dq2 (6 x 6, row-major) of one stencil: accumulateVertexForce / Hessian
// (energy.h:240-258).  E = 1/2 K C^2 through second-order duals, except the anchor, whose energy / gradient / Hessian the
// reference overrides with the point-map form J^T k (p - a), J^T k J, J = [I | skew(-R local)]
// (rigidBodySpringAnchorEnergy.cpp:45-66, :99-123).
template <int KIND>
__device__ void rbVertexGradHess(const RbState &s, const int32_t *ids, const double *p, int local, double *g6, double *H36)
{
    using EK = RbEnergyKind<KIND>;
    if constexpr (KIND == RB_SPRING_ANCHOR) {
        const int id = ids[0];
        const V3d r = mat3Vec(s.R + 9 * id, V3d(p[0], p[1], p[2]));
        const V3d pt = ld3(s.x, id) + r;
        const double k = p[6];
        const V3d d(pt.x - p[3], pt.y - p[4], pt.z - p[5]);
        // J = [I | skew(-r)], skew(-r) = [[0, r.z, -r.y], [-r.z, 0, r.x], [r.y, -r.x, 0]]
        double J[3][6] = {{1, 0, 0, 0, r.z, -r.y}, {0, 1, 0, -r.z, 0, r.x}, {0, 0, 1, r.y, -r.x, 0}};
        const double e[3] = {k * d.x, k * d.y, k * d.z};
        for (int a = 0; a < 6; ++a) {
            g6[a] += J[0][a] * e[0] + J[1][a] * e[1] + J[2][a] * e[2];
            for (int b = 0; b < 6; ++b) H36[6 * a + b] += k * (J[0][a] * J[0][b] + J[1][a] * J[1][b] + J[2][a] * J[2][b]);
        }
    } else {
        using D2 = Dual<6, true>;
        const D2 C = rbConstraint<KIND, D2>(s, ids, p, local);
        const D2 E = 0.5 * (C * (C * EK::K(p)));
        for (int a = 0; a < 6; ++a) {
            g6[a] += E.g[a];
            for (int b = 0; b < 6; ++b) H36[6 * a + b] += E.hess(a, b);
        }
    }
}

// 6 x 6: determinant and solution by LU with partial pivoting (what Eigen's determinant() / inverse() do beyond 4 x 4)
__device__ bool solve6(const double *H36, const double *f6, double eps, double *dx6)
{
    double A[6][7];
    for (int r = 0; r < 6; ++r) {
        for (int c = 0; c < 6; ++c) A[r][c] = H36[6 * r + c];
        A[r][6] = f6[r];
    }
    double det = 1.0;
    for (int k = 0; k < 6; ++k) {
        int piv = k;
        for (int r = k + 1; r < 6; ++r)
            if (::fabs(A[r][k]) > ::fabs(A[piv][k])) piv = r;
        if (A[piv][k] == 0.0) return false;
        if (piv != k) {
            for (int c = 0; c < 7; ++c) {
                const double t = A[k][c];
                A[k][c] = A[piv][c];
                A[piv][c] = t;
            }
            det = -det;
        }
        det *= A[k][k];
        for (int r = k + 1; r < 6; ++r) {
            const double f = A[r][k] / A[k][k];
            for (int c = k + 1; c < 7; ++c) A[r][c] -= f * A[k][c];
        }
    }
    if (!(det > eps)) return false;  // vbd.cpp:139
    for (int k = 5; k >= 0; --k) {
        double acc = A[k][6];
        for (int c = k + 1; c < 6; ++c) acc -= A[k][c] * dx6[c];
        dx6[k] = acc / A[k][k];
    }
    return true;
}

struct RbVbdArgs {
    RbState s;
    int nEnergies;
    RbEnergyDev e[kRbMaxEnergies];
    const int32_t *order;  // bodies sorted by (colour, id)
};

__device__ void rbAccumulate(const RbVbdArgs &a, int body, double *g6, double *H36)
{
    for (int e = 0; e < a.nEnergies; ++e) {
        const RbEnergyDev &en = a.e[e];
        for (uint32_t q = en.rowptr[body]; q < en.rowptr[body + 1]; ++q) {
            const uint32_t ent = en.entries[q];
            const int st = static_cast<int>(ent >> 1), local = static_cast<int>(ent & 1u);
            switch (en.kind) {
            case RB_SPRING:
                rbVertexGradHess<RB_SPRING>(a.s, en.stencils + 2 * st, en.records + RbEnergyKind<RB_SPRING>::NP * st, local, g6, H36);
                break;
            case RB_SPRING_SQUARED:
                rbVertexGradHess<RB_SPRING_SQUARED>(a.s, en.stencils + 2 * st, en.records + RbEnergyKind<RB_SPRING_SQUARED>::NP * st, local, g6, H36);
                break;
            case RB_SPRING_ANCHOR:
                rbVertexGradHess<RB_SPRING_ANCHOR>(a.s, en.stencils + st, en.records + RbEnergyKind<RB_SPRING_ANCHOR>::NP * st, local, g6, H36);
                break;
            case RB_ORIENTATION_ALIGNMENT:
                rbVertexGradHess<RB_ORIENTATION_ALIGNMENT>(a.s, en.stencils + 2 * st, en.records + RbEnergyKind<RB_ORIENTATION_ALIGNMENT>::NP * st, local,
                                                           g6, H36);
                break;
            default: break;
            }
        }
    }
}

// l_vertexDescent for a rigid body (vbd.cpp:99-147): bodies [begin, end) of `order` share no stencil
__global__ void rbVbdPassKernel(const RbVbdArgs a, int begin, int end, double invdt)
{
    const int t = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= end) return;
    const int i = a.order[t];
    const RbState &s = a.s;
    const double invdtSquared = invdt * invdt;
    const double mOverdtSquared = s.mass[i] * invdtSquared;
    double g[6] = {0, 0, 0, 0, 0, 0}, H[36];
    for (int k = 0; k < 36; ++k) H[k] = 0.0;
    rbAccumulate(a, i, g, H);
    double f[6];
    const V3d dxI = ld3(s.x, i) - ld3(s.xInertial, i);
    f[0] = -mOverdtSquared * dxI.x - g[0];
    f[1] = -mOverdtSquared * dxI.y - g[1];
    f[2] = -mOverdtSquared * dxI.z - g[2];
    const V3d dOm = integratedOmega(s.R + 9 * i, ld3(s.thetaOld, i), invdt) - ld3(s.omega, i);
    const V3d IdOm = mat3Vec(s.inertia + 9 * i, dOm);
    f[3] = -invdt * IdOm.x - g[3];
    f[4] = -invdt * IdOm.y - g[4];
    f[5] = -invdt * IdOm.z - g[5];
    for (int k = 0; k < 3; ++k) H[7 * k] += mOverdtSquared;
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) H[6 * (3 + r) + 3 + c] += invdtSquared * s.inertia[9 * i + 3 * r + c];
    double dx[6];
    if (solve6(H, f, 1e-10, dx)) applyDx(s, i, dx);
}

// projectPosition for one stencil (energy.h:117-153 with the rigid-body W block)
template <int KIND>
__device__ void rbProject(const RbState &s, const int32_t *ids, const double *p, double dt)
{
    using EK = RbEnergyKind<KIND>;
    constexpr int S = EK::S, K = 6 * S;
    using D1 = Dual<K, false>;
    const D1 C = rbConstraint<KIND, D1>(s, ids, p, -1);
    double Wg[K];
    double A = 0.0;
    for (int b = 0; b < S; ++b) {
        const int id = ids[b];
        const double w = s.w[id];
        for (int m = 0; m < 3; ++m) Wg[6 * b + m] = w * C.g[6 * b + m];
        const V3d ga(C.g[6 * b + 3], C.g[6 * b + 4], C.g[6 * b + 5]);
        const V3d Ig = mat3Vec(s.inertiaInv + 9 * id, ga);
        Wg[6 * b + 3] = Ig.x, Wg[6 * b + 4] = Ig.y, Wg[6 * b + 5] = Ig.z;
    }
    for (int k = 0; k < K; ++k) A += C.g[k] * Wg[k];
    A += EK::alpha(p) / (dt * dt) + 1e-10;
    const double dl = -(1.0 / A) * C.v;
    for (int b = 0; b < S; ++b) {
        double dx[6];
        for (int m = 0; m < 6; ++m) dx[m] = Wg[6 * b + m] * dl;
        applyDx(s, ids[b], dx);
    }
}
__device__ void rbProjectKind(int kind, const RbState &s, const int32_t *stencils, const double *records, int i, double dt)
{
    switch (kind) {
    case RB_SPRING: rbProject<RB_SPRING>(s, stencils + 2 * i, records + RbEnergyKind<RB_SPRING>::NP * i, dt); break;
    case RB_SPRING_SQUARED: rbProject<RB_SPRING_SQUARED>(s, stencils + 2 * i, records + RbEnergyKind<RB_SPRING_SQUARED>::NP * i, dt); break;
    case RB_SPRING_ANCHOR: rbProject<RB_SPRING_ANCHOR>(s, stencils + i, records + RbEnergyKind<RB_SPRING_ANCHOR>::NP * i, dt); break;
    case RB_ORIENTATION_ALIGNMENT:
        rbProject<RB_ORIENTATION_ALIGNMENT>(s, stencils + 2 * i, records + RbEnergyKind<RB_ORIENTATION_ALIGNMENT>::NP * i, dt);
        break;
    default: break;
    }
}
// one stencil colour of one energy (xpbd.cpp:88-95); serial: one thread walks [begin, end) in index order (:96-102)
__global__ void rbXpbdPassKernel(const RbState s, int kind, const int32_t *stencils, const double *records, int begin, int end,
                                 double dt, int serial)
{
    if (serial) {
        if (blockIdx.x == 0 && threadIdx.x == 0)
            for (int i = begin; i < end; ++i) {
                rbProjectKind(kind, s, stencils, records, i, dt);
                __threadfence();
            }
        return;
    }
    const int i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i < end) rbProjectKind(kind, s, stencils, records, i, dt);
}

// single-element evaluators for the parity tests
__global__ void rbEvalConstraintKernel(const RbState s, int kind, const int32_t *ids, const double *p, double *out)
{
    // out[0] = C, out[1 + k] = dC/dq_k (6 S)
    switch (kind) {
#define SPGX_RB_CASE(KIND)                                                                                             \
    case KIND: {                                                                                                       \
        using D1 = Dual<6 * RbEnergyKind<KIND>::S, false>;                                                             \
        const D1 C = rbConstraint<KIND, D1>(s, ids, p, -1);                                                            \
        out[0] = C.v;                                                                                                  \
        for (int k = 0; k < 6 * RbEnergyKind<KIND>::S; ++k) out[1 + k] = C.g[k];                                       \
        break;                                                                                                         \
    }
        SPGX_RB_CASE(RB_SPRING)
        SPGX_RB_CASE(RB_SPRING_SQUARED)
        SPGX_RB_CASE(RB_SPRING_ANCHOR)
        SPGX_RB_CASE(RB_ORIENTATION_ALIGNMENT)
#undef SPGX_RB_CASE
    default: break;
    }
}
__global__ void rbEvalVertexKernel(const RbState s, int kind, const int32_t *ids, const double *p, int local, double *out42)
{
    double g[6] = {0, 0, 0, 0, 0, 0}, H[36];
    for (int k = 0; k < 36; ++k) H[k] = 0.0;
    switch (kind) {
    case RB_SPRING: rbVertexGradHess<RB_SPRING>(s, ids, p, local, g, H); break;
    case RB_SPRING_SQUARED: rbVertexGradHess<RB_SPRING_SQUARED>(s, ids, p, local, g, H); break;
    case RB_SPRING_ANCHOR: rbVertexGradHess<RB_SPRING_ANCHOR>(s, ids, p, local, g, H); break;
    case RB_ORIENTATION_ALIGNMENT: rbVertexGradHess<RB_ORIENTATION_ALIGNMENT>(s, ids, p, local, g, H); break;
    default: break;
    }
    for (int k = 0; k < 6; ++k) out42[k] = -g[k];  // force
    for (int k = 0; k < 36; ++k) out42[6 + k] = H[k];
}
}  // namespace

struct RbEnergyState {
    int kind = 0, S = 0, NP = 0, n = 0;
    std::vector<int32_t> stencils;  // body ids, caller's order
    std::vector<double> records;
    std::vector<int32_t> stencilColors, order, positionOf;
    std::vector<int64_t> colorOffsets;
    DevBuf<int32_t> d_stencils;
    DevBuf<double> d_records;
    DevBuf<uint32_t> d_rowptr, d_entries;
};
}  // namespace spgx

using namespace spgx;

struct spgx_rb_group {
    spgx_ctx *ctx = nullptr;
    int n = 0;
    std::vector<double> xInitial, thetaInitial, mass, localI, localIinv;
    std::vector<double> xPending, thetaPending, vPending, omegaPending;
    std::vector<RbEnergyState> energies;
    int solverKind = -1;
    bool finalized = false;
    std::vector<int32_t> bodyColors, order;
    std::vector<int64_t> colorOffsets;
    DevBuf<int32_t> d_order;
    DevBuf<double> d_x, d_theta, d_v, d_omega, d_R, d_inertia, d_inertiaInv, d_xOld, d_thetaOld, d_xInertial, d_mass, d_w, d_localI,
        d_localIinv, d_eval;
    cudaGraphExec_t graphExec = nullptr;
    struct Key {
        double dt = 0, g[3] = {0, 0, 0};
        int nsub = 0, iters = 0;
        bool operator==(const Key &o) const
        {
            return dt == o.dt && nsub == o.nsub && iters == o.iters && g[0] == o.g[0] && g[1] == o.g[1] && g[2] == o.g[2];
        }
    } key;
    int64_t launches = 0;
    ~spgx_rb_group()
    {
        if (graphExec) cudaGraphExecDestroy(graphExec);
    }
    RbState state() const
    {
        RbState s{};
        s.x = d_x.ptr, s.theta = d_theta.ptr, s.v = d_v.ptr, s.omega = d_omega.ptr, s.R = d_R.ptr, s.inertia = d_inertia.ptr;
        s.inertiaInv = d_inertiaInv.ptr, s.xOld = d_xOld.ptr, s.thetaOld = d_thetaOld.ptr, s.xInertial = d_xInertial.ptr;
        s.mass = d_mass.ptr, s.w = d_w.ptr, s.localI = d_localI.ptr, s.localIinv = d_localIinv.ptr;
        s.n = n;
        return s;
    }
};

namespace
{
void inverse3(const double *m, double *inv)
{
    const double c00 = m[4] * m[8] - m[5] * m[7], c01 = m[5] * m[6] - m[3] * m[8], c02 = m[3] * m[7] - m[4] * m[6];
    const double det = m[0] * c00 + m[1] * c01 + m[2] * c02;
    const double id = 1.0 / det;
    inv[0] = c00 * id, inv[1] = (m[2] * m[7] - m[1] * m[8]) * id, inv[2] = (m[1] * m[5] - m[2] * m[4]) * id;
    inv[3] = c01 * id, inv[4] = (m[0] * m[8] - m[2] * m[6]) * id, inv[5] = (m[2] * m[3] - m[0] * m[5]) * id;
    inv[6] = c02 * id, inv[7] = (m[1] * m[6] - m[0] * m[7]) * id, inv[8] = (m[0] * m[4] - m[1] * m[3]) * id;
}

// addStencil of the four energies: validation + model -> record
void rbPrepare(int kind, const double *model, double *p)
{
    switch (kind) {
    case RB_SPRING:  // rigidBodySpringEnergy.cpp:36-51
        if (model[6] <= 0) throw std::runtime_error("Zero/negative rest length invalid");
        std::copy(model, model + 6, p);
        p[6] = model[6], p[7] = model[7], p[8] = 1.0 / model[7];
        break;
    case RB_SPRING_SQUARED:  // rigidBodySpringSquaredEnergy.cpp:35-50
        if (model[6] < 0) throw std::runtime_error("Negative rest length invalid");
        std::copy(model, model + 6, p);
        p[6] = model[6] * model[6], p[7] = model[7], p[8] = 1.0 / model[7];
        break;
    case RB_SPRING_ANCHOR:  // rigidBodySpringAnchorEnergy.cpp:68-81
        std::copy(model, model + 6, p);
        p[6] = model[6], p[7] = 1.0 / model[6];
        break;
    case RB_ORIENTATION_ALIGNMENT:  // rigidBodyOrientationAlignmentEnergy.cpp:30-37
        p[0] = model[0], p[1] = 1.0 / model[0];
        break;
    default: throw std::runtime_error("unknown rigid-body energy kind");
    }
}

int blocksFor(int n)
{
    return (n + 63) / 64;
}

void enqueueRbFrame(spgx_rb_group *grp, double dtStep, int nsub, int iters, const double *g, cudaStream_t st)
{
    const double dt = dtStep / nsub, invdt = 1.0 / dt;
    const RbState s = grp->state();
    const bool vbd = grp->solverKind == SPGX_VBD;
    RbVbdArgs va{};
    va.s = s;
    va.nEnergies = static_cast<int>(grp->energies.size());
    for (int e = 0; e < va.nEnergies; ++e) {
        auto &en = grp->energies[e];
        va.e[e] = {en.kind, en.n, en.d_stencils.ptr, en.d_records.ptr, en.d_rowptr.ptr, en.d_entries.ptr};
    }
    va.order = grp->d_order.ptr;
    for (int sub = 0; sub < nsub; ++sub) {
        rbPredictKernel<<<blocksFor(grp->n), 64, 0, st>>>(s, dt, g[0], g[1], g[2]);
        grp->launches++;
        for (int it = 0; it < iters; ++it) {
            if (vbd) {
                for (size_t c = 0; c + 1 < grp->colorOffsets.size(); ++c) {
                    const int b = static_cast<int>(grp->colorOffsets[c]), e = static_cast<int>(grp->colorOffsets[c + 1]);
                    if (e > b) {
                        rbVbdPassKernel<<<blocksFor(e - b), 64, 0, st>>>(va, b, e, invdt);
                        grp->launches++;
                    }
                }
            } else {
                const int serial = grp->solverKind == SPGX_XPBD_SERIAL ? 1 : 0;
                for (auto &en : grp->energies)
                    for (size_t c = 0; c + 1 < en.colorOffsets.size(); ++c) {
                        const int b = static_cast<int>(en.colorOffsets[c]), e = static_cast<int>(en.colorOffsets[c + 1]);
                        if (e > b) {
                            rbXpbdPassKernel<<<serial ? 1 : blocksFor(e - b), serial ? 1 : 64, 0, st>>>(s, en.kind, en.d_stencils.ptr,
                                                                                                      en.d_records.ptr, b, e, dt, serial);
                            grp->launches++;
                        }
                    }
            }
        }
        rbFinalizeKernel<<<blocksFor(grp->n), 64, 0, st>>>(s, invdt);
        grp->launches++;
    }
}

// rotation: when given, the matrices are restored as they are instead of being rebuilt from theta
void rbUploadState(spgx_rb_group *grp, const double *x, const double *theta, const double *v, const double *omega,
                   const double *rotation = nullptr)
{
    cudaStream_t st = grp->ctx->stream;
    const size_t bytes = 3 * static_cast<size_t>(grp->n) * sizeof(double);
    if (x) SPGX_CUDA(cudaMemcpyAsync(grp->d_x.ptr, x, bytes, cudaMemcpyHostToDevice, st));
    if (theta) SPGX_CUDA(cudaMemcpyAsync(grp->d_theta.ptr, theta, bytes, cudaMemcpyHostToDevice, st));
    if (v) SPGX_CUDA(cudaMemcpyAsync(grp->d_v.ptr, v, bytes, cudaMemcpyHostToDevice, st));
    if (omega) SPGX_CUDA(cudaMemcpyAsync(grp->d_omega.ptr, omega, bytes, cudaMemcpyHostToDevice, st));
    if (rotation) {
        SPGX_CUDA(cudaMemcpyAsync(grp->d_R.ptr, rotation, 3 * bytes, cudaMemcpyHostToDevice, st));
        if (grp->n > 0) rbInertiaKernel<<<blocksFor(grp->n), 64, 0, st>>>(grp->state());
    } else if (theta && grp->n > 0) {
        rbFromThetaKernel<<<blocksFor(grp->n), 64, 0, st>>>(grp->state());
    }
    SPGX_CUDA(cudaStreamSynchronize(st));
    SPGX_CUDA(cudaGetLastError());
}
}  // namespace

extern "C" {

int spgx_rb_kind_stencil_size(int kind)
{
    return rbKindStencilSize(kind);
}
int spgx_rb_kind_num_model_params(int kind)
{
    return rbKindNumModelParams(kind);
}

int spgx_rb_group_create(spgx_ctx *ctx, int nBodies, const double *x, const double *theta, const double *mass,
                         const double *localInertia, spgx_rb_group **out)
{
    return guarded([&] {
        if (!ctx || !out || nBodies < 0 || (nBodies > 0 && (!x || !theta || !mass || !localInertia)))
            throw std::runtime_error("invalid rigid-body group arguments");
        auto g = std::make_unique<spgx_rb_group>();
        g->ctx = ctx;
        g->n = nBodies;
        const size_t n3 = 3 * static_cast<size_t>(nBodies);
        g->xInitial.assign(x, x + n3);
        g->thetaInitial.assign(theta, theta + n3);
        g->mass.assign(mass, mass + nBodies);
        g->localI.assign(localInertia, localInertia + 9 * static_cast<size_t>(nBodies));
        g->localIinv.resize(g->localI.size());
        for (int i = 0; i < nBodies; ++i) inverse3(g->localI.data() + 9 * i, g->localIinv.data() + 9 * i);  // addBody, rigidBodyGroup.cpp:156
        g->xPending = g->xInitial;
        g->thetaPending = g->thetaInitial;
        g->vPending.assign(n3, 0.0);
        g->omegaPending.assign(n3, 0.0);
        *out = g.release();
    });
}

void spgx_rb_group_destroy(spgx_rb_group *grp)
{
    if (!grp) return;
    cudaSetDevice(grp->ctx->device);
    delete grp;
}

int spgx_rb_group_add_energy(spgx_rb_group *grp, int kind, int nStencils, const int32_t *stencils, const double *modelParams,
                             int *energyId)
{
    return guarded([&] {
        if (!grp) throw StateError("null group");
        if (kind < 0 || kind >= RB_KIND_COUNT) throw std::runtime_error("unknown rigid-body energy kind");
        if (static_cast<int>(grp->energies.size()) >= kRbMaxEnergies) throw std::runtime_error("too many energies in one rigid-body group");
        RbEnergyState e;
        e.kind = kind, e.S = rbKindStencilSize(kind), e.NP = rbKindNumParams(kind), e.n = nStencils;
        const int NM = rbKindNumModelParams(kind);
        e.stencils.assign(stencils, stencils + static_cast<size_t>(nStencils) * e.S);
        for (int32_t b : e.stencils)
            if (b < 0 || b >= grp->n) throw std::runtime_error("stencil body index out of range");
        e.records.resize(static_cast<size_t>(nStencils) * e.NP);
        for (int i = 0; i < nStencils; ++i) rbPrepare(kind, modelParams + static_cast<size_t>(i) * NM, e.records.data() + static_cast<size_t>(i) * e.NP);
        grp->energies.push_back(std::move(e));
        grp->finalized = false;
        if (energyId) *energyId = static_cast<int>(grp->energies.size()) - 1;
    });
}

int spgx_rb_group_remove_energy(spgx_rb_group *grp, int energyId)
{
    return guarded([&] {
        if (!grp || energyId < 0 || energyId >= static_cast<int>(grp->energies.size())) throw std::runtime_error("energy id out of range");
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        SPGX_CUDA(cudaStreamSynchronize(grp->ctx->stream));
        grp->energies.erase(grp->energies.begin() + energyId);
        grp->finalized = false;
    });
}

// RBSpringAnchorEnergy::updateAnchor (rigidBodySpringAnchorEnergy.h:17): takes effect at the next step
int spgx_rb_energy_update_anchor(spgx_rb_group *grp, int energyId, int stencil, const double *anchor3)
{
    return guarded([&] {
        auto &e = grp->energies.at(energyId);
        if (e.kind != RB_SPRING_ANCHOR || stencil < 0 || stencil >= e.n) throw std::runtime_error("not an anchor stencil");
        std::copy(anchor3, anchor3 + 3, e.records.begin() + static_cast<size_t>(stencil) * e.NP + 3);
        if (grp->finalized) {
            SPGX_CUDA(cudaSetDevice(grp->ctx->device));
            const int pos = e.positionOf.empty() ? stencil : e.positionOf[stencil];
            SPGX_CUDA(cudaMemcpyAsync(e.d_records.ptr + static_cast<size_t>(pos) * e.NP + 3, anchor3, 3 * sizeof(double), cudaMemcpyHostToDevice,
                                      grp->ctx->stream));
            SPGX_CUDA(cudaStreamSynchronize(grp->ctx->stream));
        }
    });
}

int spgx_rb_group_finalize(spgx_rb_group *grp, int solverKind)
{
    return guarded([&] {
        if (!grp) throw StateError("null group");
        if (solverKind != SPGX_VBD && solverKind != SPGX_XPBD && solverKind != SPGX_XPBD_SERIAL)
            throw std::runtime_error("rigid-body groups run under VBD and XPBD");
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        cudaStream_t st = grp->ctx->stream;
        const int n = grp->n;
        const size_t n3 = 3 * static_cast<size_t>(n);
        // the state survives a re-finalize (live edits: addEnergy / removeEnergy on a stepping solver)
        std::vector<double> xh, th, vh, oh, Rh;
        if (grp->d_x.ptr) {
            xh.resize(n3), th.resize(n3), vh.resize(n3), oh.resize(n3), Rh.resize(3 * n3);
            SPGX_CUDA(cudaMemcpyAsync(Rh.data(), grp->d_R.ptr, 3 * n3 * sizeof(double), cudaMemcpyDeviceToHost, st));
            SPGX_CUDA(cudaMemcpyAsync(xh.data(), grp->d_x.ptr, n3 * sizeof(double), cudaMemcpyDeviceToHost, st));
            SPGX_CUDA(cudaMemcpyAsync(th.data(), grp->d_theta.ptr, n3 * sizeof(double), cudaMemcpyDeviceToHost, st));
            SPGX_CUDA(cudaMemcpyAsync(vh.data(), grp->d_v.ptr, n3 * sizeof(double), cudaMemcpyDeviceToHost, st));
            SPGX_CUDA(cudaMemcpyAsync(oh.data(), grp->d_omega.ptr, n3 * sizeof(double), cudaMemcpyDeviceToHost, st));
            SPGX_CUDA(cudaStreamSynchronize(st));
        } else {
            xh = grp->xPending, th = grp->thetaPending, vh = grp->vPending, oh = grp->omegaPending;
        }
        if (grp->graphExec) {
            cudaGraphExecDestroy(grp->graphExec);
            grp->graphExec = nullptr;
        }
        grp->solverKind = solverKind;
        std::vector<FlatStencilsView> sets;
        for (auto &e : grp->energies) sets.push_back({e.stencils.data(), static_cast<int64_t>(e.stencils.size()), e.S});
        if (solverKind == SPGX_VBD) {
            // computeParallelVertexGroups / computeStencilInfoPerVertex over bodies (vbd.cpp:203-268)
            std::vector<Incidence> inc;
            for (auto &fs : sets) inc.push_back(buildIncidence(n, fs));
            grp->bodyColors = n > 0 ? colorVertices(n, sets, &inc) : std::vector<int32_t>();
            groupByKey(grp->bodyColors, grp->order, grp->colorOffsets);
            grp->d_order.upload(grp->order, st);
            for (size_t ei = 0; ei < grp->energies.size(); ++ei) {
                auto &e = grp->energies[ei];
                std::vector<uint32_t> rowptr(static_cast<size_t>(n) + 1, 0), entries;
                for (int b = 0; b < n; ++b) {
                    rowptr[b] = static_cast<uint32_t>(entries.size());
                    for (int64_t q = inc[ei].rowptr[b]; q < inc[ei].rowptr[b + 1]; ++q) {
                        const int64_t flat = inc[ei].flat[q];
                        entries.push_back((static_cast<uint32_t>(flat / e.S) << 1) | static_cast<uint32_t>(flat % e.S));
                    }
                }
                rowptr[n] = static_cast<uint32_t>(entries.size());
                e.positionOf.clear();
                e.d_stencils.upload(e.stencils, st);
                e.d_records.upload(e.records, st);
                e.d_rowptr.upload(rowptr, st);
                e.d_entries.upload(entries, st);
                SPGX_CUDA(cudaStreamSynchronize(st));
            }
        } else {
            grp->bodyColors.clear();
            for (size_t ei = 0; ei < grp->energies.size(); ++ei) {
                auto &e = grp->energies[ei];
                if (solverKind == SPGX_XPBD) {
                    e.stencilColors = colorStencils(sets[ei]);  // xpbd.cpp:133-160
                    groupByKey(e.stencilColors, e.order, e.colorOffsets);
                } else {
                    e.stencilColors.assign(e.n, 0);
                    e.order.resize(e.n);
                    for (int i = 0; i < e.n; ++i) e.order[i] = i;
                    e.colorOffsets = {0, e.n};
                }
                e.positionOf.resize(e.n);
                std::vector<int32_t> st2(e.stencils.size());
                std::vector<double> rec2(e.records.size());
                for (int pos = 0; pos < e.n; ++pos) {
                    const int i = e.order[pos];
                    e.positionOf[i] = pos;
                    std::copy(e.stencils.begin() + static_cast<size_t>(i) * e.S, e.stencils.begin() + static_cast<size_t>(i + 1) * e.S,
                              st2.begin() + static_cast<size_t>(pos) * e.S);
                    std::copy(e.records.begin() + static_cast<size_t>(i) * e.NP, e.records.begin() + static_cast<size_t>(i + 1) * e.NP,
                              rec2.begin() + static_cast<size_t>(pos) * e.NP);
                }
                e.d_stencils.upload(st2, st);
                e.d_records.upload(rec2, st);
                e.d_rowptr.release();
                e.d_entries.release();
                SPGX_CUDA(cudaStreamSynchronize(st));
            }
        }
        for (DevBuf<double> *b : {&grp->d_x, &grp->d_theta, &grp->d_v, &grp->d_omega, &grp->d_xOld, &grp->d_thetaOld, &grp->d_xInertial})
            b->alloc(std::max<size_t>(n3, 1));
        for (DevBuf<double> *b : {&grp->d_R, &grp->d_inertia, &grp->d_inertiaInv}) b->alloc(std::max<size_t>(3 * n3, 1));
        std::vector<double> w(static_cast<size_t>(n));
        for (int i = 0; i < n; ++i) w[i] = grp->mass[i] == 0 ? 0 : 1 / grp->mass[i];  // rigidBodyGroup.cpp:151
        grp->d_mass.upload(grp->mass, st);
        grp->d_w.upload(w, st);
        grp->d_localI.upload(grp->localI, st);
        grp->d_localIinv.upload(grp->localIinv, st);
        grp->d_eval.alloc(64);
        SPGX_CUDA(cudaStreamSynchronize(st));
        grp->finalized = true;
        rbUploadState(grp, xh.data(), th.data(), vh.data(), oh.data(), Rh.empty() ? nullptr : Rh.data());
        grp->xPending.clear(), grp->thetaPending.clear(), grp->vPending.clear(), grp->omegaPending.clear();
    });
}

int spgx_rb_step(spgx_rb_group *grp, double dt, int nsubsteps, int iterations, const double *gravity, int nFrames)
{
    return guarded([&] {
        if (!grp || !grp->finalized) throw StateError("rigid-body group is not finalized (call spgx_rb_group_finalize first)");
        if (nsubsteps < 1 || iterations < 1 || nFrames < 0) throw std::runtime_error("invalid step parameters");
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        cudaStream_t st = grp->ctx->stream;
        static const double defaultG[3] = {0, -9.8, 0};  // baseSolver.h:39
        const double *g = gravity ? gravity : defaultG;
        spgx_rb_group::Key key;
        key.dt = dt, key.nsub = nsubsteps, key.iters = iterations, key.g[0] = g[0], key.g[1] = g[1], key.g[2] = g[2];
        if (grp->n == 0) return;
        if (!grp->graphExec || !(key == grp->key)) {
            if (grp->graphExec) {
                cudaGraphExecDestroy(grp->graphExec);
                grp->graphExec = nullptr;
            }
            grp->key = key;
            grp->launches = 0;
            cudaGraph_t graph = nullptr;
            SPGX_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
            try {
                enqueueRbFrame(grp, dt, nsubsteps, iterations, g, st);
            } catch (...) {
                cudaStreamEndCapture(st, &graph);
                if (graph) cudaGraphDestroy(graph);
                throw;
            }
            SPGX_CUDA(cudaStreamEndCapture(st, &graph));
            cudaError_t err = cudaGraphInstantiate(&grp->graphExec, graph, 0);
            cudaGraphDestroy(graph);
            if (err != cudaSuccess) throw CudaError(std::string("cudaGraphInstantiate: ") + cudaGetErrorString(err));
        }
        for (int f = 0; f < nFrames; ++f) SPGX_CUDA(cudaGraphLaunch(grp->graphExec, st));
        SPGX_CUDA(cudaStreamSynchronize(st));
        SPGX_CUDA(cudaGetLastError());
    });
}

int spgx_rb_download(spgx_rb_group *grp, double *x, double *theta, double *v, double *omega, double *rotation)
{
    return guarded([&] {
        if (!grp) throw StateError("null group");
        const size_t n3 = 3 * static_cast<size_t>(grp->n);
        if (!grp->d_x.ptr) {
            if (x) std::copy(grp->xPending.begin(), grp->xPending.end(), x);
            if (theta) std::copy(grp->thetaPending.begin(), grp->thetaPending.end(), theta);
            if (v) std::copy(grp->vPending.begin(), grp->vPending.end(), v);
            if (omega) std::copy(grp->omegaPending.begin(), grp->omegaPending.end(), omega);
            if (rotation) throw StateError("rotation matrices exist once the group is finalized");
            return;
        }
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        cudaStream_t st = grp->ctx->stream;
        if (x) SPGX_CUDA(cudaMemcpyAsync(x, grp->d_x.ptr, n3 * sizeof(double), cudaMemcpyDeviceToHost, st));
        if (theta) SPGX_CUDA(cudaMemcpyAsync(theta, grp->d_theta.ptr, n3 * sizeof(double), cudaMemcpyDeviceToHost, st));
        if (v) SPGX_CUDA(cudaMemcpyAsync(v, grp->d_v.ptr, n3 * sizeof(double), cudaMemcpyDeviceToHost, st));
        if (omega) SPGX_CUDA(cudaMemcpyAsync(omega, grp->d_omega.ptr, n3 * sizeof(double), cudaMemcpyDeviceToHost, st));
        if (rotation) SPGX_CUDA(cudaMemcpyAsync(rotation, grp->d_R.ptr, 3 * n3 * sizeof(double), cudaMemcpyDeviceToHost, st));
        SPGX_CUDA(cudaStreamSynchronize(st));
    });
}

int spgx_rb_upload(spgx_rb_group *grp, const double *x, const double *theta, const double *v, const double *omega)
{
    return guarded([&] {
        if (!grp) throw StateError("null group");
        const size_t n3 = 3 * static_cast<size_t>(grp->n);
        if (!grp->finalized) {
            if (x) grp->xPending.assign(x, x + n3);
            if (theta) grp->thetaPending.assign(theta, theta + n3);
            if (v) grp->vPending.assign(v, v + n3);
            if (omega) grp->omegaPending.assign(omega, omega + n3);
            return;
        }
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        rbUploadState(grp, x, theta, v, omega);
    });
}

// RigidBodyGroup::reset (rigidBodyGroup.cpp:180-192)
int spgx_rb_reset(spgx_rb_group *grp)
{
    return guarded([&] {
        if (!grp) throw StateError("null group");
        std::vector<double> zeros(3 * static_cast<size_t>(grp->n), 0.0);
        if (spgx_rb_upload(grp, grp->xInitial.data(), grp->thetaInitial.data(), zeros.data(), zeros.data()) != SPGX_OK)
            throw CudaError(spgx_last_error());
    });
}

int spgx_rb_get_body_colors(spgx_rb_group *grp, int32_t *colors)
{
    return guarded([&] {
        if (!grp || !grp->finalized || grp->solverKind != SPGX_VBD) throw StateError("rigid-body group is not finalized for VBD");
        std::copy(grp->bodyColors.begin(), grp->bodyColors.end(), colors);
    });
}

int spgx_rb_get_stencil_colors(spgx_rb_group *grp, int energyId, int32_t *colors)
{
    return guarded([&] {
        if (!grp || !grp->finalized || grp->solverKind == SPGX_VBD) throw StateError("rigid-body group is not finalized for XPBD");
        const auto &e = grp->energies.at(energyId);
        std::copy(e.stencilColors.begin(), e.stencilColors.end(), colors);
    });
}

// constraintsAndGradient of one stencil on the current device state: C (1) and dC/dq (6 S)
int spgx_rb_eval_constraint(spgx_rb_group *grp, int energyId, int stencil, double *C, double *grad)
{
    return guarded([&] {
        if (!grp || !grp->finalized) throw StateError("rigid-body group is not finalized");
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        auto &e = grp->energies.at(energyId);
        if (stencil < 0 || stencil >= e.n) throw std::runtime_error("stencil out of range");
        const int pos = e.positionOf.empty() ? stencil : e.positionOf[stencil];
        cudaStream_t st = grp->ctx->stream;
        rbEvalConstraintKernel<<<1, 1, 0, st>>>(grp->state(), e.kind, e.d_stencils.ptr + static_cast<size_t>(pos) * e.S,
                                                e.d_records.ptr + static_cast<size_t>(pos) * e.NP, grp->d_eval.ptr);
        double out[13];
        SPGX_CUDA(cudaMemcpyAsync(out, grp->d_eval.ptr, sizeof(out), cudaMemcpyDeviceToHost, st));
        SPGX_CUDA(cudaStreamSynchronize(st));
        SPGX_CUDA(cudaGetLastError());
        *C = out[0];
        std::copy(out + 1, out + 1 + 6 * e.S, grad);
    });
}

// accumulateVertexForce / accumulateVertexHessian of one (stencil, body) incidence: force (6), Hessian block (6 x 6 row-major)
int spgx_rb_eval_vertex(spgx_rb_group *grp, int energyId, int stencil, int local, double *force6, double *hess36)
{
    return guarded([&] {
        if (!grp || !grp->finalized) throw StateError("rigid-body group is not finalized");
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        auto &e = grp->energies.at(energyId);
        if (stencil < 0 || stencil >= e.n || local < 0 || local >= e.S) throw std::runtime_error("stencil out of range");
        const int pos = e.positionOf.empty() ? stencil : e.positionOf[stencil];
        cudaStream_t st = grp->ctx->stream;
        rbEvalVertexKernel<<<1, 1, 0, st>>>(grp->state(), e.kind, e.d_stencils.ptr + static_cast<size_t>(pos) * e.S,
                                            e.d_records.ptr + static_cast<size_t>(pos) * e.NP, local, grp->d_eval.ptr);
        double out[42];
        SPGX_CUDA(cudaMemcpyAsync(out, grp->d_eval.ptr, sizeof(out), cudaMemcpyDeviceToHost, st));
        SPGX_CUDA(cudaStreamSynchronize(st));
        SPGX_CUDA(cudaGetLastError());
        std::copy(out, out + 6, force6);
        std::copy(out + 6, out + 42, hess36);
    });
}

}  // extern "C"
