// XPBD per-colour constraint projection (reference: spg/source/sim/solver/xpbd.cpp:82-104, generic block
// projection spg/include/spg/sim/energy/energy.h:117-153, SpringEnergy override springEnergy.cpp:48-68).
//
// One thread owns one stencil of the colour: it gathers the S positions and inverse masses, evaluates C
// and dC/dx with first-order duals over the 3S stencil DOFs, forms A = gradC^T W gradC + alpha/dt^2 + 1e-10 I
// (W is diagonal for particles, so the dense W products of the reference collapse), solves the C x C system
// in registers and scatters x_s += w_s gradC_s dlambda.  Stencils of one colour share no vertex, so the
// scatter is race free.  With `lambda` the full XPBD right-hand side -C - alpha~ lambda is used and lambda is
// accumulated on device; without it this is the reference's "small steps" form (dlambda = -A^-1 C).
#include "energies.cuh"
#include "kernels.h"

namespace spgx
{
namespace
{
template <int KIND>
__device__ __forceinline__ void projectStencil(double *__restrict__ x, const double *__restrict__ invMass,
                                               const int32_t *__restrict__ stencils, const double *__restrict__ records,
                                               double *__restrict__ lambda, int i, double dt)
{
    using EK = EnergyKind<KIND>;
    constexpr int S = EK::S, NC = EK::NC, K = 3 * S;
    using D1 = Dual<K, false>;
    const int32_t *ids = stencils + static_cast<size_t>(i) * S;
    const double *p = records + static_cast<size_t>(i) * EK::NP;
    int id[S];
    double w[S];
    V3d px[S];
#pragma unroll
    for (int s = 0; s < S; ++s) {
        id[s] = ids[s];
        w[s] = invMass[id[s]];
        const size_t o = 3 * static_cast<size_t>(id[s]);
        px[s] = V3d(x[o], x[o + 1], x[o + 2]);
    }
    const double dt2 = dt * dt;

    if constexpr (KIND == K_SPRING) {
        // springEnergy.cpp:48-68 (early-outs at 1e-6, no +1e-10 regularisation)
        constexpr double eps = 1.0e-6;
        const double wSum = w[0] + w[1];
        if (wSum < eps) return;
        const double alphaTilde = p[2] / dt2;
        V3d n = px[1] - px[0];
        const double d = ::sqrt(dot(n, n));
        if (d < eps) return;
        n = V3d(n.x / d, n.y / d, n.z / d);
        double c = d - p[0];
        if (lambda) {
            const double lam = lambda[i];
            const double dl = (-c - alphaTilde * lam) / (wSum + alphaTilde);
            lambda[i] = lam + dl;
            c = -dl * (wSum + alphaTilde);
        }
        const double scale = c / (wSum + alphaTilde);
        const size_t o0 = 3 * static_cast<size_t>(id[0]), o1 = 3 * static_cast<size_t>(id[1]);
        x[o0] = px[0].x + w[0] * (n.x * scale);
        x[o0 + 1] = px[0].y + w[0] * (n.y * scale);
        x[o0 + 2] = px[0].z + w[0] * (n.z * scale);
        x[o1] = px[1].x - w[1] * (n.x * scale);
        x[o1 + 1] = px[1].y - w[1] * (n.y * scale);
        x[o1 + 2] = px[1].z - w[1] * (n.z * scale);
        return;
    } else {
        V3<D1> xs[S];
#pragma unroll
        for (int s = 0; s < S; ++s)
            xs[s] = V3<D1>(D1(px[s].x, 3 * s), D1(px[s].y, 3 * s + 1), D1(px[s].z, 3 * s + 2));
        D1 C[NC];
        EK::constraints(xs, p, C);
        double A[NC][NC];
        double rhs[NC];
#pragma unroll
        for (int a = 0; a < NC; ++a) {
#pragma unroll
            for (int b = a; b < NC; ++b) {
                double acc = 0.0;
#pragma unroll
                for (int k = 0; k < K; ++k) acc += C[a].g[k] * (w[k / 3] * C[b].g[k]);
                acc += EK::alpha(p, a, b) / dt2;
                if (a == b) acc += 1e-10;
                A[a][b] = acc;
                A[b][a] = acc;
            }
            rhs[a] = -C[a].v;
        }
        if (lambda) {
#pragma unroll
            for (int a = 0; a < NC; ++a)
#pragma unroll
                for (int b = 0; b < NC; ++b) rhs[a] -= (EK::alpha(p, a, b) / dt2) * lambda[static_cast<size_t>(i) * NC + b];
        }
        solveSmall<NC>(A, rhs);  // rhs <- dlambda
        if (lambda) {
#pragma unroll
            for (int a = 0; a < NC; ++a) lambda[static_cast<size_t>(i) * NC + a] += rhs[a];
        }
#pragma unroll
        for (int s = 0; s < S; ++s) {
            double d[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                double acc = 0.0;
#pragma unroll
                for (int a = 0; a < NC; ++a) acc += (w[s] * C[a].g[3 * s + c]) * rhs[a];
                d[c] = acc;
            }
            const size_t o = 3 * static_cast<size_t>(id[s]);
            x[o] = px[s].x + d[0];
            x[o + 1] = px[s].y + d[1];
            x[o + 2] = px[s].z + d[2];
        }
    }
}

template <int KIND>
__global__ void __launch_bounds__(128) xpbdPassKernel(double *x, const double *invMass, const int32_t *stencils,
                                                      const double *records, double *lambda, int begin, int end,
                                                      double dt)
{
    const int i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i < end) projectStencil<KIND>(x, invMass, stencils, records, lambda, i, dt);
}

// serial Gauss-Seidel path (xpbd.cpp:96-102): one thread walks the stencils in index order
template <int KIND>
__global__ void xpbdSerialKernel(double *x, const double *invMass, const int32_t *stencils, const double *records,
                                 double *lambda, int begin, int end, double dt)
{
    for (int i = begin; i < end; ++i) {
        projectStencil<KIND>(x, invMass, stencils, records, lambda, i, dt);
        __threadfence();
    }
}

template <int KIND>
void launchKind(double *x, const double *invMass, const int32_t *stencils, const double *records, double *lambda,
                int begin, int end, double dt, bool serial, cudaStream_t stream)
{
    const int n = end - begin;
    if (n <= 0) return;
    if (serial) {
        xpbdSerialKernel<KIND><<<1, 1, 0, stream>>>(x, invMass, stencils, records, lambda, begin, end, dt);
    } else {
        const int threads = 128;
        xpbdPassKernel<KIND><<<(n + threads - 1) / threads, threads, 0, stream>>>(x, invMass, stencils, records, lambda,
                                                                                 begin, end, dt);
    }
}

template <int KIND>
__device__ void evalConstraintsOne(const double *x, const int32_t *ids, const double *rec, double *outC, double *outGrad)
{
    using EK = EnergyKind<KIND>;
    constexpr int S = EK::S, NC = EK::NC, K = 3 * S;
    using D1 = Dual<K, false>;
    V3<D1> xs[S];
    for (int s = 0; s < S; ++s) {
        const int id = ids[s];
        xs[s] = V3<D1>(D1(x[3 * id], 3 * s), D1(x[3 * id + 1], 3 * s + 1), D1(x[3 * id + 2], 3 * s + 2));
    }
    D1 C[NC];
    EK::constraints(xs, rec, C);
    for (int a = 0; a < NC; ++a) {
        outC[a] = C[a].v;
        for (int k = 0; k < K; ++k) outGrad[k * NC + a] = C[a].g[k];
    }
}

__global__ void evalConstraintsKernel(int kind, const double *x, const int32_t *ids, const double *rec, double *outC,
                                      double *outGrad)
{
    switch (kind) {
    case K_SPRING: evalConstraintsOne<K_SPRING>(x, ids, rec, outC, outGrad); break;
    case K_SPRING_CONTINUUM: evalConstraintsOne<K_SPRING_CONTINUUM>(x, ids, rec, outC, outGrad); break;
    case K_DISCRETE_BENDING: evalConstraintsOne<K_DISCRETE_BENDING>(x, ids, rec, outC, outGrad); break;
    case K_QUADRATIC_BENDING: evalConstraintsOne<K_QUADRATIC_BENDING>(x, ids, rec, outC, outGrad); break;
    case K_MEMBRANE_STVK: evalConstraintsOne<K_MEMBRANE_STVK>(x, ids, rec, outC, outGrad); break;
    case K_MEMBRANE_CHOI: evalConstraintsOne<K_MEMBRANE_CHOI>(x, ids, rec, outC, outGrad); break;
    case K_MEMBRANE_BW: evalConstraintsOne<K_MEMBRANE_BW>(x, ids, rec, outC, outGrad); break;
    case K_TET_STVK: evalConstraintsOne<K_TET_STVK>(x, ids, rec, outC, outGrad); break;
    case K_TET_SNH: evalConstraintsOne<K_TET_SNH>(x, ids, rec, outC, outGrad); break;
    case K_ANCHOR: evalConstraintsOne<K_ANCHOR>(x, ids, rec, outC, outGrad); break;
    case K_BENDING_BW: evalConstraintsOne<K_BENDING_BW>(x, ids, rec, outC, outGrad); break;
    case K_SPRING_SQUARED_CONTINUUM: evalConstraintsOne<K_SPRING_SQUARED_CONTINUUM>(x, ids, rec, outC, outGrad); break;
    case K_TET_SSNH: evalConstraintsOne<K_TET_SSNH>(x, ids, rec, outC, outGrad); break;
    default: break;
    }
}
}  // namespace

void launchXpbdPass(int kind, double *x, const double *invMass, const int32_t *stencils, const double *records,
                    double *lambda, int begin, int end, double dt, bool serial, cudaStream_t stream)
{
    switch (kind) {
    case K_SPRING: launchKind<K_SPRING>(x, invMass, stencils, records, lambda, begin, end, dt, serial, stream); break;
    case K_SPRING_CONTINUUM:
        launchKind<K_SPRING_CONTINUUM>(x, invMass, stencils, records, lambda, begin, end, dt, serial, stream);
        break;
    case K_DISCRETE_BENDING:
        launchKind<K_DISCRETE_BENDING>(x, invMass, stencils, records, lambda, begin, end, dt, serial, stream);
        break;
    case K_QUADRATIC_BENDING:
        launchKind<K_QUADRATIC_BENDING>(x, invMass, stencils, records, lambda, begin, end, dt, serial, stream);
        break;
    case K_MEMBRANE_STVK:
        launchKind<K_MEMBRANE_STVK>(x, invMass, stencils, records, lambda, begin, end, dt, serial, stream);
        break;
    case K_MEMBRANE_CHOI:
        launchKind<K_MEMBRANE_CHOI>(x, invMass, stencils, records, lambda, begin, end, dt, serial, stream);
        break;
    case K_MEMBRANE_BW:
        launchKind<K_MEMBRANE_BW>(x, invMass, stencils, records, lambda, begin, end, dt, serial, stream);
        break;
    case K_TET_STVK: launchKind<K_TET_STVK>(x, invMass, stencils, records, lambda, begin, end, dt, serial, stream); break;
    case K_TET_SNH: launchKind<K_TET_SNH>(x, invMass, stencils, records, lambda, begin, end, dt, serial, stream); break;
    case K_ANCHOR: launchKind<K_ANCHOR>(x, invMass, stencils, records, lambda, begin, end, dt, serial, stream); break;
    case K_BENDING_BW: launchKind<K_BENDING_BW>(x, invMass, stencils, records, lambda, begin, end, dt, serial, stream); break;
    case K_SPRING_SQUARED_CONTINUUM:
        launchKind<K_SPRING_SQUARED_CONTINUUM>(x, invMass, stencils, records, lambda, begin, end, dt, serial, stream);
        break;
    case K_TET_SSNH: launchKind<K_TET_SSNH>(x, invMass, stencils, records, lambda, begin, end, dt, serial, stream); break;
    default: break;
    }
}

void launchEvalConstraints(int kind, const double *x, const int32_t *stencil, const double *record, double *outC,
                           double *outGrad, cudaStream_t stream)
{
    evalConstraintsKernel<<<1, 1, 0, stream>>>(kind, x, stencil, record, outC, outGrad);
}
}  // namespace spgx
