// XPBD per-colour constraint projection (reference: spg/source/sim/solver/xpbd.cpp:82-104, generic block
// projection spg/include/spg/sim/energy/energy.h:117-153, SpringEnergy override springEnergy.cpp:48-68).
//
// One thread owns one stencil of the colour: it gathers the S positions and inverse masses, evaluates C
// and dC/dx with first-order duals over the 3S stencil DOFs, forms A = gradC^T W gradC + alpha/dt^2 + 1e-10 I
// (W is diagonal for particles, so the dense W products of the reference collapse), solves the C x C system
// in registers and scatters x_s += w_s gradC_s dlambda.  Stencils of one colour share no vertex, so the
// scatter is race free.  With `lambda` the full XPBD right-hand side -C - alpha~ lambda is used and lambda is
// accumulated on device; without it this is the reference's "small steps" form (dlambda = -A^-1 C).
#include <algorithm>
#include <cstdlib>

#include "energies.cuh"
#include "halo.cuh"
#include "kernels.h"
#include "streams.cuh"

namespace spgx
{
namespace
{
// the generic block projection (energy.h:117-153): C and dC/dx with first-order duals over the 3S stencil DOFs
template <int KIND>
__device__ __forceinline__ void projectGeneric(const double *w, V3d *px, const double *p, double *__restrict__ lambda, int i, double dt2)
{
    using EK = EnergyKind<KIND>;
    constexpr int S = EK::S, NC = EK::NC, K = 3 * S;
    using D1 = Dual<K, false>;
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
        px[s] = V3d(px[s].x + d[0], px[s].y + d[1], px[s].z + d[2]);
    }
}

// What a projection computes from the inverse masses and the record alone (no position): requested ahead of the pass's wait
// for the previous pass, so that the divisions are off the dependent chain gather -> projection -> store that bounds a pass
// (config 2: 540 passes a frame, each one that chain).
struct ProjectPre {
    double v[2];
};
template <int KIND>
__device__ __forceinline__ ProjectPre projectPrepare(const double *w, const double *p, double dt, bool handProject)
{
    using EK = EnergyKind<KIND>;
    ProjectPre pre{{0.0, 0.0}};
    const double dt2 = dt * dt;
    if constexpr (KIND == K_SPRING) {
        const double alphaTilde = p[2] / dt2;
        pre.v[0] = alphaTilde;
        pre.v[1] = 1.0 / (w[0] + w[1] + alphaTilde);
    } else if constexpr (HasHandProject<EK>::value) {
        if (handProject) EK::projectPrepare(w, p, dt2, pre.v);
    }
    return pre;
}

// The projection of one stencil on operands that are already in registers: px <- the projected positions.  Returns false
// when the stencil moves nothing (the spring's early-outs).  Shared by the per-pass kernels and the single-thread serial path.
template <int KIND>
__device__ __forceinline__ bool projectLoaded(V3d *px, const double *w, const double *p, const ProjectPre &pre, double *__restrict__ lambda,
                                              int i, double dt, bool handProject)
{
    using EK = EnergyKind<KIND>;
    constexpr int NC = EK::NC;
    const double dt2 = dt * dt;
    if constexpr (KIND == K_SPRING) {
        // springEnergy.cpp:48-68 (early-outs at 1e-6, no +1e-10 regularisation).  The reference divides by the length and by
        // w0 + w1 + alpha~; here the length comes from one reciprocal square root (d = |n|^2 / sqrt(|n|^2), n / d = n rsqrt) and the
        // second denominator's reciprocal from projectPrepare: the same values to a unit in the last place, a dependent chain
        // of ~15 instead of ~60 FP64 instructions behind the gather
        constexpr double eps = 1.0e-6;
        const double wSum = w[0] + w[1];
        if (wSum < eps) return false;
        const double alphaTilde = pre.v[0], invDen = pre.v[1];
        V3d n = px[1] - px[0];
        const double dd = dot(n, n);
        if (dd < eps * eps) return false;  // d < eps
        const double invd = fastRsqrt(dd);
        const double d = dd * invd;
        n = V3d(n.x * invd, n.y * invd, n.z * invd);
        double c = d - p[0];
        if (lambda) {
            const double lam = lambda[i];
            const double dl = (-c - alphaTilde * lam) * invDen;
            lambda[i] = lam + dl;
            c = -dl * (wSum + alphaTilde);
        }
        const double scale = c * invDen;
        px[0] = V3d(px[0].x + w[0] * (n.x * scale), px[0].y + w[0] * (n.y * scale), px[0].z + w[0] * (n.z * scale));
        px[1] = V3d(px[1].x - w[1] * (n.x * scale), px[1].y - w[1] * (n.y * scale), px[1].z - w[1] * (n.z * scale));
        return true;
    } else if constexpr (HasHandProject<EK>::value) {
        if (handProject)
            EK::projectPrepared(px, w, p, pre.v, lambda ? lambda + static_cast<size_t>(i) * NC : nullptr);
        else
            projectGeneric<KIND>(w, px, p, lambda, i, dt2);
        return true;
    } else {
        projectGeneric<KIND>(w, px, p, lambda, i, dt2);
        return true;
    }
}

// PDL: the caller was launched with programmatic stream serialization; the stencil's vertex ids and inverse masses (never
// written by a pass) are fetched before griddepcontrol.wait, i.e. while the previous pass may still be draining
template <int KIND, bool PDL = false>
__device__ __forceinline__ void projectStencil(double *__restrict__ x, const double *__restrict__ stencilW,
                                               const int32_t *__restrict__ stencils, const double *__restrict__ records,
                                               double *__restrict__ lambda, int i, double dt, bool handProject = true)
{
    using EK = EnergyKind<KIND>;
    constexpr int S = EK::S;
    const int32_t *ids = stencils + static_cast<size_t>(i) * S;
    int id[S];
    double w[S];
    V3d px[S];
    // everything a pass never writes -- vertex ids, inverse masses, the stencil's record -- is requested up front (under PDL:
    // before the dependency wait, i.e. while the previous pass may still be draining)
    double p[EK::NP];
#pragma unroll
    for (int k = 0; k < EK::NP; ++k) p[k] = __ldg(records + static_cast<size_t>(i) * EK::NP + k);
#pragma unroll
    for (int s = 0; s < S; ++s) {
        id[s] = __ldg(ids + s);
        w[s] = __ldg(stencilW + static_cast<size_t>(i) * S + s);
    }
    const ProjectPre pre = projectPrepare<KIND>(w, p, dt, handProject);
    if constexpr (PDL) asm volatile("griddepcontrol.wait;" ::: "memory");
#pragma unroll
    for (int s = 0; s < S; ++s) {
        const size_t o = 3 * static_cast<size_t>(id[s]);
        px[s] = V3d(x[o], x[o + 1], x[o + 2]);
    }
    if (!projectLoaded<KIND>(px, w, p, pre, lambda, i, dt, handProject)) return;
#pragma unroll
    for (int s = 0; s < S; ++s) {
        const size_t o = 3 * static_cast<size_t>(id[s]);
        x[o] = px[s].x;
        x[o + 1] = px[s].y;
        x[o + 2] = px[s].z;
    }
}

template <int KIND>
__global__ void __launch_bounds__(128) xpbdPassKernel(double *x, const double *stencilW, const int32_t *stencils,
                                                      const double *records, double *lambda, int begin, int end,
                                                      double dt, bool handProject, const VbdHaloDev h)
{
    if (h.nPeers > 0) {
        if (threadIdx.x == 0) haloWaitAll(h, haloEpochBase(h) + static_cast<uint32_t>(h.event));
        __syncthreads();
    }
    const int i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i < end) projectStencil<KIND>(x, stencilW, stencils, records, lambda, i, dt, handProject);
}
// the same pass under programmatic dependent launch: the next pass may become resident and run its prologue while this
// one drains (config 2 is 540 passes of 10-30 k stencils per frame: the kernel boundary is most of a pass)
template <int KIND>
__global__ void __launch_bounds__(128) xpbdPassKernelPdl(double *x, const double *stencilW, const int32_t *stencils,
                                                         const double *records, double *lambda, int begin, int end,
                                                         double dt, bool handProject)
{
    asm volatile("griddepcontrol.launch_dependents;");
    const int i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i < end) projectStencil<KIND, true>(x, stencilW, stencils, records, lambda, i, dt, handProject);
    // threads past the end leave without waiting: the grid as a whole still completes only after its waiting threads
}

__global__ void xpbdEmptyKernel(double *) {}

// serial Gauss-Seidel path (xpbd.cpp:96-102): one thread walks the stencils in index order
template <int KIND>
__global__ void xpbdSerialKernel(double *x, const double *stencilW, const int32_t *stencils, const double *records,
                                 double *lambda, int begin, int end, double dt)
{
    for (int i = begin; i < end; ++i) {
        projectStencil<KIND>(x, stencilW, stencils, records, lambda, i, dt);
        __threadfence();
    }
}

template <int KIND>
void launchKind(double *x, const double *stencilW, const int32_t *stencils, const double *records, double *lambda,
                int begin, int end, double dt, bool serial, cudaStream_t stream, const VbdHaloDev *halo)
{
    const int n = end - begin;
    if (n <= 0) return;
    const VbdHaloDev h = (halo && halo->nPeers > 0) ? *halo : VbdHaloDev{};
    // SPGX_XPBD_HAND=0: kinds with a hand-derived projection (EnergyKind<>::project) take the generic dual-number path instead
    static const bool hand = !(getenv("SPGX_XPBD_HAND") && atoi(getenv("SPGX_XPBD_HAND")) == 0);
    if (serial) {
        xpbdSerialKernel<KIND><<<1, 1, 0, stream>>>(x, stencilW, stencils, records, lambda, begin, end, dt);
    } else {
        // measurement knobs: SPGX_XPBD_THREADS (block size, <= 128), SPGX_XPBD_DEBUG_EMPTY=1 (same grids, empty kernels: the
        // launch floor of the pass sequence; results are wrong)
        // one warp per block: a pass of 10-30 k stencils is one partial wave either way, and 4 x as many blocks spread it over
        // every SM (config 2: 1.59 / 1.52 / 1.45 ms per frame at 128 / 64 / 32 threads, profiles/r2m_xpbd_sweep_config2.txt)
        static const int threads = getenv("SPGX_XPBD_THREADS") ? std::min(128, std::max(32, atoi(getenv("SPGX_XPBD_THREADS")))) : 32;
        static const bool empty = getenv("SPGX_XPBD_DEBUG_EMPTY") && atoi(getenv("SPGX_XPBD_DEBUG_EMPTY")) != 0;
        if (empty) {
            xpbdEmptyKernel<<<(n + threads - 1) / threads, threads, 0, stream>>>(x);
            return;
        }
        // programmatic dependent launch: the next pass is resident and has its ids, inverse masses and record in registers
        // when this one completes (config 2: 1.45 -> 1.41 ms per frame at 32 threads; SPGX_XPBD_PDL=0 turns it off)
        static const bool pdl = !(getenv("SPGX_XPBD_PDL") && atoi(getenv("SPGX_XPBD_PDL")) == 0);
        if (pdl && h.nPeers == 0) {
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3((n + threads - 1) / threads);
            cfg.blockDim = dim3(threads);
            cfg.stream = stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            cfg.numAttrs = 1;
            cudaLaunchKernelEx(&cfg, xpbdPassKernelPdl<KIND>, x, stencilW, stencils, records, lambda, begin, end, dt, hand);
        } else {
            xpbdPassKernel<KIND><<<(n + threads - 1) / threads, threads, 0, stream>>>(x, stencilW, stencils, records, lambda,
                                                                                     begin, end, dt, hand, h);
        }
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
#define SPGX_KIND(NAME, ID, BYTES)                                                                                     \
    case NAME:                                                                                                         \
        evalConstraintsOne<NAME>(x, ids, rec, outC, outGrad);                     \
        break;
#include "kinds.def"
#undef SPGX_KIND
    default: break;
    }
}
}  // namespace

void launchXpbdPass(int kind, double *x, const double *stencilW, const int32_t *stencils, const double *records,
                    double *lambda, int begin, int end, double dt, bool serial, cudaStream_t stream, const VbdHaloDev *halo)
{
    switch (kind) {
        // (a static load has no constraint form: its body is a stub and spgx_xpbd_step refuses it, staticLoadEnergy.cpp:8-13)
#define SPGX_KIND(NAME, ID, BYTES)                                                                                     \
    case NAME:                                                                                                         \
        launchKind<NAME>(x, stencilW, stencils, records, lambda, begin, end, dt, serial, stream, halo); \
        break;
#include "kinds.def"
#undef SPGX_KIND
    default: break;
    }
}

void launchEvalConstraints(int kind, const double *x, const int32_t *stencil, const double *record, double *outC,
                           double *outGrad, cudaStream_t stream)
{
    evalConstraintsKernel<<<1, 1, 0, stream>>>(kind, x, stencil, record, outC, outGrad);
}
}  // namespace spgx
