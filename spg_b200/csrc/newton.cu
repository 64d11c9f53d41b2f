// Device side of the implicit-Euler family: system force vector, element Hessians, block-sparse
// Hessian assembly, and a device-resident preconditioned conjugate gradient.
//
// Reference: spg/source/sim/solver/implicitEulerBase.cpp:102-143 (getSystemForce), :176-221 (getSystemStiffnessMatrix),
// :223-238 (solveLinearSystem = Eigen::ConjugateGradient, external/eigen/Eigen/src/IterativeLinearSolvers/
// ConjugateGradient.h:29-87 with the DiagonalPreconditioner of BasicPreconditioners.h:62-73);
// spg/include/spg/sim/energy/energy.h:196-237 (accumulateForces, negativeHessianTriplets), :348-362 (energyGradient /
// energyHessian by autodiff).
//
// What changes on the GPU:
//  * forces are gathered per vertex over its incidence table with first-order duals seeded on the owner (no atomics,
//    fixed order) instead of scattered per stencil under `omp atomic`;
//  * the k x k element Hessian (k = 3S) is not produced by one 1 + k + k^2 scalar per operation: one thread per
//    (stencil, vertex pair) runs MIXED second-order duals (value, two 3-gradients, the 3 x 3 mixed partials: 16 doubles,
//    dual.cuh DualMixed) for the off-diagonal blocks, one thread per (stencil, vertex) second-order duals seeded on that
//    vertex (10 doubles) for the diagonal blocks;
//  * there is no triplet list and no sort: the sparse pattern (one block row per vertex) is built once at finalize and
//    every block gathers its element contributions in a fixed order, so the assembled matrix is reproducible bit for bit;
//  * the CG loop never leaves the device: reductions go through a fixed grid of partial sums that every block re-reduces
//    in the same order, the scalars (alpha, beta, convergence flag, iteration count) live in a two-entry state ring, and
//    iterations launched after convergence are no-ops.
#include <cfloat>

#include "energies.cuh"
#include "hinge_hessian.cuh"
#include "membrane_hessian.cuh"
#include "kernels.h"

namespace spgx
{
namespace
{
// ---------------------------------------------------------------------------------------------------------------------
// system force
// ---------------------------------------------------------------------------------------------------------------------
using D1 = Dual<3, false>;

template <int KIND>
__device__ __forceinline__ void forceKind(const NewtonEnergyDev &en, const double *__restrict__ x, int v, int sub, int G,
                                          V3d &g)
{
    using EK = EnergyKind<KIND>;
    constexpr int S = EK::S;
    const uint32_t rb = en.rowptr[v], re = en.rowptr[v + 1];
    for (uint32_t q = rb + sub; q < re; q += G) {
        const uint32_t ent = en.entries[q];
        const uint32_t st = ent >> 2;
        const int local = ent & 3u;
        const int32_t *ids = en.stencils + static_cast<size_t>(st) * S;
        const double *rec = en.records + static_cast<size_t>(st) * EK::NP;
        V3<D1> xs[S];
#pragma unroll
        for (int s = 0; s < S; ++s) {
            const size_t o = 3 * static_cast<size_t>(ids[s]);
            const bool own = (s == local);
            xs[s] = V3<D1>(D1(x[o], own ? 0 : -1), D1(x[o + 1], own ? 1 : -1), D1(x[o + 2], own ? 2 : -1));
        }
        const D1 E = stencilEnergy<KIND, D1>(xs, rec);
        g.x += E.g[0];
        g.y += E.g[1];
        g.z += E.g[2];
    }
}


constexpr int kForceLanes = 4;  // lanes per vertex

__global__ void __launch_bounds__(128) systemForceKernel(const NewtonArgs a, double gx, double gy, double gz,
                                                         double *__restrict__ f)
{
    constexpr int G = kForceLanes, VW = 32 / G;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const int sub = lane / VW;
    const int v = warp * VW + (lane % VW);
    if (warp * VW >= a.nVerts) return;
    const bool valid = v < a.nVerts;
    V3d g(0.0, 0.0, 0.0);
    if (valid) {
        for (int e = 0; e < a.nEnergies; ++e) {
            const NewtonEnergyDev &en = a.e[e];
            switch (en.kind) {
#define SPGX_KIND(NAME, ID, BYTES)                                                                                     \
    case NAME: forceKind<NAME>(en, a.x, v, sub, G, g); break;
#include "kinds.def"
#undef SPGX_KIND
            default: break;
            }
        }
    }
#pragma unroll
    for (int o = VW; o < 32; o <<= 1) {
        g.x += __shfl_xor_sync(0xffffffffu, g.x, o);
        g.y += __shfl_xor_sync(0xffffffffu, g.y, o);
        g.z += __shfl_xor_sync(0xffffffffu, g.z, o);
    }
    if (valid && sub == 0) {
        const double m = a.mass[v];
        const size_t o = 3 * static_cast<size_t>(v);
        f[o] = gx * m - g.x;
        f[o + 1] = gy * m - g.y;
        f[o + 2] = gz * m - g.z;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// element Hessians: one thread per (stencil, pair of stencil vertices)
// ---------------------------------------------------------------------------------------------------------------------
template <int S>
struct Pairs;
template <>
struct Pairs<1> {
    static constexpr int N = 1;
    __device__ static void get(int, int &a, int &b) { a = 0, b = -1; }
};
template <>
struct Pairs<2> {
    static constexpr int N = 1;
    __device__ static void get(int, int &a, int &b) { a = 0, b = 1; }
};
template <>
struct Pairs<3> {
    static constexpr int N = 3;
    __device__ static void get(int p, int &a, int &b)
    {
        a = p == 2 ? 1 : 0;
        b = p == 0 ? 1 : 2;
    }
};
template <>
struct Pairs<4> {
    static constexpr int N = 6;
    __device__ static void get(int p, int &a, int &b)
    {
        // (0,1) (0,2) (0,3) (1,2) (1,3) (2,3)
        a = p < 3 ? 0 : (p < 5 ? 1 : 2);
        b = p < 3 ? p + 1 : (p < 5 ? p - 1 : 3);
    }
};

// Work items of one stencil: its S (S - 1) / 2 vertex pairs (a < b), each evaluated with mixed second-order duals (16
// doubles per scalar) for the off-diagonal blocks (a, b) and (b, a) = transpose, then its S vertices, each evaluated with
// second-order duals seeded on that vertex alone (10 doubles per scalar) for the diagonal block.
template <int KIND>
__global__ void __launch_bounds__(128) elementHessianKernel(const NewtonEnergyDev en, const double *__restrict__ x)
{
    using EK = EnergyKind<KIND>;
    constexpr int S = EK::S;
    using P = Pairs<S>;
    constexpr int NP = S > 1 ? P::N : 0, NW = NP + S;
    const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    const long long st = t / NW;
    if (st >= en.nStencils) return;
    const int item = static_cast<int>(t % NW);
    const int32_t *ids = en.stencils + static_cast<size_t>(st) * S;
    const double *rec = en.records + static_cast<size_t>(st) * EK::NP;
    double *out = en.elemH + static_cast<size_t>(st) * (9 * S * S);
    if (item < NP) {
        int a, b;
        P::get(item, a, b);
        V3<DualMixed> xs[S];
#pragma unroll
        for (int s = 0; s < S; ++s) {
            const size_t o = 3 * static_cast<size_t>(ids[s]);
            const int base = s == a ? 0 : (s == b ? 3 : -4);  // negative seeds are passive
            xs[s] = V3<DualMixed>(DualMixed(x[o], base), DualMixed(x[o + 1], base + 1), DualMixed(x[o + 2], base + 2));
        }
        const DualMixed E = stencilEnergy<KIND, DualMixed>(xs, rec);
        double *bab = out + (a * S + b) * 9, *bba = out + (b * S + a) * 9;
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                bab[3 * i + j] = E.h[3 * i + j];
                bba[3 * j + i] = E.h[3 * i + j];
            }
    } else {
        const int a = item - NP;
        using D = Dual<3, true>;
        V3<D> xs[S];
#pragma unroll
        for (int s = 0; s < S; ++s) {
            const size_t o = 3 * static_cast<size_t>(ids[s]);
            const int base = s == a ? 0 : -4;
            xs[s] = V3<D>(D(x[o], base), D(x[o + 1], base + 1), D(x[o + 2], base + 2));
        }
        const D E = stencilEnergy<KIND, D>(xs, rec);
        double *baa = out + (a * S + a) * 9;
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) baa[3 * i + j] = E.hess(i, j);
    }
}

// discrete bending: the hand-derived 12 x 12 Hessian, one thread per hinge (hinge_hessian.cuh).  A thread's 16 blocks are
// 1152 contiguous bytes, so direct stores would scatter every warp-wide store over 32 sectors (measured: 452 M store
// sectors for 14 M requests, 2.3 GB of DRAM reads to fill partly written sectors, 5.3 ms).  One warp per block instead
// stages its 32 hinges in shared memory (row stride 145 doubles: conflict-free) and copies the block's contiguous
// 36 KB out with 16-byte coalesced stores.
constexpr int kHingeStride = 145;
__global__ void __launch_bounds__(32) hingeHessianKernel(const NewtonEnergyDev en, const double *__restrict__ x)
{
    extern __shared__ double stage[];
    const int lane = threadIdx.x;
    const long long base = static_cast<long long>(blockIdx.x) * 32;
    const long long st = base + lane;
    if (st < en.nStencils) {
        const int32_t *ids = en.stencils + static_cast<size_t>(st) * 4;
        V3d xs[4];
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            const size_t o = 3 * static_cast<size_t>(ids[s]);
            xs[s] = V3d(x[o], x[o + 1], x[o + 2]);
        }
        hingeElementHessian(xs, en.records + static_cast<size_t>(st) * EnergyKind<K_DISCRETE_BENDING>::NP, stage + lane * kHingeStride);
    }
    __syncwarp();
    const long long left = en.nStencils - base;
    const int total = static_cast<int>(left < 32 ? left : 32) * 144;  // doubles of this block's output (even)
    double *dst = en.elemH + static_cast<size_t>(base) * 144;
    for (int q = 2 * lane; q < total; q += 64) {
        const int e = q / 144, k = q - e * 144;  // q and q + 1 belong to the same hinge (144 is even)
        const double *src = stage + e * kHingeStride + k;
        *reinterpret_cast<double2 *>(dst + q) = make_double2(src[0], src[1]);
    }
}

// StVK membrane: the hand-derived 9 x 9 Hessian (membrane_hessian.cuh), one thread per triangle, staged like the hinge
// kernel (a triangle's 9 blocks are 648 bytes: 8-byte coalesced stores, the stage stride 81 is odd: conflict-free).
// Opt-in (SPGX_MEMBRANE_HAND=1) until it has been measured and parity-checked on the device; the arithmetic is checked on
// the host against Dual<9, true> (tests/native/host_membrane_hessian.cpp).
__global__ void __launch_bounds__(32) membraneStvkHessianKernel(const NewtonEnergyDev en, const double *__restrict__ x)
{
    extern __shared__ double stage[];
    const int lane = threadIdx.x;
    const long long base = static_cast<long long>(blockIdx.x) * 32;
    const long long st = base + lane;
    if (st < en.nStencils) {
        const int32_t *ids = en.stencils + static_cast<size_t>(st) * 3;
        V3d xs[3];
#pragma unroll
        for (int s = 0; s < 3; ++s) {
            const size_t o = 3 * static_cast<size_t>(ids[s]);
            xs[s] = V3d(x[o], x[o + 1], x[o + 2]);
        }
        membraneStvkElementHessian(xs, en.records + static_cast<size_t>(st) * EnergyKind<K_MEMBRANE_STVK>::NP, stage + lane * 81);
    }
    __syncwarp();
    const long long left = en.nStencils - base;
    const int total = static_cast<int>(left < 32 ? left : 32) * 81;
    double *dst = en.elemH + static_cast<size_t>(base) * 81;
    for (int q = lane; q < total; q += 32) dst[q] = stage[q];  // (stage rows are contiguous at stride 81: a flat copy)
}

template <int KIND>
void launchElementHessianKind(const NewtonEnergyDev &en, const double *x, cudaStream_t stream)
{
    if (KIND == K_MEMBRANE_STVK) {
        const char *env = getenv("SPGX_MEMBRANE_HAND");  // read per launch: a test can switch it inside one process
        const bool hand = env && atoi(env) != 0;
        if (hand) {
            if (en.nStencils > 0)
                membraneStvkHessianKernel<<<static_cast<unsigned>((en.nStencils + 31) / 32), 32, 32 * 81 * sizeof(double), stream>>>(en, x);
            return;
        }
    }
    if (KIND == K_DISCRETE_BENDING) {
        // SPGX_HINGE_HAND=0: the generic mixed-dual evaluation (what the hand-derived blocks are tested against)
        static const bool hand = !(getenv("SPGX_HINGE_HAND") && atoi(getenv("SPGX_HINGE_HAND")) == 0);
        if (hand) {
            if (en.nStencils > 0)
                hingeHessianKernel<<<static_cast<unsigned>((en.nStencils + 31) / 32), 32, 32 * kHingeStride * sizeof(double), stream>>>(en, x);
            return;
        }
    }
    constexpr int S = EnergyKind<KIND>::S;
    const long long threads = static_cast<long long>(en.nStencils) * ((S > 1 ? Pairs<S>::N : 0) + S);
    if (threads == 0) return;
    elementHessianKernel<KIND><<<static_cast<unsigned>((threads + 127) / 128), 128, 0, stream>>>(en, x);
}

// ---------------------------------------------------------------------------------------------------------------------
// block-sparse assembly: one thread per block (row, col)
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) bsrAssembleKernel(const NewtonArgs a)
{
    const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (t >= a.nnzb) return;
    const int row = a.bsrRow[t], col = a.bsrCol[t];
    double acc[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int e = 0; e < a.nEnergies; ++e) {
        const NewtonEnergyDev &en = a.e[e];
        const int S = kindStencilSize(en.kind);
        const uint32_t rb = en.rowptr[row], re = en.rowptr[row + 1];
        for (uint32_t q = rb; q < re; ++q) {
            const uint32_t ent = en.entries[q];
            const uint32_t st = ent >> 2;
            const int la = ent & 3u;
            const int32_t *ids = en.stencils + static_cast<size_t>(st) * S;
            for (int lb = 0; lb < S; ++lb) {
                if (ids[lb] != col) continue;
                const double *blk = en.elemH + static_cast<size_t>(st) * (9 * S * S) + (la * S + lb) * 9;
#pragma unroll
                for (int k = 0; k < 9; ++k) acc[k] += blk[k];
            }
        }
    }
    // component-major storage (component k of block t at bsrH[k * nnzb + t]): the write here and the reads of the sparse
    // product are unit-stride across the lanes that own consecutive blocks
#pragma unroll
    for (int k = 0; k < 9; ++k) a.bsrH[k * a.nnzb + t] = acc[k];
}

// the same sums from the precomputed gather lists (api_newton.cu newtonBuild): no incidence walk, no stencil-id
// compares, four element blocks requested per trip; the order of the additions is the one above, bit for bit
__global__ void __launch_bounds__(128) bsrAssembleGatherKernel(const NewtonArgs a)
{
    const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (t >= a.nnzb) return;
    const uint32_t b = a.asmPtr[t], e = a.asmPtr[t + 1];
    double acc[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    uint32_t k = b;
    for (; k + 4 <= e; k += 4) {
        const double *p0 = a.elemArena + 9 * static_cast<size_t>(a.asmSrc[k]);
        const double *p1 = a.elemArena + 9 * static_cast<size_t>(a.asmSrc[k + 1]);
        const double *p2 = a.elemArena + 9 * static_cast<size_t>(a.asmSrc[k + 2]);
        const double *p3 = a.elemArena + 9 * static_cast<size_t>(a.asmSrc[k + 3]);
        double v0[9], v1[9], v2[9], v3[9];
#pragma unroll
        for (int c = 0; c < 9; ++c) v0[c] = p0[c], v1[c] = p1[c], v2[c] = p2[c], v3[c] = p3[c];  // (plain loads: the barrier below orders them)
        asm volatile("" ::: "memory");  // all 36 loads are requested before the first addition waits for one
#pragma unroll
        for (int c = 0; c < 9; ++c) acc[c] = ((acc[c] + v0[c]) + v1[c]) + v2[c] + v3[c];
    }
    for (; k < e; ++k) {
        const double *p0 = a.elemArena + 9 * static_cast<size_t>(a.asmSrc[k]);
#pragma unroll
        for (int c = 0; c < 9; ++c) acc[c] += p0[c];
    }
#pragma unroll
    for (int c = 0; c < 9; ++c) a.bsrH[c * a.nnzb + t] = acc[c];
}

// ---------------------------------------------------------------------------------------------------------------------
// conjugate gradient
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kRowLanes = 8;  // lanes per block row in the sparse product

// sum over the block, result valid in thread 0 (fixed tree: deterministic)
__device__ __forceinline__ double blockSum(double v, double *smem)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) smem[warp] = v;
    __syncthreads();
    double s = 0.0;
    if (threadIdx.x == 0)
        for (int w = 0; w < kCgThreads / 32; ++w) s += smem[w];
    return s;
}
// every block reduces the kCgBlocks partial sums itself, in the same order: result in every thread
__device__ __forceinline__ double reducePartials(const double *__restrict__ partials, double *smem)
{
    double v = 0.0;
    for (int i = threadIdx.x; i < kCgBlocks; i += kCgThreads) v += partials[i];
    const double s = blockSum(v, smem);
    __shared__ double bcast;
    if (threadIdx.x == 0) bcast = s;
    __syncthreads();
    const double r = bcast;
    __syncthreads();
    return r;
}

// y = (cM diag(mass) + cH H) x over block rows; partial sums of x . y when `partials` is given.
// VARIANT 0: one block per lane and trip.  1: the row pointers of the next sweep are requested before this sweep's blocks
// (takes one dependent round trip out of every sweep).  2: as 1, and the first two trips of a row (up to 16 blocks: every
// row of a triangle mesh) are requested together.  The per-lane summation order is the same in all three: results are
// bit-identical.
struct SpmvBlock {
    double h0, h1, h2, h3, h4, h5, h6, h7, h8;
    size_t c;
};
__device__ __forceinline__ SpmvBlock spmvLoad(const CgArgs &a, uint32_t k)
{
    const size_t nb = static_cast<size_t>(a.nnzb);
    const double *h = a.bsrH + k;
    SpmvBlock b;
    b.c = 3 * static_cast<size_t>(a.bsrCol[k]);
    b.h0 = __ldg(h), b.h1 = __ldg(h + nb), b.h2 = __ldg(h + 2 * nb), b.h3 = __ldg(h + 3 * nb), b.h4 = __ldg(h + 4 * nb);
    b.h5 = __ldg(h + 5 * nb), b.h6 = __ldg(h + 6 * nb), b.h7 = __ldg(h + 7 * nb), b.h8 = __ldg(h + 8 * nb);
    return b;
}
__device__ __forceinline__ SpmvBlock spmvLoadH(const CgArgs &a, uint32_t k, int32_t col)
{
    const size_t nb = static_cast<size_t>(a.nnzb);
    const double *h = a.bsrH + k;
    SpmvBlock b;
    b.c = 3 * static_cast<size_t>(col);
    b.h0 = __ldg(h), b.h1 = __ldg(h + nb), b.h2 = __ldg(h + 2 * nb), b.h3 = __ldg(h + 3 * nb), b.h4 = __ldg(h + 4 * nb);
    b.h5 = __ldg(h + 5 * nb), b.h6 = __ldg(h + 6 * nb), b.h7 = __ldg(h + 7 * nb), b.h8 = __ldg(h + 8 * nb);
    return b;
}
__device__ __forceinline__ void spmvAccumulate(const SpmvBlock &b, double x0, double x1, double x2, double &y0, double &y1, double &y2)
{
    y0 += b.h0 * x0 + b.h1 * x1 + b.h2 * x2;
    y1 += b.h3 * x0 + b.h4 * x1 + b.h5 * x2;
    y2 += b.h6 * x0 + b.h7 * x1 + b.h8 * x2;
}

template <int VARIANT>
__global__ void __launch_bounds__(kCgThreads) bsrSpmvKernel(const CgArgs a, const double *__restrict__ x,
                                                           double *__restrict__ y, double *__restrict__ partials,
                                                           const CgState *__restrict__ state)
{
    __shared__ double smem[kCgThreads / 32];
    if (state && state->done) return;
    constexpr int RPB = kCgThreads / kRowLanes;  // rows per block per sweep
    const int sub = threadIdx.x % kRowLanes;
    const int rowInBlock = threadIdx.x / kRowLanes;
    const long long stride = static_cast<long long>(gridDim.x) * RPB;
    double dotAcc = 0.0;
    uint32_t rbNext = 0, reNext = 0;
    if (VARIANT >= 1) {
        const long long row = static_cast<long long>(blockIdx.x) * RPB + rowInBlock;
        if (row < a.nVerts) rbNext = a.bsrRowptr[row], reNext = a.bsrRowptr[row + 1];
    }
    for (long long base = static_cast<long long>(blockIdx.x) * RPB; base < a.nVerts; base += stride) {
        const long long row = base + rowInBlock;
        double y0 = 0.0, y1 = 0.0, y2 = 0.0;
        uint32_t rb = rbNext, re = reNext;
        if (VARIANT >= 1) {
            rbNext = reNext = 0;
            if (row + stride < a.nVerts) rbNext = a.bsrRowptr[row + stride], reNext = a.bsrRowptr[row + stride + 1];
        } else if (row < a.nVerts) {
            rb = a.bsrRowptr[row], re = a.bsrRowptr[row + 1];
        }
        if (row < a.nVerts) {
            // (two blocks per lane and trip together with a 148 x 8 grid measured slower: 90.8 vs 81.1 ms per 300
            // iterations at 1 M vertices)
            uint32_t k = rb + sub;
            if (VARIANT >= 2) {
                const bool p1 = k < re, p2 = k + kRowLanes < re;
                if (p1) {
                    const SpmvBlock b1 = spmvLoad(a, k);
                    const SpmvBlock b2 = spmvLoad(a, p2 ? k + kRowLanes : k);
                    const double u0 = x[b1.c], u1 = x[b1.c + 1], u2 = x[b1.c + 2];
                    const double w0 = x[b2.c], w1 = x[b2.c + 1], w2 = x[b2.c + 2];
                    spmvAccumulate(b1, u0, u1, u2, y0, y1, y2);
                    if (p2) spmvAccumulate(b2, w0, w1, w2, y0, y1, y2);
                }
                k += 2 * kRowLanes;
            }
            for (; k < re; k += kRowLanes) {
                const SpmvBlock b = spmvLoad(a, k);
                spmvAccumulate(b, x[b.c], x[b.c + 1], x[b.c + 2], y0, y1, y2);
            }
        }
#pragma unroll
        for (int o = 1; o < kRowLanes; o <<= 1) {
            y0 += __shfl_xor_sync(0xffffffffu, y0, o);
            y1 += __shfl_xor_sync(0xffffffffu, y1, o);
            y2 += __shfl_xor_sync(0xffffffffu, y2, o);
        }
        if (row < a.nVerts && sub == 0) {
            const size_t o = 3 * static_cast<size_t>(row);
            const double m = a.cM * a.mass[row];
            const double x0 = x[o], x1 = x[o + 1], x2 = x[o + 2];
            y0 = a.cH * y0 + m * x0;
            y1 = a.cH * y1 + m * x1;
            y2 = a.cH * y2 + m * x2;
            y[o] = y0, y[o + 1] = y1, y[o + 2] = y2;
            dotAcc += x0 * y0 + x1 * y1 + x2 * y2;
        }
    }
    if (partials) {
        const double s = blockSum(dotAcc, smem);
        if (threadIdx.x == 0) partials[blockIdx.x] = s;
    }
}

// VARIANT 3: software pipeline over the sweeps of a block: the row pointers are requested two sweeps ahead and the column
// ids one sweep ahead, so a sweep's matrix blocks and its gathered x entries are requested together (one exposed memory
// round trip per sweep instead of three).  Same lanes, same per-lane order: bit-identical to the other variants.
__global__ void __launch_bounds__(kCgThreads, 4) bsrSpmvPipelinedKernel(const CgArgs a, const double *__restrict__ x,
                                                                    double *__restrict__ y, double *__restrict__ partials,
                                                                    const CgState *__restrict__ state)
{
    __shared__ double smem[kCgThreads / 32];
    if (state && state->done) return;
    constexpr int RPB = kCgThreads / kRowLanes;
    const int sub = threadIdx.x % kRowLanes;
    const int rowInBlock = threadIdx.x / kRowLanes;
    const long long stride = static_cast<long long>(gridDim.x) * RPB;
    const long long n = a.nVerts;
    double dotAcc = 0.0;
    long long row = static_cast<long long>(blockIdx.x) * RPB + rowInBlock;
    uint32_t rb = 0, re = 0, rbN = 0, reN = 0;
    if (row < n) rb = a.bsrRowptr[row], re = a.bsrRowptr[row + 1];
    if (row + stride < n) rbN = a.bsrRowptr[row + stride], reN = a.bsrRowptr[row + stride + 1];
    int32_t c1 = 0, c2 = 0;
    if (rb + sub < re) c1 = a.bsrCol[rb + sub];
    if (rb + sub + kRowLanes < re) c2 = a.bsrCol[rb + sub + kRowLanes];
    for (long long base = static_cast<long long>(blockIdx.x) * RPB; base < n; base += stride, row += stride) {
        // requests for the sweeps to come
        uint32_t rbNN = 0, reNN = 0;
        if (row + 2 * stride < n) rbNN = a.bsrRowptr[row + 2 * stride], reNN = a.bsrRowptr[row + 2 * stride + 1];
        int32_t c1N = 0, c2N = 0;
        if (rbN + sub < reN) c1N = a.bsrCol[rbN + sub];
        if (rbN + sub + kRowLanes < reN) c2N = a.bsrCol[rbN + sub + kRowLanes];
        // this sweep (the row's own x entries and mass are requested up front too: after the lane reduction they would be
        // one more exposed round trip)
        double y0 = 0.0, y1 = 0.0, y2 = 0.0;
        double xr0 = 0.0, xr1 = 0.0, xr2 = 0.0, mr = 0.0;
        if (row < n && sub == 0) {
            const size_t o = 3 * static_cast<size_t>(row);
            xr0 = x[o], xr1 = x[o + 1], xr2 = x[o + 2], mr = a.mass[row];
        }
        uint32_t k = rb + sub;
        if (k < re) {
            const bool p2 = k + kRowLanes < re;
            const SpmvBlock b1 = spmvLoadH(a, k, c1);
            const SpmvBlock b2 = spmvLoadH(a, p2 ? k + kRowLanes : k, p2 ? c2 : c1);
            const double u0 = x[b1.c], u1 = x[b1.c + 1], u2 = x[b1.c + 2];
            const double w0 = x[b2.c], w1 = x[b2.c + 1], w2 = x[b2.c + 2];
            spmvAccumulate(b1, u0, u1, u2, y0, y1, y2);
            if (p2) spmvAccumulate(b2, w0, w1, w2, y0, y1, y2);
            for (k += 2 * kRowLanes; k < re; k += kRowLanes) {
                const SpmvBlock b = spmvLoad(a, k);
                spmvAccumulate(b, x[b.c], x[b.c + 1], x[b.c + 2], y0, y1, y2);
            }
        }
#pragma unroll
        for (int o = 1; o < kRowLanes; o <<= 1) {
            y0 += __shfl_xor_sync(0xffffffffu, y0, o);
            y1 += __shfl_xor_sync(0xffffffffu, y1, o);
            y2 += __shfl_xor_sync(0xffffffffu, y2, o);
        }
        if (row < n && sub == 0) {
            const size_t o = 3 * static_cast<size_t>(row);
            const double m = a.cM * mr;
            const double x0 = xr0, x1 = xr1, x2 = xr2;
            y0 = a.cH * y0 + m * x0;
            y1 = a.cH * y1 + m * x1;
            y2 = a.cH * y2 + m * x2;
            y[o] = y0, y[o + 1] = y1, y[o + 2] = y2;
            dotAcc += x0 * y0 + x1 * y1 + x2 * y2;
        }
        rb = rbN, re = reN, rbN = rbNN, reN = reNN, c1 = c1N, c2 = c2N;
    }
    if (partials) {
        const double s = blockSum(dotAcc, smem);
        if (threadIdx.x == 0) partials[blockIdx.x] = s;
    }
}

void launchSpmvVariant(const CgArgs &a, const double *x, double *y, double *partials, const CgState *state, cudaStream_t stream)
{
    static const int variant = getenv("SPGX_SPMV_VARIANT") ? atoi(getenv("SPGX_SPMV_VARIANT")) : 3;
    if (variant == 3)
        bsrSpmvPipelinedKernel<<<kCgBlocks, kCgThreads, 0, stream>>>(a, x, y, partials, state);
    else if (variant == 1)
        bsrSpmvKernel<1><<<kCgBlocks, kCgThreads, 0, stream>>>(a, x, y, partials, state);
    else if (variant == 2)
        bsrSpmvKernel<2><<<kCgBlocks, kCgThreads, 0, stream>>>(a, x, y, partials, state);
    else
        bsrSpmvKernel<0><<<kCgBlocks, kCgThreads, 0, stream>>>(a, x, y, partials, state);
}

// Jacobi preconditioner (BasicPreconditioners.h:62-73): 1 / diagonal, 1 where the diagonal is zero
__global__ void invDiagKernel(const CgArgs a)
{
    const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= 3LL * a.nVerts) return;
    const long long v = i / 3;
    const int c = static_cast<int>(i % 3);
    const double d = a.cM * a.mass[v] + a.cH * a.bsrH[static_cast<size_t>(4 * c) * a.nnzb + a.diagPos[v]];
    a.invdiag[i] = d != 0.0 ? 1.0 / d : 1.0;
}

// x = 0, r = b, p = invdiag r; partial sums of b . b and r . p
__global__ void __launch_bounds__(kCgThreads) cgInitKernel(const CgArgs a, const double *__restrict__ b)
{
    __shared__ double smem[kCgThreads / 32];
    const long long n = 3LL * a.nVerts;
    double bb = 0.0, rp = 0.0;
    for (long long i = static_cast<long long>(blockIdx.x) * kCgThreads + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * kCgThreads) {
        const double bi = b[i];
        const double pi = a.invdiag[i] * bi;
        a.x[i] = 0.0;
        a.r[i] = bi;
        a.p[i] = pi;
        bb += bi * bi;
        rp += bi * pi;
    }
    const double s0 = blockSum(bb, smem);
    const double s1 = blockSum(rp, smem);
    if (threadIdx.x == 0) {
        a.partials[blockIdx.x] = s0;
        a.partials[kCgBlocks + blockIdx.x] = s1;
    }
}
__global__ void __launch_bounds__(kCgThreads) cgInitFinishKernel(const CgArgs a, double tol)
{
    __shared__ double smem[kCgThreads / 32];
    const double rhsNorm2 = reducePartials(a.partials, smem);
    const double absNew = reducePartials(a.partials + kCgBlocks, smem);
    if (threadIdx.x == 0) {
        CgState s;
        s.absNew = absNew;
        s.rhsNorm2 = rhsNorm2;
        s.threshold = fmax(tol * tol * rhsNorm2, DBL_MIN);
        s.residualNorm2 = rhsNorm2;
        // rhs == 0: x = 0 and no iteration (ConjugateGradient.h:44-49); r = b, so the initial-residual test of :52 can
        // only pass when tol >= 1
        s.done = (rhsNorm2 == 0.0 || rhsNorm2 < s.threshold) ? 1 : 0;
        s.iters = 0;
        a.state[0] = s;
    }
}
// alpha = absNew / p . tmp; x += alpha p; r -= alpha tmp; z = invdiag r; partial sums of r . r and r . z
__global__ void __launch_bounds__(kCgThreads) cgUpdateKernel(const CgArgs a, int it)
{
    __shared__ double smem[kCgThreads / 32];
    const long long n = 3LL * a.nVerts;
    double rr = 0.0, rz = 0.0;
    // four independent elements per trip: 20 loads in flight per thread (the loop is pure latency otherwise); the first
    // trip is requested BEFORE the reduction of the p . tmp partials, which does not depend on it
    constexpr int U = 4;
    const long long stride = static_cast<long long>(gridDim.x) * kCgThreads;
    long long i0 = static_cast<long long>(blockIdx.x) * kCgThreads + threadIdx.x;
    double xv[U], pv[U], rv[U], tv[U], dv[U];
    auto request = [&](long long at) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long i = at + u * stride;
            if (i < n) xv[u] = a.x[i], pv[u] = a.p[i], rv[u] = a.r[i], tv[u] = a.tmp[i], dv[u] = a.invdiag[i];
        }
    };
    request(i0);
    const CgState st = a.state[it & 1];
    if (st.done) return;
    const double pAp = reducePartials(a.partials, smem);
    const double alpha = st.absNew / pAp;
    while (i0 < n) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long i = i0 + u * stride;
            if (i < n) {
                a.x[i] = xv[u] + alpha * pv[u];
                const double ri = rv[u] - alpha * tv[u];
                const double zi = dv[u] * ri;
                a.r[i] = ri;
                a.z[i] = zi;
                rr += ri * ri;
                rz += ri * zi;
            }
        }
        i0 += U * stride;
        request(i0);
    }
    const double s0 = blockSum(rr, smem);
    const double s1 = blockSum(rz, smem);
    if (threadIdx.x == 0) {
        a.partials[kCgBlocks + blockIdx.x] = s1;
        // r . r goes to a third range so that the p . tmp partials stay intact for blocks still reducing them
        a.partials[2 * kCgBlocks + blockIdx.x] = s0;
    }
}
// convergence test, beta = absNew' / absNew, p = z + beta p; block 0 writes the next state
__global__ void __launch_bounds__(kCgThreads) cgDirectionKernel(const CgArgs a, int it)
{
    __shared__ double smem[kCgThreads / 32];
    const long long n = 3LL * a.nVerts;
    constexpr int U = 4;
    const long long stride = static_cast<long long>(gridDim.x) * kCgThreads;
    long long i0 = static_cast<long long>(blockIdx.x) * kCgThreads + threadIdx.x;
    double zv[U], pv[U];
    auto request = [&](long long at) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long i = at + u * stride;
            if (i < n) zv[u] = a.z[i], pv[u] = a.p[i];
        }
    };
    request(i0);  // before the state and the two reductions, which do not depend on it
    const CgState st = a.state[it & 1];
    if (st.done) {
        if (blockIdx.x == 0 && threadIdx.x == 0) a.state[(it + 1) & 1] = st;
        return;
    }
    const double rr = reducePartials(a.partials + 2 * kCgBlocks, smem);
    const double rz = reducePartials(a.partials + kCgBlocks, smem);
    const bool converged = rr < st.threshold;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        CgState nx = st;
        nx.residualNorm2 = rr;
        nx.done = converged ? 1 : 0;
        if (!converged) {
            nx.absNew = rz;
            nx.iters = st.iters + 1;
        }
        a.state[(it + 1) & 1] = nx;
    }
    if (converged) return;
    const double beta = rz / st.absNew;
    while (i0 < n) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long i = i0 + u * stride;
            if (i < n) a.p[i] = zv[u] + beta * pv[u];
        }
        i0 += U * stride;
        request(i0);
    }
}

__global__ void axpbyKernel(double *out, double alpha, const double *a, double beta, const double *b, long long n)
{
    const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) out[i] = alpha * a[i] + beta * b[i];
}
__global__ void __launch_bounds__(kCgThreads) residualKernel(double *__restrict__ out, double c1, const double *__restrict__ f, double c2,
                                                            const double *__restrict__ mass, const double *__restrict__ a,
                                                            const double *__restrict__ b, long long n3, double *__restrict__ partials)
{
    __shared__ double smem[kCgThreads / 32];
    double acc = 0.0;
    for (long long i = static_cast<long long>(blockIdx.x) * kCgThreads + threadIdx.x; i < n3; i += static_cast<long long>(gridDim.x) * kCgThreads) {
        const double r = c1 * f[i] - c2 * (mass[i / 3] * (a[i] - b[i]));
        out[i] = r;
        acc += r * r;
    }
    const double s = blockSum(acc, smem);
    if (threadIdx.x == 0) partials[blockIdx.x] = s;
}
__global__ void __launch_bounds__(kCgThreads) normFinishKernel(const double *__restrict__ partials, double *__restrict__ normOut)
{
    __shared__ double smem[kCgThreads / 32];
    const double s = reducePartials(partials, smem);
    if (threadIdx.x == 0) *normOut = ::sqrt(s);
}
__global__ void addMassDiffKernel(double *__restrict__ out, double c, const double *__restrict__ mass, const double *__restrict__ a,
                                  const double *__restrict__ b, long long n3)
{
    const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n3) out[i] += c * (mass[i / 3] * (a[i] - b[i]));
}
__global__ void applyDxKernel(double *__restrict__ x, double *__restrict__ v, const double *__restrict__ dx,
                              const double *__restrict__ x0, double invdt, long long n)
{
    const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double xn = x[i] + dx[i];
    x[i] = xn;
    v[i] = (xn - x0[i]) * invdt;
}
__global__ void symplecticKernel(double *__restrict__ x, double *__restrict__ v, const double *__restrict__ f,
                                 const double *__restrict__ invMass, long long n3, double dt, double gx, double gy, double gz)
{
    const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n3) return;
    const int c = static_cast<int>(i % 3);
    double vn = v[i] + dt * (c == 0 ? gx : (c == 1 ? gy : gz));
    vn += invMass[i / 3] * f[i] * dt;
    v[i] = vn;
    x[i] += vn * dt;
}
}  // namespace

void launchSystemForce(const NewtonArgs &a, const double *g3, double *f, cudaStream_t stream)
{
    if (a.nVerts == 0) return;
    constexpr int VW = 32 / kForceLanes;
    const int warps = (a.nVerts + VW - 1) / VW;
    const int blocks = (warps * 32 + 127) / 128;
    systemForceKernel<<<blocks, 128, 0, stream>>>(a, g3[0], g3[1], g3[2], f);
}

void launchElementHessians(const NewtonArgs &a, cudaStream_t stream)
{
    for (int e = 0; e < a.nEnergies; ++e) {
        switch (a.e[e].kind) {
#define SPGX_KIND(NAME, ID, BYTES)                                                                                     \
    case NAME: launchElementHessianKind<NAME>(a.e[e], a.x, stream); break;
#include "kinds.def"
#undef SPGX_KIND
        default: break;
        }
    }
}

void launchBsrAssemble(const NewtonArgs &a, cudaStream_t stream)
{
    if (a.nnzb == 0) return;
    // SPGX_ASM_GATHER=0: the incidence-walking kernel (3.8 ms at config 6)
    static const bool gather = !(getenv("SPGX_ASM_GATHER") && atoi(getenv("SPGX_ASM_GATHER")) == 0);
    if (gather && a.asmPtr)
        bsrAssembleGatherKernel<<<static_cast<unsigned>((a.nnzb + 127) / 128), 128, 0, stream>>>(a);
    else
        bsrAssembleKernel<<<static_cast<unsigned>((a.nnzb + 127) / 128), 128, 0, stream>>>(a);
}

void launchBsrSpmv(const CgArgs &a, const double *x, double *y, cudaStream_t stream)
{
    if (a.nVerts == 0) return;
    launchSpmvVariant(a, x, y, nullptr, nullptr, stream);
}

void launchCgInit(const CgArgs &a, const double *b, double tol, cudaStream_t stream)
{
    const long long n = 3LL * a.nVerts;
    invDiagKernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(a);
    cgInitKernel<<<kCgBlocks, kCgThreads, 0, stream>>>(a, b);
    cgInitFinishKernel<<<1, kCgThreads, 0, stream>>>(a, tol);
}

void launchCgIteration(const CgArgs &a, int it, cudaStream_t stream)
{
    launchSpmvVariant(a, a.p, a.tmp, a.partials, a.state + (it & 1), stream);
    cgUpdateKernel<<<kCgBlocks, kCgThreads, 0, stream>>>(a, it);
    cgDirectionKernel<<<kCgBlocks, kCgThreads, 0, stream>>>(a, it);
}

void launchAxpby(double *out, double alpha, const double *a, double beta, const double *b, long long n, cudaStream_t stream)
{
    if (n == 0) return;
    axpbyKernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(out, alpha, a, beta, b, n);
}

void launchResidualNorm(double *out, double c1, const double *f, double c2, const double *mass, const double *a, const double *b,
                        int nVerts, double *partials, double *normOut, cudaStream_t stream)
{
    residualKernel<<<kCgBlocks, kCgThreads, 0, stream>>>(out, c1, f, c2, mass, a, b, 3LL * nVerts, partials);
    normFinishKernel<<<1, kCgThreads, 0, stream>>>(partials, normOut);
}

void launchAddMassDiff(double *out, double c, const double *mass, const double *a, const double *b, int nVerts, cudaStream_t stream)
{
    const long long n3 = 3LL * nVerts;
    if (n3 == 0) return;
    addMassDiffKernel<<<static_cast<unsigned>((n3 + 255) / 256), 256, 0, stream>>>(out, c, mass, a, b, n3);
}

void launchApplyDx(double *x, double *v, const double *dx, const double *x0, double invdt, long long n, cudaStream_t stream)
{
    if (n == 0) return;
    applyDxKernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(x, v, dx, x0, invdt, n);
}

void launchSymplecticUpdate(double *x, double *v, const double *fInternal, const double *invMass, int nVerts, double dt,
                            const double *g3, cudaStream_t stream)
{
    const long long n3 = 3LL * nVerts;
    if (n3 == 0) return;
    symplecticKernel<<<static_cast<unsigned>((n3 + 255) / 256), 256, 0, stream>>>(x, v, fInternal, invMass, n3, dt, g3[0],
                                                                                  g3[1], g3[2]);
}
}  // namespace spgx
