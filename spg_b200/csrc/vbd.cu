// Vertex Block Descent colour pass (reference: spg/source/sim/solver/vbd.cpp:99-157).
//
// One thread group of G lanes owns one vertex of the colour.  Lanes stride over the vertex's incident
// (stencil, local) entries energy by energy (all lanes of a warp are inside the same energy kind at the
// same time, so kinds never diverge), each lane evaluates the element energy with second-order duals
// seeded on the OWNING vertex only (K = 3: value, 3-gradient, packed 3x3 Hessian) -- the diagonal block the
// reference extracts from a full 3S x 3S autodiff pass (energy.h:240-258) -- and the lanes' partial
// force / Hessian are reduced with warp shuffles.  Lane 0 adds the inertia term, applies the
// det(H) > 1e-10 gate and the cofactor solve, and writes the vertex.
#include "energies.cuh"
#include "kernels.h"

namespace spgx
{
namespace
{
using D2 = Dual<3, true>;

template <int KIND>
__device__ __forceinline__ void accumulateKind(const VbdEnergyDev &en, const double *__restrict__ x, int slot, int lane,
                                               int G, V3d &g, Sym3 &H)
{
    using EK = EnergyKind<KIND>;
    constexpr int S = EK::S;
    const uint32_t rb = en.rowptr[slot], re = en.rowptr[slot + 1];
    for (uint32_t q = rb + lane; q < re; q += G) {
        const uint32_t ent = en.entries[q];
        const uint32_t st = ent >> 2;
        const int local = ent & 3u;
        const int32_t *ids = en.stencils + static_cast<size_t>(st) * S;
        const double *rec = en.records + static_cast<size_t>(st) * EK::NP;
        if constexpr (HasHandDerived<EK>::value) {
            V3d xs[S];
#pragma unroll
            for (int s = 0; s < S; ++s) {
                const size_t o = 3 * static_cast<size_t>(ids[s]);
                xs[s] = V3d(x[o], x[o + 1], x[o + 2]);
            }
            EK::vertexGradHess(xs, rec, local, g, H);
        } else {
            V3<D2> xs[S];
#pragma unroll
            for (int s = 0; s < S; ++s) {
                const size_t o = 3 * static_cast<size_t>(ids[s]);
                const bool own = (s == local);
                xs[s] = V3<D2>(D2(x[o], own ? 0 : -1), D2(x[o + 1], own ? 1 : -1), D2(x[o + 2], own ? 2 : -1));
            }
            const D2 E = stencilEnergy<KIND, D2>(xs, rec);
            g.x += E.g[0];
            g.y += E.g[1];
            g.z += E.g[2];
            H.xx += E.h[0];
            H.xy += E.h[1];
            H.xz += E.h[2];
            H.yy += E.h[3];
            H.yz += E.h[4];
            H.zz += E.h[5];
        }
    }
}

// streaming path: the slice's step records are read fully coalesced (one 128 B / 256 B line per field per warp).
// The loop is software-pipelined by hand: while step t is being evaluated, the neighbour positions of step t + 1
// and the record of step t + 2 are already in flight, so the DRAM (record) and L2 (gather) latencies overlap the
// FP64 work instead of serialising with it -- the kernel is latency-bound otherwise (profiles/r1_notes.md).
template <int KIND>
struct EllStep {
    using EK = EnergyKind<KIND>;
    uint32_t id[EK::NI];
    double q[EK::NQ];
    __device__ __forceinline__ void load(const VbdEnergyDev &en, uint32_t t, int lane)
    {
        const uint32_t *idp = en.ellIds + static_cast<size_t>(t) * (EK::NI * 32) + lane;
        const double *qp = en.ellParams + static_cast<size_t>(t) * (EK::NQ * 32) + lane;
#pragma unroll
        for (int k = 0; k < EK::NI; ++k) id[k] = __ldg(idp + k * 32);
#pragma unroll
        for (int k = 0; k < EK::NQ; ++k) q[k] = __ldg(qp + k * 32);
    }
    __device__ __forceinline__ bool padding() const { return id[0] == 0xFFFFFFFFu; }
    __device__ __forceinline__ void gather(const double *__restrict__ x, V3d *xn) const
    {
#pragma unroll
        for (int k = 0; k < EK::NI; ++k) {
            // padding lanes read slot 0 (harmless) so the gather needs no branch
            const uint32_t v = padding() ? 0u : (k == 0 ? (id[0] & 0x3FFFFFFFu) : id[k]);
            const size_t o = 3 * static_cast<size_t>(v);
            xn[k] = V3d(x[o], x[o + 1], x[o + 2]);
        }
    }
};

template <int KIND>
__device__ __forceinline__ void accumulateEll(const VbdEnergyDev &en, const double *__restrict__ x, int slice, int lane,
                                              const V3d &xo, V3d &g, Sym3 &H)
{
    using EK = EnergyKind<KIND>;
    const uint32_t tb = en.ellOffsets[slice], te = en.ellOffsets[slice + 1];
    if (tb >= te) return;
    EllStep<KIND> r0, r1, r2;
    V3d p0[EK::NI], p1[EK::NI];
    r0.load(en, tb, lane);
    r1 = r0;
    if (tb + 1 < te) r1.load(en, tb + 1, lane);
    r0.gather(x, p0);
    for (uint32_t t = tb; t < te; ++t) {
        r2 = r1;
        if (t + 2 < te) r2.load(en, t + 2, lane);
        if (t + 1 < te) r1.gather(x, p1);
        if (!r0.padding()) EK::incidenceGradHess(xo, p0, r0.q, static_cast<int>(r0.id[0] >> 30), g, H);
        r0 = r1;
        r1 = r2;
#pragma unroll
        for (int k = 0; k < EK::NI; ++k) p0[k] = p1[k];
    }
}

template <unsigned MASK, int KIND>
__device__ __forceinline__ void accumulateIfPresent(const VbdEnergyDev &en, const double *__restrict__ x, int slot,
                                                    int slice, int lane, int sub, int G, bool valid, const V3d &xo,
                                                    V3d &g, Sym3 &H)
{
    if constexpr ((MASK >> KIND) & 1u) {
        if (en.kind == KIND) {
            if constexpr (HasIncidenceRecord<EnergyKind<KIND>>::value) {
                if (en.ellOffsets) {
                    accumulateEll<KIND>(en, x, slice, lane, xo, g, H);
                    return;
                }
            }
            if (valid) accumulateKind<KIND>(en, x, slot, sub, G, g, H);
        }
    }
}

template <unsigned MASK, int G>
__global__ void __launch_bounds__(128) vbdPassKernel(const VbdArgs a, int begin, int end, int sliceBase,
                                                     double invdtSquared)
{
    constexpr int VW = 32 / G;  // vertices per warp
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const int sub = lane / VW;  // which of the vertex's G lanes
    const int slot = begin + warp * VW + (lane % VW);
    if (begin + warp * VW >= end) return;  // whole warp past the colour
    const bool valid = slot < end;
    const int slice = sliceBase + warp;
    V3d xo(0.0, 0.0, 0.0);
    if (valid) {
        const size_t o = 3 * static_cast<size_t>(slot);
        xo = V3d(a.x[o], a.x[o + 1], a.x[o + 2]);
    }
    V3d f(0.0, 0.0, 0.0);
    Sym3 H{0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    for (int e = 0; e < a.nEnergies; ++e) {
        const VbdEnergyDev &en = a.e[e];
        accumulateIfPresent<MASK, K_SPRING>(en, a.x, slot, slice, lane, sub, G, valid, xo, f, H);
        accumulateIfPresent<MASK, K_SPRING_CONTINUUM>(en, a.x, slot, slice, lane, sub, G, valid, xo, f, H);
        accumulateIfPresent<MASK, K_DISCRETE_BENDING>(en, a.x, slot, slice, lane, sub, G, valid, xo, f, H);
        accumulateIfPresent<MASK, K_QUADRATIC_BENDING>(en, a.x, slot, slice, lane, sub, G, valid, xo, f, H);
        accumulateIfPresent<MASK, K_MEMBRANE_STVK>(en, a.x, slot, slice, lane, sub, G, valid, xo, f, H);
        accumulateIfPresent<MASK, K_MEMBRANE_CHOI>(en, a.x, slot, slice, lane, sub, G, valid, xo, f, H);
        accumulateIfPresent<MASK, K_MEMBRANE_BW>(en, a.x, slot, slice, lane, sub, G, valid, xo, f, H);
        accumulateIfPresent<MASK, K_TET_STVK>(en, a.x, slot, slice, lane, sub, G, valid, xo, f, H);
        accumulateIfPresent<MASK, K_TET_SNH>(en, a.x, slot, slice, lane, sub, G, valid, xo, f, H);
        accumulateIfPresent<MASK, K_ANCHOR>(en, a.x, slot, slice, lane, sub, G, valid, xo, f, H);
    }
    // lanes of one vertex sit VW apart: butterfly over the sub-lane bits
#pragma unroll
    for (int o = VW; o < 32; o <<= 1) {
        f.x += __shfl_xor_sync(0xffffffffu, f.x, o);
        f.y += __shfl_xor_sync(0xffffffffu, f.y, o);
        f.z += __shfl_xor_sync(0xffffffffu, f.z, o);
        H.xx += __shfl_xor_sync(0xffffffffu, H.xx, o);
        H.xy += __shfl_xor_sync(0xffffffffu, H.xy, o);
        H.xz += __shfl_xor_sync(0xffffffffu, H.xz, o);
        H.yy += __shfl_xor_sync(0xffffffffu, H.yy, o);
        H.yz += __shfl_xor_sync(0xffffffffu, H.yz, o);
        H.zz += __shfl_xor_sync(0xffffffffu, H.zz, o);
    }
    if (valid && sub == 0) {
        const size_t o = 3 * static_cast<size_t>(slot);
        const double mOverdtSquared = a.mass[slot] * invdtSquared;
        // f holds the accumulated energy gradient: force = -gradient; inertia term of vbd.cpp:109-112, :123
        f.x = -f.x - mOverdtSquared * (xo.x - a.xInertial[o]);
        f.y = -f.y - mOverdtSquared * (xo.y - a.xInertial[o + 1]);
        f.z = -f.z - mOverdtSquared * (xo.z - a.xInertial[o + 2]);
        H.xx += mOverdtSquared;
        H.yy += mOverdtSquared;
        H.zz += mOverdtSquared;
        const double d = det(H);
        if (d > 1e-10) {  // vbd.cpp:139
            const V3d dx = solveSym3(H, d, f);
            a.x[o] = xo.x + dx.x;
            a.x[o + 1] = xo.y + dx.y;
            a.x[o + 2] = xo.z + dx.z;
        }
    }
}

template <unsigned MASK, int G>
void launchMaskedG(const VbdArgs &args, int begin, int end, int sliceBase, double invdtSquared, cudaStream_t stream)
{
    constexpr int VW = 32 / G;
    const int n = end - begin;
    if (n <= 0) return;
    const int warps = (n + VW - 1) / VW;
    const int threads = 128;
    const int blocks = (warps * 32 + threads - 1) / threads;
    vbdPassKernel<MASK, G><<<blocks, threads, 0, stream>>>(args, begin, end, sliceBase, invdtSquared);
}

template <unsigned MASK>
void launchMasked(const VbdArgs &args, int lanes, int begin, int end, int sliceBase, double invdtSquared,
                  cudaStream_t stream)
{
    switch (lanes) {
    case 1: launchMaskedG<MASK, 1>(args, begin, end, sliceBase, invdtSquared, stream); break;
    case 2: launchMaskedG<MASK, 2>(args, begin, end, sliceBase, invdtSquared, stream); break;
    case 8: launchMaskedG<MASK, 8>(args, begin, end, sliceBase, invdtSquared, stream); break;
    default: launchMaskedG<MASK, 4>(args, begin, end, sliceBase, invdtSquared, stream); break;
    }
}

constexpr unsigned bit(int k)
{
    return 1u << k;
}
constexpr unsigned kMaskCloth = bit(K_DISCRETE_BENDING) | bit(K_MEMBRANE_STVK) | bit(K_ANCHOR);
constexpr unsigned kMaskClothQuad = bit(K_QUADRATIC_BENDING) | bit(K_SPRING) | bit(K_ANCHOR);
constexpr unsigned kMaskSnh = bit(K_TET_SNH) | bit(K_ANCHOR);
constexpr unsigned kMaskTetStvk = bit(K_TET_STVK) | bit(K_ANCHOR);
constexpr unsigned kMaskAll = (1u << K_COUNT) - 1u;
}  // namespace

void launchVbdPass(const VbdArgs &args, unsigned kindMask, int lanes, int begin, int end, int sliceBase,
                   double invdtSquared, cudaStream_t stream)
{
    auto subset = [&](unsigned m) { return (kindMask & ~m) == 0; };
    if (subset(kMaskCloth))
        launchMasked<kMaskCloth>(args, lanes, begin, end, sliceBase, invdtSquared, stream);
    else if (subset(kMaskSnh))
        launchMasked<kMaskSnh>(args, lanes, begin, end, sliceBase, invdtSquared, stream);
    else if (subset(kMaskTetStvk))
        launchMasked<kMaskTetStvk>(args, lanes, begin, end, sliceBase, invdtSquared, stream);
    else if (subset(kMaskClothQuad))
        launchMasked<kMaskClothQuad>(args, lanes, begin, end, sliceBase, invdtSquared, stream);
    else
        launchMasked<kMaskAll>(args, lanes, begin, end, sliceBase, invdtSquared, stream);
}

// ---- single-thread evaluator for the per-element parity tests (energy.h:240-258) -----------------
namespace
{
template <int KIND>
__device__ void evalOne(const double *x, const int32_t *ids, const double *rec, int local, double *out12, bool fast)
{
    using EK = EnergyKind<KIND>;
    if constexpr (HasHandDerived<EK>::value) {
        if (fast) {
            V3d xs[EK::S];
            for (int s = 0; s < EK::S; ++s) xs[s] = V3d(x[3 * ids[s]], x[3 * ids[s] + 1], x[3 * ids[s] + 2]);
            V3d g(0.0, 0.0, 0.0);
            Sym3 H{0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
            EK::vertexGradHess(xs, rec, local, g, H);
            out12[0] = -g.x, out12[1] = -g.y, out12[2] = -g.z;
            out12[3] = H.xx, out12[4] = H.xy, out12[5] = H.xz;
            out12[6] = H.xy, out12[7] = H.yy, out12[8] = H.yz;
            out12[9] = H.xz, out12[10] = H.yz, out12[11] = H.zz;
            return;
        }
    }
    V3<D2> xs[EK::S];
    for (int s = 0; s < EK::S; ++s) {
        const int id = ids[s];
        const bool own = (s == local);
        xs[s] = V3<D2>(D2(x[3 * id], own ? 0 : -1), D2(x[3 * id + 1], own ? 1 : -1), D2(x[3 * id + 2], own ? 2 : -1));
    }
    const D2 E = stencilEnergy<KIND, D2>(xs, rec);
    for (int i = 0; i < 3; ++i) out12[i] = -E.g[i];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) out12[3 + 3 * i + j] = E.hess(i, j);
}

__global__ void evalVertexKernel(int kind, const double *x, const int32_t *ids, const double *rec, int local,
                                 double *out12, bool fast)
{
    switch (kind) {
    case K_SPRING: evalOne<K_SPRING>(x, ids, rec, local, out12, fast); break;
    case K_SPRING_CONTINUUM: evalOne<K_SPRING_CONTINUUM>(x, ids, rec, local, out12, fast); break;
    case K_DISCRETE_BENDING: evalOne<K_DISCRETE_BENDING>(x, ids, rec, local, out12, fast); break;
    case K_QUADRATIC_BENDING: evalOne<K_QUADRATIC_BENDING>(x, ids, rec, local, out12, fast); break;
    case K_MEMBRANE_STVK: evalOne<K_MEMBRANE_STVK>(x, ids, rec, local, out12, fast); break;
    case K_MEMBRANE_CHOI: evalOne<K_MEMBRANE_CHOI>(x, ids, rec, local, out12, fast); break;
    case K_MEMBRANE_BW: evalOne<K_MEMBRANE_BW>(x, ids, rec, local, out12, fast); break;
    case K_TET_STVK: evalOne<K_TET_STVK>(x, ids, rec, local, out12, fast); break;
    case K_TET_SNH: evalOne<K_TET_SNH>(x, ids, rec, local, out12, fast); break;
    case K_ANCHOR: evalOne<K_ANCHOR>(x, ids, rec, local, out12, fast); break;
    default: break;
    }
}
}  // namespace

void launchEvalVertexForceHessian(int kind, const double *x, const int32_t *stencil, const double *record, int local,
                                  double *out12, bool handDerived, cudaStream_t stream)
{
    evalVertexKernel<<<1, 1, 0, stream>>>(kind, x, stencil, record, local, out12, handDerived);
}
}  // namespace spgx
