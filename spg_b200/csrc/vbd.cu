// Vertex Block Descent colour pass, generic CSR form (reference: spg/source/sim/solver/vbd.cpp:99-157), and the
// dispatcher between it and the streaming kernels (vbd_stream.cuh).
//
// One thread group of G lanes owns one vertex of the colour.  Lanes stride over the vertex's incident
// (stencil, local) entries energy by energy, each lane evaluates the element energy with second-order duals
// seeded on the OWNING vertex only (K = 3: value, 3-gradient, packed 3x3 Hessian) -- the diagonal block the
// reference extracts from a full 3S x 3S autodiff pass (energy.h:240-258) -- or the kind's hand-derived block,
// and the lanes' partial force / Hessian are reduced with warp shuffles.  This is the path every energy kind
// can take; kinds with a stream (streams.cuh) normally run through vbdStreamKernel instead.
#include "vbd_stream.cuh"

namespace spgx
{
namespace
{
template <unsigned MASK, int G>
__global__ void __launch_bounds__(128) vbdPassKernel(const VbdArgs a, int begin, int end, double invdtSquared)
{
    constexpr int VW = 32 / G;  // vertices per warp
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const int sub = lane / VW;  // which of the vertex's G lanes
    const int slot = begin + warp * VW + (lane % VW);
    if (begin + warp * VW >= end) return;  // whole warp past the colour
    const bool valid = slot < end;
    V3d xo(0.0, 0.0, 0.0);
    V3d f(0.0, 0.0, 0.0);
    Sym3 H{0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    if (valid) {
        const size_t o = 3 * static_cast<size_t>(slot);
        xo = V3d(a.x[o], a.x[o + 1], a.x[o + 2]);
        accumulateCsrEnergies<MASK>(a, slot, sub, G, f, H);
    }
    reduceSolveWrite<G>(a, slot, valid, sub, xo, f, H, invdtSquared);
}

template <unsigned MASK, int G>
void launchMaskedG(const VbdArgs &args, int begin, int end, double invdtSquared, cudaStream_t stream)
{
    constexpr int VW = 32 / G;
    const int n = end - begin;
    if (n <= 0) return;
    const int warps = (n + VW - 1) / VW;
    const int threads = 128;
    const int blocks = (warps * 32 + threads - 1) / threads;
    vbdPassKernel<MASK, G><<<blocks, threads, 0, stream>>>(args, begin, end, invdtSquared);
}

template <unsigned MASK>
void launchMasked(const VbdArgs &args, int lanes, int begin, int end, double invdtSquared, cudaStream_t stream)
{
    switch (lanes) {
    case 1: launchMaskedG<MASK, 1>(args, begin, end, invdtSquared, stream); break;
    case 2: launchMaskedG<MASK, 2>(args, begin, end, invdtSquared, stream); break;
    case 8: launchMaskedG<MASK, 8>(args, begin, end, invdtSquared, stream); break;
    default: launchMaskedG<MASK, 4>(args, begin, end, invdtSquared, stream); break;
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

bool vbdStreamComboSupported(unsigned streamMask, unsigned csrKindMask)
{
    if (streamMask == 0 || (csrKindMask & ~kKindsAnchor) != 0) return false;
    return (streamMask & ~kStreamsCloth) == 0 || (streamMask & ~kStreamsTets) == 0;
}

void launchVbdPass(const VbdArgs &args, unsigned kindMask, unsigned streamMask, int lanes, int occ, int begin, int end,
                   int sliceBase, double invdtSquared, cudaStream_t stream)
{
    if (args.nStreams > 0) {
        if ((streamMask & ~kStreamsClothSingle) == 0)
            launchStreamCloth(args, lanes, occ, begin, end, sliceBase, invdtSquared, stream);
        else if ((streamMask & ~kStreamsCloth) == 0)
            launchStreamClothWedge(args, lanes, occ, begin, end, sliceBase, invdtSquared, stream);
        else
            launchStreamTets(args, lanes, occ, begin, end, sliceBase, invdtSquared, stream);
        return;
    }
    auto subset = [&](unsigned m) { return (kindMask & ~m) == 0; };
    if (subset(kMaskCloth))
        launchMasked<kMaskCloth>(args, lanes, begin, end, invdtSquared, stream);
    else if (subset(kMaskSnh))
        launchMasked<kMaskSnh>(args, lanes, begin, end, invdtSquared, stream);
    else if (subset(kMaskTetStvk))
        launchMasked<kMaskTetStvk>(args, lanes, begin, end, invdtSquared, stream);
    else if (subset(kMaskClothQuad))
        launchMasked<kMaskClothQuad>(args, lanes, begin, end, invdtSquared, stream);
    else
        launchMasked<kMaskAll>(args, lanes, begin, end, invdtSquared, stream);
}

// ---- single-thread evaluator for the per-element parity tests (energy.h:240-258) -----------------
namespace
{
template <int KIND>
__device__ void evalOne(const double *x, const int32_t *ids, const double *rec, int local, double *out12, int fast)
{
    using EK = EnergyKind<KIND>;
    if (fast == 2 && streamTypeOf(KIND, local) >= 0) {
        // the canonical stream evaluator the streaming kernel runs, fed through the host-side canonicalisation
        const int st = streamTypeOf(KIND, local);
        int32_t nbr[3];
        double q[16];
        makeStreamRecord(KIND, ids, rec, local, nbr, q);
        V3d xn[3];
        for (int k = 0; k < streamNI(st); ++k) xn[k] = V3d(x[3 * nbr[k]], x[3 * nbr[k] + 1], x[3 * nbr[k] + 2]);
        V3d g(0.0, 0.0, 0.0);
        Sym3 H{0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        evalStream(st, V3d(x[3 * ids[local]], x[3 * ids[local] + 1], x[3 * ids[local] + 2]), xn, q, g, H);
        out12[0] = -g.x, out12[1] = -g.y, out12[2] = -g.z;
        out12[3] = H.xx, out12[4] = H.xy, out12[5] = H.xz;
        out12[6] = H.xy, out12[7] = H.yy, out12[8] = H.yz;
        out12[9] = H.xz, out12[10] = H.yz, out12[11] = H.zz;
        return;
    }
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
                                 double *out12, int fast)
{
    switch (kind) {
#define SPGX_KIND(NAME, ID, BYTES)                                                                                     \
    case NAME: evalOne<NAME>(x, ids, rec, local, out12, fast); break;
#include "kinds.def"
#undef SPGX_KIND
    default: break;
    }
}
}  // namespace

void launchEvalVertexForceHessian(int kind, const double *x, const int32_t *stencil, const double *record, int local,
                                  double *out12, int handDerived, cudaStream_t stream)
{
    evalVertexKernel<<<1, 1, 0, stream>>>(kind, x, stencil, record, local, out12, handDerived);
}
}  // namespace spgx
