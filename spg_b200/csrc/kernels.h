// Launcher interface between the C-ABI layer (api.cu) and the kernel translation units.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace spgx
{
constexpr int kMaxEnergies = 8;

// One energy of a group as the VBD kernels see it.
struct VbdEnergyDev {
    int kind;
    int nStencils;
    const int32_t *stencils;   // S slot ids per stencil
    const double *records;     // NP doubles per stencil (energies.cuh)
    const uint32_t *rowptr;    // nVerts + 1, indexed by slot
    const uint32_t *entries;   // (stencil << 2) | local, ascending per row
    // streaming layout (kinds with an owner-centric incidence record; null otherwise): a sliced ELL over
    // warp-slices of 32 / G vertices.  Step t of a slice holds one incidence per lane:
    //   ids    [t][NI][32] uint32 (0xFFFFFFFF in field 0 = padding), params [t][NQ][32] double
    const uint32_t *ellOffsets;  // per warp-slice first step, nSlices + 1
    const uint32_t *ellIds;
    const double *ellParams;
};

struct VbdArgs {
    double *x;                 // 3 doubles per slot
    const double *xInertial;   // 3 doubles per slot
    const double *mass;        // per slot
    int nEnergies;
    VbdEnergyDev e[kMaxEnergies];
};

// vbd.cpp:99-147 for the slots [begin, end) of one colour
// lanes: threads per vertex (1, 2, 4 or 8); sliceBase: index of the colour's first warp-slice
void launchVbdPass(const VbdArgs &args, unsigned kindMask, int lanes, int begin, int end, int sliceBase,
                   double invdtSquared, cudaStream_t stream);

// vbd.cpp:49-75 / xpbd.cpp:48-74: xOld <- x; v += dt g; x += v dt; (xInertial <- x when not null)
void launchPredict(double *x, double *v, double *xOld, double *xInertial, int nVerts, double dt, const double *g3,
                   cudaStream_t stream);
// particleGroup.cpp:65-72: v <- (x - xOld) * invdt
void launchFinalize(const double *x, const double *xOld, double *v, int nVerts, double invdt, cudaStream_t stream);
// dst[perm[i]] = src[i] (scatter) or dst[i] = src[perm[i]] (gather) over 3-vectors
void launchPermute3(const double *src, double *dst, const int32_t *perm, int n, bool gather, cudaStream_t stream);

// energy.h:117-153 (springEnergy.cpp:48-68 for K_SPRING) for stencils [begin, end) of one colour.
// lambda == nullptr: reference small-steps form; otherwise NC doubles per stencil are accumulated.
void launchXpbdPass(int kind, double *x, const double *invMass, const int32_t *stencils, const double *records,
                    double *lambda, int begin, int end, double dt, bool serial, cudaStream_t stream);

// single-thread evaluators for the per-element parity tests
void launchEvalVertexForceHessian(int kind, const double *x, const int32_t *stencil, const double *record, int local,
                                  double *out12, bool handDerived, cudaStream_t stream);
void launchEvalConstraints(int kind, const double *x, const int32_t *stencil, const double *record, double *outC,
                           double *outGrad, cudaStream_t stream);
}  // namespace spgx
