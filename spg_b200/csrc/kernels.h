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
};

struct VbdArgs {
    double *x;                 // 3 doubles per slot
    const double *xInertial;   // 3 doubles per slot
    const double *mass;        // per slot
    int nEnergies;
    VbdEnergyDev e[kMaxEnergies];
};

// vbd.cpp:99-147 for the slots [begin, end) of one colour
void launchVbdPass(const VbdArgs &args, unsigned kindMask, int begin, int end, double invdtSquared,
                   cudaStream_t stream);

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
