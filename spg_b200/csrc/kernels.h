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
    int streamed;              // 1: every incidence of this energy lives in a stream below (the CSR path skips it)
};

constexpr int kNumStreamTypes = 6;  // StreamType (streams.cuh)
constexpr int kSliceDescWords = 8;
// One owner-centric incidence stream (streams.cuh): a sliced ELL over warp-slices of VW = 32 / G consecutive slots of
// one colour.  Step t of a slice holds one incidence per lane (lane = sub * VW + v carries incidence t * G + sub of the
// slice's vertex v):  idx [t][32] uint32 = the NI neighbour positions as 8-bit indices into the owner's unique-neighbour
// list (0xFFFFFFFF = padding),  params [t][NQ][32] double.  A group has at most one stream per type.
struct VbdStreamDev {
    const uint32_t *idx;  // null: the group has no stream of this type
    const double *params;
};

constexpr int kMaxPeers = 8;
// Peer-memory halo exchange fused into the colour pass (comm.cu): every owned boundary vertex is stored straight into
// the halo slot of each neighbouring rank's position array (CUDA IPC mapping, NVLink), the last warp of the pass to
// finish its stores raises an event counter in the neighbours' flag arrays, and the slices of the next pass that read
// halo slots wait for the neighbours' counters.  Events of one substep: 0 = predict done, j = colour pass j done
// (j = 1..P); the value written for event j is *epochBase + j + 1; *epochBase advances by P + 1 per substep.
struct VbdHaloDev {
    int nPeers;                      // 0: single domain, or the NCCL exchange is used instead
    int event;                       // event this pass raises (waits for event - 1)
    int pushWarps;                   // warps of this pass that hold push entries (the last one to finish raises the event)
    double *peerX[kMaxPeers];        // the neighbours' position arrays
    uint32_t *peerFlag[kMaxPeers];   // this rank's entry in each neighbour's flag array
    const uint32_t *myFlag;          // this rank's flag array, one entry per neighbour
    const uint2 *push;               // (local slot, neighbour << 28 | slot on the neighbour), in slice order
    uint32_t *pushDone;              // warps of the running pass that have finished their stores
    const uint32_t *epochBase;
    uint32_t *error;                 // set to 1 when a wait ran into the timeout
    unsigned long long timeoutNs;
    int debug;                       // timing experiments only (SPGX_DEBUG_HALO, results are wrong): bit 0 skip the waits,
                                     // bit 1 skip the system fences of the pushing warps, bit 2 skip the remote stores, bit 3 skip the
                                     // signal / wait-advance kernels around the sweeps, bit 4 skip the push code of boundary slices
};

struct VbdArgs {
    double *x;                 // 3 doubles per slot
    const double *xInertial;   // 3 doubles per slot
    const double *mass;        // per slot
    int nEnergies;
    VbdEnergyDev e[kMaxEnergies];
    // streaming layout (nStreams == 0: CSR path only)
    int nStreams;
    VbdStreamDev s[kNumStreamTypes];  // indexed by StreamType
    // per warp-slice descriptor, kSliceDescWords uint32 each, nSlices + 1 entries (the last one closes the ranges):
    //   [0] first slot, [1] first row of the unique-neighbour table, [streamDescWord(t)] first step of stream type t (2 + t;
    //       the wedge stream of a cloth takes the word of ST_TET_STVK),
    //   [7] bit 0: some vertex of the slice has incidences on the CSR path; bit 1: the slice reads halo slots;
    //       bit 2: uniform slice (full, every vertex the same number of incidences per stream, a multiple of 2 G);
    //       bits 3..31: number of push entries (VbdHaloDev::push) of all earlier slices
    const uint32_t *sliceDesc;
    const uint32_t *nbrIds;  // [row][VW] slot ids; row k of a slice = k-th unique neighbour of each of its vertices
    int maxNbr;              // largest unique-neighbour count of any vertex (sizes the shared-memory tile)
    int nbrFixed;            // 1: nbrIds has a fixed stride of maxNbr rows per slice (rows past a slice's count hold slot 0)
    int nSlices;             // all colours
    int lookahead;           // slices ahead whose streams are pulled into L2 (0: off)
    int pdl;                 // 1: launch colour passes with programmatic stream serialization
    int carveout;            // preferred shared-memory carveout of the streaming kernel in percent (-1: maximum shared)
    size_t l2PersistBytes;   // > 0: the position array is given a persisting L2 access-policy window of this many bytes
    float l2HitRatio;        // fraction of the window that may persist (set-aside / window when the array is larger)
    VbdHaloDev halo;
};

// can launchVbdPass run this combination of stream types + CSR kinds through the streaming kernel?
bool vbdStreamComboSupported(unsigned streamMask, unsigned csrKindMask);

// vbd.cpp:99-147 for the slots [begin, end) of one colour
// lanes: threads per vertex (1, 2, 4 or 8); sliceBase: index of the colour's first warp-slice
// kindMask: kinds that go through the CSR path; streamMask: stream types present; occ: resident blocks per SM hint
void launchVbdPass(const VbdArgs &args, unsigned kindMask, unsigned streamMask, int lanes, int occ, int begin, int end,
                   int sliceBase, double invdtSquared, cudaStream_t stream);

// vbd.cpp:49-75 / xpbd.cpp:48-74: xOld <- x; v += dt g; x += v dt; (xInertial <- x when not null)
void launchPredict(double *x, double *v, double *xOld, double *xInertial, int nVerts, double dt, const double *g3,
                   cudaStream_t stream);
// particleGroup.cpp:65-72: v <- (x - xOld) * invdt
void launchFinalize(const double *x, const double *xOld, double *v, int nVerts, double invdt, cudaStream_t stream);
// launchFinalize of one substep and launchPredict (xInertial == nullptr) of the next in one launch, same roundings
void launchFinalizePredict(double *x, double *v, double *xOld, int nVerts, double dt, double invdt, const double *g3,
                           cudaStream_t stream);
// peer-memory halo exchange outside the colour passes (comm.cu): raise event `event` / wait for it and advance the epoch
void launchHaloSignal(const VbdHaloDev &h, int event, cudaStream_t stream);
void launchHaloWaitAdvance(const VbdHaloDev &h, int event, uint32_t *epochBase, cudaStream_t stream);
// XPBD: store entries [begin, end) of h.push into the neighbours' halo slots, then raise event `event` (one block that only
// raises it when the range is empty)
void launchXpbdHaloPush(const VbdHaloDev &h, const double *x, int begin, int end, int event, cudaStream_t stream);
// non-finite components of x / v over the first nVerts slots: out3 = [count x, count v, smallest offending vertex id + 1];
// the caller initialises out3 to {0, 0, ~0}
void launchCheckFinite(const double *x, const double *v, const int32_t *slot2vert, int nVerts, unsigned long long *out3,
                       cudaStream_t stream);
// dst[i] = src[idx[i]]
void launchGatherScalar(const double *src, const int32_t *idx, long long n, double *dst, cudaStream_t stream);
// dst[perm[i]] = src[i] (scatter) or dst[i] = src[perm[i]] (gather) over 3-vectors
void launchPermute3(const double *src, double *dst, const int32_t *perm, int n, bool gather, cudaStream_t stream);

// energy.h:117-153 (springEnergy.cpp:48-68 for K_SPRING) for stencils [begin, end) of one colour.
// lambda == nullptr: reference small-steps form; otherwise NC doubles per stencil are accumulated.
// stencilW: the inverse masses of every stencil's vertices, S per stencil in the order of `stencils` (gathered once at
// finalize: a pass then requests ids, inverse masses and record TOGETHER instead of the masses after the ids).
// halo (may be null): a partitioned group under the peer-memory exchange; every block waits for the neighbours' event
// halo->event - 1 before it reads positions.
void launchXpbdPass(int kind, double *x, const double *stencilW, const int32_t *stencils, const double *records,
                    double *lambda, int begin, int end, double dt, bool serial, cudaStream_t stream, const VbdHaloDev *halo = nullptr);

// ---- implicit-Euler family and symplectic Euler (newton.cu; reference: spg/source/sim/solver/implicitEulerBase.cpp) ----
struct NewtonEnergyDev {
    int kind;
    int nStencils;
    const int32_t *stencils;  // S vertex ids per stencil (identity ordering: slot == vertex)
    const double *records;    // NP doubles per stencil
    const uint32_t *rowptr;   // nVerts + 1
    const uint32_t *entries;  // (stencil << 2) | local, ascending per vertex
    double *elemH;            // per stencil S x S blocks of 3 x 3, block (la, lb) at ((la * S + lb) * 9), row-major inside
};
struct NewtonArgs {
    const double *x;
    const double *mass;
    int nVerts;
    int nEnergies;
    NewtonEnergyDev e[kMaxEnergies];
    // block-sparse rows of the assembled Hessian: row v holds the blocks of v and of every vertex sharing a stencil
    // with it, columns ascending
    const uint32_t *bsrRowptr;  // nVerts + 1
    const int32_t *bsrCol;      // per block
    const int32_t *bsrRow;      // per block
    const uint32_t *diagPos;    // per vertex: its diagonal block
    double *bsrH;               // 9 x nnzb doubles, component-major: entry (i, j) of block t at bsrH[(3 i + j) * nnzb + t]
    long long nnzb;
    // gather lists of the assembly (built once with the pattern): block t sums the arena blocks asmSrc[asmPtr[t] .. asmPtr[t + 1])
    const uint32_t *asmPtr;
    const uint32_t *asmSrc;
    const double *elemArena;
};
// getSystemForce (implicitEulerBase.cpp:102-143): f_v = g * m_v - sum over the incident stencils of dE/dx_v (owner-centric
// gather over the vertex incidence table: no atomics, fixed summation order). g3: host pointer, 3 doubles.
void launchSystemForce(const NewtonArgs &a, const double *g3, double *f, cudaStream_t stream);
// energyHessian of every stencil of every energy (energy.h:354-362) into e[].elemH
void launchElementHessians(const NewtonArgs &a, cudaStream_t stream);
// getSystemStiffnessMatrix (:176-221) without the triplet list: every block of the sparse Hessian gathers its
// contributions from the element Hessians in a fixed order (energies in insertion order, stencils ascending)
void launchBsrAssemble(const NewtonArgs &a, cudaStream_t stream);

constexpr int kCgBlocks = 592;   // 148 SMs x 4: fixed grid of the reductions, so partial sums have a fixed order
constexpr int kCgThreads = 256;
struct CgState {
    double absNew, rhsNorm2, threshold, residualNorm2;
    int done, iters;
};
struct CgArgs {
    // A = cM * diag(mass) + cH * H (H: the assembled block-sparse Hessian)
    const uint32_t *bsrRowptr;
    const int32_t *bsrCol;
    const double *bsrH;  // component-major (NewtonArgs::bsrH)
    long long nnzb;
    const uint32_t *diagPos;
    const double *mass;
    double cM, cH;
    int nVerts;
    double *x, *r, *p, *z, *tmp, *invdiag;  // 3 * nVerts each
    double *partials;                       // 3 * kCgBlocks: [p . Ap | r . z | r . r]
    CgState *state;                         // 2 entries (read [it & 1], written [(it + 1) & 1])
};
// y = (cM diag(mass) + cH H) x
void launchBsrSpmv(const CgArgs &a, const double *x, double *y, cudaStream_t stream);
// Eigen::ConjugateGradient with the Jacobi preconditioner (ConjugateGradient.h:29-87), device-resident:
// launchCgInit: invdiag, x = 0, r = b, p = invdiag r, state[0]; launchCgIteration: iteration `it` (3 kernels; a no-op
// once state.done is set)
void launchCgInit(const CgArgs &a, const double *b, double tol, cudaStream_t stream);
void launchCgIteration(const CgArgs &a, int it, cudaStream_t stream);
// out = alpha * a + beta * b over n doubles (out may alias a or b)
void launchAxpby(double *out, double alpha, const double *a, double beta, const double *b, long long n, cudaStream_t stream);
// out = c1 * f - c2 * mass[vertex] * (a - b) and its Euclidean norm (deterministic two-stage reduction through `partials`,
// kCgBlocks doubles) into *normOut: the Newton residuals dt f - M (v - v0) (implicitEulerNewtonRobust.cpp:66, :88) and
// f - M (x - x0) / dt^2 (quasiStaticNewtonRobust.cpp:67, :86)
void launchResidualNorm(double *out, double c1, const double *f, double c2, const double *mass, const double *a, const double *b,
                        int nVerts, double *partials, double *normOut, cudaStream_t stream);
// out += c * mass[vertex] * (a - b): the M (v0 - v-1) term of BDF2 (bdf2.cpp:53) without a mass matrix
void launchAddMassDiff(double *out, double c, const double *mass, const double *a, const double *b, int nVerts, cudaStream_t stream);
// x += dx; v = (x - x0) * invdt  (integrateObjectsVelocitiesFromDx, implicitEulerBase.cpp:76-88)
void launchApplyDx(double *x, double *v, const double *dx, const double *x0, double invdt, long long n, cudaStream_t stream);
// SimplecticEuler::step velocity update (simplecticEuler.cpp:23-48): v += dt g; v += w * fInternal * dt; then x += v dt
void launchSymplecticUpdate(double *x, double *v, const double *fInternal, const double *invMass, int nVerts, double dt,
                            const double *g3, cudaStream_t stream);

// single-thread evaluators for the per-element parity tests
void launchEvalVertexForceHessian(int kind, const double *x, const int32_t *stencil, const double *record, int local,
                                  double *out12, int handDerived, cudaStream_t stream);
void launchEvalConstraints(int kind, const double *x, const int32_t *stencil, const double *record, double *outC,
                           double *outGrad, cudaStream_t stream);
}  // namespace spgx
