/* spgx — C ABI of the B200-native SPG solver hot path (VBD sweep, XPBD projection, shared element evaluators).
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++ or torch types.  The reference
 * (alexrodag/spg @ 75122ce) has no FFI of its own; every entry point below names the reference C++
 * interface it stands in for (paths relative to the reference root).  The C++ facade in
 * include/spgx/spgx.hpp re-creates the reference's class surface (ParticleGroup, the energy classes'
 * addStencil / preparePrecomputations, solver::VBD / solver::XPBD with step / addObject / setDt / ...)
 * on top of these calls.
 *
 * Conventions: every function returning int returns 0 on success and a non-zero status otherwise;
 * spgx_last_error() then returns the message a std::runtime_error in the reference would carry (the
 * facade re-throws it).  Nothing throws across this boundary.  All host arrays are caller-owned and
 * copied during the call; all device memory is owned by the library behind the opaque handles.
 * Vertex data is fp64, 3 doubles per vertex (x, y, z); indices are int32.
 * There is no CPU fallback: without a CUDA device every compute call fails with SPGX_ERR_CUDA.
 */
#ifndef SPGX_H
#define SPGX_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPGX_OK 0
#define SPGX_ERR_INVALID 1 /* std::runtime_error in the reference (validation) */
#define SPGX_ERR_CUDA 2    /* CUDA runtime / driver failure, or no device */
#define SPGX_ERR_STATE 3   /* call order violation (e.g. step before finalize) */
#define SPGX_ERR_COMM 4    /* NCCL failure */

/* Energy kinds (one per reference energy class on the hot path). Model parameters per stencil:
 *  SPRING            spg/source/sim/energy/springEnergy.cpp:33            {restLength, stiffness}
 *  SPRING_CONTINUUM  .../springContinuumEnergy.cpp:19                     {restLength, stiffness}
 *  DISCRETE_BENDING  .../discreteBendingEnergy.cpp:33                     {restTheta, stiffness}
 *  QUADRATIC_BENDING .../quadraticBendingEnergy.cpp:32                    {restTheta (must be 0), stiffness}
 *  MEMBRANE_STVK     .../membraneStvkEnergy.cpp:31                        {young, poisson}
 *  MEMBRANE_CHOI     .../membraneChoiEnergy.cpp:31                        {kWeft, kWarp, kBias}
 *  MEMBRANE_BW       .../membraneBaraffWitkinEnergy.cpp:29                {kWeft, kWarp, kBias}
 *  TET_STVK          .../stvkEnergy.cpp:68                                {young, poisson}
 *  TET_SNH           .../stableNeoHookeanEnergy.cpp:64                    {young, poisson}
 *  ANCHOR            .../springAnchorEnergy.cpp:29                        {anchor.x, anchor.y, anchor.z, stiffness}
 *  BENDING_BW        .../baraffWitkinBendingEnergy.cpp:40                 {restTheta, stiffness}
 *  SPRING_SQUARED_CONTINUUM .../springSquaredContinuumEnergy.cpp:20       {restLength, stiffness}
 *  TET_SSNH          .../stableSquaredNeoHookeanEnergy.cpp:34             {young, poisson}
 *  STATIC_LOAD       spg/include/spg/sim/energy/staticLoadEnergy.h:11     {dir.x, dir.y, dir.z, stiffness}  (VBD only: its
 *                    constraint form throws in the reference, staticLoadEnergy.cpp:8-13, and so does spgx_xpbd_step)
 */
enum spgx_kind {
    SPGX_SPRING = 0,
    SPGX_SPRING_CONTINUUM = 1,
    SPGX_DISCRETE_BENDING = 2,
    SPGX_QUADRATIC_BENDING = 3,
    SPGX_MEMBRANE_STVK = 4,
    SPGX_MEMBRANE_CHOI = 5,
    SPGX_MEMBRANE_BW = 6,
    SPGX_TET_STVK = 7,
    SPGX_TET_SNH = 8,
    SPGX_ANCHOR = 9,
    SPGX_BENDING_BW = 10,
    SPGX_SPRING_SQUARED_CONTINUUM = 11,
    SPGX_TET_SSNH = 12,
    SPGX_STATIC_LOAD = 13,
    SPGX_KIND_COUNT = 14
};

enum spgx_solver_kind {
    SPGX_VBD = 0,        /* spg/source/sim/solver/vbd.cpp:12-195 */
    SPGX_XPBD = 1,       /* spg/source/sim/solver/xpbd.cpp:16-131, useParallelGaussSeidel = true */
    SPGX_XPBD_SERIAL = 2, /* same with useParallelGaussSeidel = false (xpbd.cpp:96-102): stencils in index order */
    SPGX_NEWTON = 3 /* the implicit-Euler family (implicitEulerBase.cpp) and symplectic Euler: spgx_implicit_step,
                       spgx_symplectic_step, spgx_system_force, spgx_system_stiffness, spgx_linear_solve */
};

/* which member of the implicit-Euler family spgx_implicit_step runs (one linearised solve per substep each) */
enum spgx_implicit_kind {
    SPGX_IE_BARAFF_WITKIN = 0, /* spg/source/sim/solver/implicitEulerBaraffWitkin.cpp:11-59: (M - dt^2 K) dv = dt (f + dt K v0) */
    SPGX_IE_NEWTON_DX = 1,     /* .../implicitEulerNewtonDx.cpp:11-64: (M / dt^2 - K) dx = f at the inertial guess */
    SPGX_IE_NEWTON_DV = 2,     /* .../implicitEulerNewtonDv.cpp:11-67: (M - dt^2 K) dv = dt f at the inertial guess */
    SPGX_IE_BDF2 = 3,          /* .../bdf2.cpp:10-76: (M - 4/9 dt^2 K) dv = 1/3 M (v0 - v-1) + 2/3 dt (f + 1/3 K (x0 - x-1 + 2 dt v0));
                                  the previous state (x-1, v-1) lives with the group, spgx_reset forgets it (bdf2.cpp:78-83) */
    SPGX_IE_STATIC_NEWTON = 4, /* .../staticNewton.cpp:10-51: K dx = -f, positions only */
    SPGX_IE_QUASI_STATIC_NEWTON = 5, /* .../quasiStaticNewton.cpp:10-55: (M - dt^2 K) dx = dt^2 f, positions only */
    /* the two adaptive solvers: up to 5 Newton iterations per substep, each with a halving line search (up to 5 trials)
     * on the residual norm; a substep that does not reach the residual threshold is rolled back and the solver's own time
     * step shrinks by 0.75, five good substeps in a row let it grow back (state kept with the group, cleared by spgx_reset) */
    SPGX_IE_NEWTON_ROBUST = 6,             /* .../implicitEulerNewtonRobust.cpp:11-130, threshold 1e-5 on |dt f - M (v - v0)| */
    SPGX_IE_QUASI_STATIC_NEWTON_ROBUST = 7 /* .../quasiStaticNewtonRobust.cpp:11-128, threshold 1e-2 on |f - M (x - x0) / dt^2| */
};

typedef struct spgx_ctx spgx_ctx;     /* one CUDA device + stream (+ optional NCCL communicator) */
typedef struct spgx_group spgx_group; /* a solver's copy of one ParticleGroup with its energies */

typedef struct spgx_timings {
    /* the four phases the reference prints under Verbosity::Performance (vbd.cpp:184-194, xpbd.cpp:125-130), ms */
    double inertial_ms;
    double solve_ms; /* "Coordinate block descent" / "Projection" */
    double update_ms;
    double total_ms;
    int64_t kernel_launches; /* kernels launched by the last step call */
} spgx_timings;

/* ---- library ------------------------------------------------------------------------------- */
const char *spgx_last_error(void); /* thread-local message of the last failed call */
const char *spgx_version(void);
int spgx_kind_stencil_size(int kind);     /* StencilBlockEnergy<S, C>::stencilSize(), energy.h:100 */
int spgx_kind_num_constraints(int kind);  /* TnConstraints of the reference class */
int spgx_kind_num_model_params(int kind); /* doubles per stencil expected by spgx_group_add_energy */
int spgx_kind_num_record_params(int kind);/* doubles per stencil of the precomputed record (tests) */

/* ---- pure host functions (no device needed) ------------------------------------------------- */
/* coloring::colorVertices, spg/source/utils/graphColoring.cpp:72-84 (bit-exact, linear time).
 * sets: nSets flat stencil arrays (entries[s], nEntries[s], stencilSizes[s]); colors: nVerts ints out. */
int spgx_color_vertices(int nVerts, int nSets, const int32_t *const *entries, const int64_t *nEntries,
                        const int32_t *stencilSizes, int32_t *colors);
/* coloring::colorStencils, graphColoring.cpp:51-70. colors: nEntries / stencilSize ints out. */
int spgx_color_stencils(const int32_t *entries, int64_t nEntries, int stencilSize, int32_t *colors);
/* spgx_color_vertices built on the device of `ctx`: the greedy colouring evaluated level by
 * level over its dependency DAG -- a node's colour depends only on its lower-indexed neighbours' -- hence the same colours
 * bit for bit (spg_b200/csrc/colour.cu). *rounds (may be NULL): the number of levels. spgx_group_finalize(SPGX_VBD) uses it
 * for groups of 500 k vertices and more (SPGX_DEVICE_COLOURING=0 / 1 overrides). */
int spgx_color_vertices_device(spgx_ctx *ctx, int nVerts, int nSets, const int32_t *const *entries, const int64_t *nEntries,
                               const int32_t *stencilSizes, int32_t *colors, int *rounds);
/* spgx_color_vertices_device for callers that may not allocate. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
 * it). It calls no cudaMalloc / cudaFree / allocating routine: its scratch is one static device array inside the library
 * (about 13 MiB on every GPU that loads it, until exit), which limits it to SPGX_NOALLOC_MAX_VERTS vertices and
 * SPGX_NOALLOC_MAX_ENTRIES stencil entries over all sets; larger inputs fail with a message naming the limit.
 * d_entries: nSets DEVICE arrays (the pointer array itself, nEntries and stencilSizes are host memory); d_colors: nVerts ints
 * on the DEVICE; *rounds (host, may be NULL). Same colours as spgx_color_vertices / _device, bit for bit. Runs on the
 * context's stream and returns after it completes. All calls on a device share the scratch: calls that overlap on the GPU
 * (other threads or contexts) race. */
#define SPGX_NOALLOC_MAX_VERTS 262144
#define SPGX_NOALLOC_MAX_ENTRIES 2097152
int spgx_color_vertices_noalloc(spgx_ctx *ctx, int nVerts, int nSets, const int32_t *const *d_entries, const int64_t *nEntries,
                                const int32_t *stencilSizes, int32_t *d_colors, int *rounds);
/* addStencil + preparePrecomputations of one energy kind on rest positions x0: writes
 * nStencils * spgx_kind_num_record_params(kind) doubles. Same validation as the reference. */
int spgx_prepare_energy(int kind, int nStencils, const int32_t *stencils, const double *modelParams,
                        int nVerts, const double *x0, double *records);

/* ---- context ------------------------------------------------------------------------------- */
int spgx_ctx_create(int device, spgx_ctx **out);
void spgx_ctx_destroy(spgx_ctx *ctx);
int spgx_ctx_sync(spgx_ctx *ctx);

/* ---- multi-GPU (one process per GPU; the reference is single-process) ------ */
/* ncclGetUniqueId: 128 bytes the caller broadcasts from rank 0 to every rank (e.g. with torch.distributed) */
int spgx_comm_unique_id(void *id128);
/* ncclCommInitRank on the context's device; collective over all ranks */
int spgx_ctx_init_comm(spgx_ctx *ctx, int rank, int nranks, const void *id128);

/* page-lock / unlock a caller-owned host buffer so spgx_upload / spgx_download run at full PCIe rate */
int spgx_host_register(void *ptr, uint64_t bytes);
int spgx_host_unregister(void *ptr);
/* page-locked host memory owned by the library (cudaHostAlloc / cudaFreeHost): what a caller without storage of its own
 * (a frame queue, a binding's staging arrays) hands to spgx_upload / spgx_download and their asynchronous forms */
int spgx_host_alloc(uint64_t bytes, void **ptr);
int spgx_host_free(void *ptr);

/* ---- particle group (BaseSolver::addObject copies the group: baseSolver.cpp:10-14) ----------- */
/* ParticleGroup::addParticle for nVerts particles (particleGroup.cpp:74-82): x current, x0 rest
 * positions, mass (inverse mass w = mass == 0 ? 0 : 1/mass). Velocities start at zero. */
int spgx_group_create(spgx_ctx *ctx, int nVerts, const double *x, const double *x0, const double *mass,
                      spgx_group **out);
void spgx_group_destroy(spgx_group *grp);
/* energy->addStencil(...) for every stencil, pGroup.addEnergy(energy), energy->preparePrecomputations(pGroup).
 * Energies are processed in the order they were added (vbd.cpp:128, xpbd.cpp:84). */
int spgx_group_add_energy(spgx_group *grp, int kind, int nStencils, const int32_t *stencils,
                          const double *modelParams, int *energyId);
/* ParticleGroup::removeEnergy (particleGroup.cpp:89-95; picking release, apps/solverCompare.cpp:1499-1502). The ids of
 * the energies added after it shift down by one, as their positions in the reference's energy vector do. The group
 * must be finalized again before the next step (requirePrecomputationUpdate, vbd.h:26 / xpbd.h:20); positions and
 * velocities are kept across it. */
int spgx_group_remove_energy(spgx_group *grp, int energyId);
/* SpringAnchorEnergy::updateAnchor (springAnchorEnergy.h:13) and friends: overwrite `count` doubles of
 * the precomputed record of one stencil starting at `offset`. Takes effect at the next step, for every solver and
 * every kind: on a group finalized for VBD whose kind runs through the streaming colour pass (discrete bending, StVK
 * membrane, tetrahedra) the stencil's canonical per-incidence records are rebuilt and patched in place as well. */
int spgx_energy_update_record(spgx_group *grp, int energyId, int stencil, int offset, const double *values,
                              int count);
int spgx_energy_get_records(spgx_group *grp, int energyId, double *records);
/* VBD::computeParallelVertexGroups + computeStencilInfoPerVertex (vbd.cpp:203-268) or
 * XPBD::computeParallelStencilGroups (xpbd.cpp:133-160), then upload. Must be called once before stepping;
 * calling it again re-does the precomputation (requirePrecomputationUpdate) and keeps positions and velocities of
 * every vertex the rank held before, owned and halo alike. (A vertex that only the NEW layout holds -- the partition
 * changed, or a new energy reaches further -- starts from the state last uploaded for it: after such a change upload
 * the global state on every rank.) */
int spgx_group_finalize(spgx_group *grp, int solverKind);

/* Domain decomposition for VBD: every rank creates the SAME global group (same vertices, same energies), then declares
 * which vertices it owns (owner[v] in [0, nranks); NULL = nranks equal contiguous index ranges, i.e. row strips of a
 * cloth grid / slabs of a beam) BEFORE spgx_group_finalize.  Colouring stays global (bit-exact with the reference);
 * the rank keeps the stencils touching its vertices, holds the other vertices of those stencils as halo copies and
 * exchanges the boundary vertices of colour c with its neighbours after colour pass c (vbd.cpp:151-157 runs the colour
 * groups one after another; vertices of one colour never share a stencil, so the result is identical to one domain).
 * spgx_upload / spgx_download keep taking global-sized arrays: upload reads the entries of the local vertices,
 * download writes the entries of the OWNED vertices only. */
int spgx_group_set_partition(spgx_group *grp, int rank, int nranks, const int32_t *owner);
/* after finalize: out8 = [nOwned, nLocal, nPeers, localLo, localHi, ownedLo, ownedHi, nColors] (global index ranges the
 * local / owned vertices span: what spgx_upload reads and spgx_download writes) */
int spgx_group_partition_info(spgx_group *grp, int64_t *out8);
/* pure host: the local view rank `rank` builds (spg_b200/csrc/partition.h). sizes: [nOwned, nLocal, nPeers, nColors,
 * nSend] out; the arrays may be NULL to query the sizes first. slot2vert: nLocal global ids (owned first);
 * peers: nPeers ranks; sendOffsets: nColors * nPeers + 1 into sendVerts (global ids, (colour, peer)-major);
 * recvSlot / recvCount: nColors * nPeers halo slot ranges. */
int spgx_partition_plan(int nVerts, int nSets, const int32_t *const *entries, const int64_t *nEntries,
                        const int32_t *stencilSizes, const int32_t *owner, int rank, int nranks, int64_t *sizes,
                        int32_t *slot2vert, int32_t *peers, int64_t *sendOffsets, int32_t *sendVerts, int64_t *recvSlot,
                        int64_t *recvCount);

/* pure host: the XPBD counterpart (spg_b200/csrc/xpbd_partition.h; SURVEY.md section 8e "next"): what rank `rank` holds and
 * exchanges when XPBD::step (xpbd.cpp:16-131) is split between ranks. A pass = one stencil colour of one energy
 * (xpbd.cpp:133-160), numbered energy-major; a rank executes every stencil touching a vertex it owns (cut stencils run on
 * every rank holding them, with identical results), and after a pass the owner of a moved vertex sends it to the ranks
 * that keep it as halo without holding the stencil that moved it. The stencil colours are computed here
 * (colorStencils) exactly as spgx_group_finalize does. sizes: [nOwned, nHalo, nPeers, nPasses, nHeld, nSend, nRecv] out; the
 * arrays may be NULL to query the sizes first. halo: nHalo global ids; peers: nPeers ranks; heldOffsets: nPasses + 1 into
 * held (stencil ids within their energy); sendOffsets / recvOffsets: nPasses * nPeers + 1 into sendVerts / recvVerts
 * (global ids, ascending, (pass, peer)-major). The device path is not built yet: XPBD groups run on one GPU. */
int spgx_xpbd_partition_plan(int nVerts, int nSets, const int32_t *const *entries, const int64_t *nEntries,
                             const int32_t *stencilSizes, const int32_t *owner, int rank, int nranks, int64_t *sizes,
                             int32_t *halo, int32_t *peers, int64_t *heldOffsets, int32_t *held, int64_t *sendOffsets,
                             int32_t *sendVerts, int64_t *recvOffsets, int32_t *recvVerts);

/* Peer-memory halo exchange (replaces the ncclSend / ncclRecv schedule once connected; the reference is single-process,
 * the dependency it serves is the colour order of vbd.cpp:151-157).  After spgx_group_finalize every rank exports a
 * blob (CUDA IPC handles of its position array and flag array + its halo layout), the caller all-gathers the blobs
 * (e.g. torch.distributed.all_gather_object) and hands ALL of them, indexed by rank, to spgx_group_p2p_connect.  From
 * then on the colour-pass kernel stores boundary vertices straight into the neighbours' halo slots over NVLink and
 * raises / waits for per-pass event counters in peer memory: no pack kernel, no NCCL launch, no host involvement.
 * blob == NULL: only *size is written.  Re-finalizing the group disconnects it. */
int spgx_group_p2p_export(spgx_group *grp, void *blob, int64_t capacity, int64_t *size);
int spgx_group_p2p_connect(spgx_group *grp, int nBlobs, const void *const *blobs, const int64_t *sizes);
/* unmap the neighbours' arrays (collective by convention: every rank disconnects, THEN groups may be destroyed, so that
 * nobody frees an array a neighbour still has mapped); the exchange falls back to NCCL */
int spgx_group_p2p_disconnect(spgx_group *grp);
enum spgx_halo_mode { SPGX_HALO_NONE = 0, SPGX_HALO_NCCL = 1, SPGX_HALO_P2P = 2 };
/* choose the exchange of the next step() (P2P needs a connected group); *mode out when not NULL */
int spgx_group_set_halo_mode(spgx_group *grp, int mode);
int spgx_group_get_halo_mode(spgx_group *grp, int *mode);
/* out8 = [stream types in use (bit mask of the streaming kernel's record types, 0: generic path) | wedge records << 8, lanes per vertex, warp-slices, largest unique-neighbour count, halo mode,
 * boundary stores per sweep, dependency levels of the device-built colouring (0: host colouring), kernel variant] */
int spgx_group_kernel_info(spgx_group *grp, int64_t *out8);

/* structure queries for the bit-exactness gates */
int spgx_group_num_colors(spgx_group *grp, int energyId /* ignored for VBD */);
int spgx_get_vertex_colors(spgx_group *grp, int32_t *colors /* nVerts */);
/* VBD incidence table in the reference's order (vbd.cpp:231-268): for vertex v and energy e the entries
 * [rowptr[v * nE + e], rowptr[v * nE + e + 1]) of (stencilIdx, localIdx) pairs. Pass entries == NULL to size. */
int spgx_get_vertex_incidence(spgx_group *grp, int64_t *rowptr, int32_t *entries, int64_t *nEntries);
int spgx_get_stencil_colors(spgx_group *grp, int energyId, int32_t *colors /* nStencils */);

/* ---- state --------------------------------------------------------------------------------- */
/* particleGroups()[i].positions() / velocities(): device -> host (either pointer may be NULL).  Waits for the steps
 * enqueued so far; a halo exchange of theirs that timed out is reported here (SPGX_ERR_COMM), as it is by
 * spgx_ctx_sync and by the next step call, so that steps enqueued with sync = 0 cannot hand back stale halo data. */
int spgx_download(spgx_group *grp, double *x, double *v);
/* The same two transfers without blocking the host, for a frame loop whose I/O overlaps the solve (the reference's loop reads
 * positions() every frame for display, apps/solverCompare.cpp:878, and writes state on user edits, :1388, :1410-1415): the
 * upload's host-to-device copy runs on its own stream into a staging pair while earlier steps still run, and is scattered
 * into place behind them; the download gathers behind the steps enqueued so far and copies out on a third stream while
 * later steps run.  x / v must be page-locked and stay untouched until spgx_transfer_wait or spgx_ctx_sync returns. */
int spgx_upload_async(spgx_group *grp, const double *x, const double *v);
int spgx_download_async(spgx_group *grp, double *x, double *v);
int spgx_transfer_wait(spgx_group *grp);
/* host -> device (callers may overwrite positions / velocities between steps, solverCompare.cpp:1388) */
int spgx_upload(spgx_group *grp, const double *x, const double *v);
/* ParticleGroup::reset (particleGroup.cpp:97-103): x <- xInitial, v <- 0 */
int spgx_reset(spgx_group *grp);
/* ParticleGroup::scaleMasses (particleGroup.cpp:105-113) */
int spgx_scale_masses(spgx_group *grp, double factor);

/* ---- solvers ------------------------------------------------------------------------------- */
/* VBD::step (vbd.cpp:12-195): nFrames calls of step(), each `nsubsteps` substeps of dt/nsubsteps with
 * `iterations` coloured sweeps. gravity: 3 doubles. Asynchronous on the context stream unless `sync`. */
int spgx_vbd_step(spgx_group *grp, double dt, int nsubsteps, int iterations, const double *gravity, int nFrames,
                  int sync);
/* XPBD::step (xpbd.cpp:16-131). iterations == 1 is the reference ("small steps": one projection pass per
 * substep, simplified numerator, energy.h:142-145); iterations > 1 keeps per-constraint lambda on device,
 * reset every substep. */
int spgx_xpbd_step(spgx_group *grp, double dt, int nsubsteps, int iterations, const double *gravity, int nFrames,
                   int sync);
int spgx_last_timings(spgx_group *grp, spgx_timings *out);
/* 0: plain launches; 1 (default): the launch sequence of one step() is captured in a CUDA graph and replayed */
int spgx_set_graph_mode(spgx_group *grp, int enabled);

/* ---- implicit-Euler family and symplectic Euler (group finalized with SPGX_NEWTON) ------------ */
/* ImplicitEulerBaraffWitkin / NewtonDx / NewtonDv ::step(): per substep the system force (getSystemForce,
 * implicitEulerBase.cpp:102-143), the energy Hessian (getSystemStiffnessMatrix, :176-221; K = -Hessian) and one
 * preconditioned conjugate-gradient solve (solveLinearSystem, :223-238: Eigen::ConjugateGradient, Jacobi preconditioner,
 * zero initial guess, stop at |r| < cgTolerance |b|), all on the device.  The reference fixes cgTolerance = 1e-5 and
 * cgMaxIterations = 300 (implicitEulerBase.h:22-23); pass those for parity.  The call synchronises with the device
 * (the host polls the convergence flag between batches of iterations); `sync` is accepted for symmetry. */
int spgx_implicit_step(spgx_group *grp, int implicitKind, double dt, int nsubsteps, const double *gravity,
                       double cgTolerance, int cgMaxIterations, int nFrames, int sync);
/* SimplecticEuler::step (simplecticEuler.cpp:12-62) with updateExplicitVelocities (energy.h:154-194) */
int spgx_symplectic_step(spgx_group *grp, double dt, int nsubsteps, const double *gravity, int nFrames, int sync);
/* CG iteration counts of the linear solves of the last spgx_implicit_step / spgx_linear_solve call (cg.iterations(),
 * implicitEulerBase.cpp:235); *n = how many there are, at most `capacity` are written */
int spgx_last_cg_iterations(spgx_group *grp, int32_t *out, int capacity, int *n);
/* getSystemForce on the current device state: f = gravity * mass - dE/dx, 3 doubles per vertex */
int spgx_system_force(spgx_group *grp, const double *gravity, double *f);
/* getSystemStiffnessMatrix on the current device state as block-sparse rows (3 x 3 blocks, row-major, columns
 * ascending within a row): K = -Hessian, with the duplicate triplets the reference hands to setFromTriplets already
 * summed.  rowptr: nVerts + 1, cols: *nBlocks, vals: 9 * *nBlocks; pass NULL arrays to query *nBlocks. */
int spgx_system_stiffness(spgx_group *grp, int64_t *nBlocks, int32_t *rowptr, int32_t *cols, double *vals);
/* energyHessian of one stencil (energy.h:354-362) on the current device state: (3S)^2 doubles, row-major */
int spgx_element_hessian(spgx_group *grp, int energyId, int stencil, double *hess);
/* solveLinearSystem for A = cM * M + cH * Hessian, the Hessian being the one assembled by the last
 * spgx_system_stiffness / spgx_implicit_step call: b, x 3 doubles per vertex */
int spgx_linear_solve(spgx_group *grp, double cM, double cH, const double *b, double *x, double cgTolerance,
                      int cgMaxIterations, int *iterations);

/* ---- per-element evaluators (tests; reference: energy.h:240-258, :117-153, :289) -------------- */
/* accumulateVertexForce + accumulateVertexHessian of (stencil, local) on the CURRENT device positions:
 * force3 = -dE/dx_v, hess9 = d2E/dx_v2 row-major. handDerived = 0: the generic dual-number evaluator;
 * 1: the kind's hand-derived override when it has one (what the VBD kernel runs), else the generic one. */
int spgx_eval_vertex_force_hessian(spgx_group *grp, int energyId, int stencil, int local, int handDerived,
                                   double *force3, double *hess9);
/* constraintsAndGradient: C (nC) and gradient (3S x nC row-major) on the current device positions */
int spgx_eval_constraints(spgx_group *grp, int energyId, int stencil, double *C, double *grad);

/* ---- rigid bodies ----------------------------------------------------------------------
 * Replaces RigidBodyGroup (spg/include/spg/sim/simObject/rigidBodyGroup.h:14-111, spg/source/sim/simObject/rigidBodyGroup.cpp)
 * under VBD::step and XPBD::step (the RigidBodyGroup branches of vbd.cpp:63-72, :104-147 and xpbd.cpp:60-70, projectPosition
 * with the 6 x 6 inverse-mass block, energy.h:117-153).  A body has 6 DOFs: centre of mass and an orientation kept as a
 * rotation matrix that is composed incrementally (rigidBodyGroup.cpp:64-78), with the axis-angle vector theta re-derived from
 * it.  Energy kinds and their addStencil parameters (modelParams, per stencil):
 *   SPGX_RB_SPRING                 rigidBodySpringEnergy.cpp:36             {local0.xyz, local1.xyz, restLength, stiffness}
 *   SPGX_RB_SPRING_SQUARED         rigidBodySpringSquaredEnergy.cpp:35      {local0.xyz, local1.xyz, restLength, stiffness}
 *   SPGX_RB_SPRING_ANCHOR          rigidBodySpringAnchorEnergy.cpp:68       {local.xyz, anchor.xyz, stiffness}
 *   SPGX_RB_ORIENTATION_ALIGNMENT  rigidBodyOrientationAlignmentEnergy.cpp:30 {stiffness}
 */
typedef struct spgx_rb_group spgx_rb_group;
enum spgx_rb_kind { SPGX_RB_SPRING = 0, SPGX_RB_SPRING_SQUARED = 1, SPGX_RB_SPRING_ANCHOR = 2, SPGX_RB_ORIENTATION_ALIGNMENT = 3 };
int spgx_rb_kind_stencil_size(int kind);
int spgx_rb_kind_num_model_params(int kind);
/* RigidBodyGroup::addBody for nBodies bodies (rigidBodyGroup.cpp:138-160): x, theta 3 doubles per body (theta: axis * angle),
 * localInertia 9 doubles per body (row-major, body frame). Velocities start at zero. */
int spgx_rb_group_create(spgx_ctx *ctx, int nBodies, const double *x, const double *theta, const double *mass,
                         const double *localInertia, spgx_rb_group **out);
void spgx_rb_group_destroy(spgx_rb_group *grp);
/* RigidBodyGroup::addEnergy / removeEnergy (rigidBodyGroup.cpp:162-178); same validation messages as the energies' addStencil */
int spgx_rb_group_add_energy(spgx_rb_group *grp, int kind, int nStencils, const int32_t *stencils, const double *modelParams,
                             int *energyId);
int spgx_rb_group_remove_energy(spgx_rb_group *grp, int energyId);
/* RBSpringAnchorEnergy::updateAnchor (rigidBodySpringAnchorEnergy.h:17) */
int spgx_rb_energy_update_anchor(spgx_rb_group *grp, int energyId, int stencil, const double *anchor3);
/* colouring over bodies (SPGX_VBD, vbd.cpp:203-268) or over stencils per energy (SPGX_XPBD, xpbd.cpp:133-160;
 * SPGX_XPBD_SERIAL: index order), upload. May be called again after add / remove: the state is kept. */
int spgx_rb_group_finalize(spgx_rb_group *grp, int solverKind);
/* nFrames x step() of the solver the group was finalized for; iterations: VBD sweeps per substep (XPBD: passes per substep,
 * the reference runs 1). gravity: 3 doubles or NULL for (0, -9.8, 0). Blocking. */
int spgx_rb_step(spgx_rb_group *grp, double dt, int nsubsteps, int iterations, const double *gravity, int nFrames);
/* any pointer may be NULL. rotation: 9 doubles per body, row-major */
int spgx_rb_download(spgx_rb_group *grp, double *x, double *theta, double *v, double *omega, double *rotation);
/* setPositions / setVelocities (rigidBodyGroup.cpp:43-63): rotation matrices and inertias are rebuilt from theta */
int spgx_rb_upload(spgx_rb_group *grp, const double *x, const double *theta, const double *v, const double *omega);
int spgx_rb_reset(spgx_rb_group *grp);
int spgx_rb_get_body_colors(spgx_rb_group *grp, int32_t *colors);
int spgx_rb_get_stencil_colors(spgx_rb_group *grp, int energyId, int32_t *colors);
/* per-element parity hooks: constraintsAndGradient (C: 1, grad: 6 S) and accumulateVertexForce / Hessian (6, 6 x 6) */
int spgx_rb_eval_constraint(spgx_rb_group *grp, int energyId, int stencil, double *C, double *grad);
int spgx_rb_eval_vertex(spgx_rb_group *grp, int energyId, int stencil, int local, double *force6, double *hess36);

/* Debug aid (what Verbosity::Debug of the facade runs after every step; the reference has no counterpart, SURVEY.md
 * section 5): counts the non-finite components of the positions and velocities this rank owns, on the device.
 * out3 = [non-finite position components, non-finite velocity components, first offending vertex id or -1]. */
int spgx_check_finite(spgx_group *grp, int64_t *out3);

/* ---- algorithmic byte model ---------------------------------- */
/* bytes one VBD iteration (all colours) / one XPBD substep must move under the gather model */
int spgx_model_bytes(spgx_group *grp, double *bytesPerIteration, double *bytesLowerBound);

#ifdef __cplusplus
}
#endif
#endif /* SPGX_H */
