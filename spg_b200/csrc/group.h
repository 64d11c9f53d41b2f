// Internal state behind the opaque C-ABI handles.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/spgx.h"
#include "kernels.h"

namespace spgx
{
struct CudaError : std::runtime_error {
    using std::runtime_error::runtime_error;
};
struct StateError : std::runtime_error {
    using std::runtime_error::runtime_error;
};
struct CommError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

#define SPGX_CUDA(call)                                                                                         \
    do {                                                                                                        \
        cudaError_t err__ = (call);                                                                             \
        if (err__ != cudaSuccess)                                                                               \
            throw spgx::CudaError(std::string(#call) + ": " + cudaGetErrorString(err__));                       \
    } while (0)

void setLastError(const std::string &msg);

// run f, translate exceptions into status codes + thread-local message
template <class F>
int guarded(F &&f)
{
    try {
        f();
        return SPGX_OK;
    } catch (const CudaError &e) {
        setLastError(e.what());
        return SPGX_ERR_CUDA;
    } catch (const StateError &e) {
        setLastError(e.what());
        return SPGX_ERR_STATE;
    } catch (const CommError &e) {
        setLastError(e.what());
        return SPGX_ERR_COMM;
    } catch (const std::exception &e) {
        setLastError(e.what());
        return SPGX_ERR_INVALID;
    } catch (...) {
        setLastError("unknown exception");
        return SPGX_ERR_INVALID;
    }
}

template <class T>
struct DevBuf {
    T *ptr = nullptr;
    size_t count = 0;
    bool owned = true;  // false: a view into another allocation (see view())
    DevBuf() = default;
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    DevBuf(DevBuf &&o) noexcept : ptr(o.ptr), count(o.count), owned(o.owned)
    {
        o.ptr = nullptr;
        o.count = 0;
        o.owned = true;
    }
    DevBuf &operator=(DevBuf &&o) noexcept
    {
        if (this != &o) {
            release();
            ptr = o.ptr;
            count = o.count;
            owned = o.owned;
            o.ptr = nullptr;
            o.count = 0;
            o.owned = true;
        }
        return *this;
    }
    ~DevBuf() { release(); }
    void release()
    {
        if (ptr && owned) cudaFree(ptr);
        ptr = nullptr;
        count = 0;
        owned = true;
    }
    // n elements of someone else's allocation starting at p (not freed here)
    void view(T *p, size_t n)
    {
        release();
        ptr = p;
        count = n;
        owned = false;
    }
    void alloc(size_t n)
    {
        release();
        if (n == 0) return;
        SPGX_CUDA(cudaMalloc(&ptr, n * sizeof(T)));
        count = n;
    }
    void upload(const std::vector<T> &h, cudaStream_t s)
    {
        if (!(ptr && !owned && count == h.size())) alloc(h.size());  // a view of the right size is written in place
        if (!h.empty()) SPGX_CUDA(cudaMemcpyAsync(ptr, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, s));
    }
};
}  // namespace spgx

struct spgx_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t h2dStream = nullptr, d2hStream = nullptr;  // spgx_upload_async / spgx_download_async (created on first use)
    cudaDeviceProp prop{};
    void *comm = nullptr;  // ncclComm_t (comm.cu)
    int rank = 0, nRanks = 1;
    std::vector<spgx_group *> groups;  // live groups of this context (spgx_ctx_sync reports their halo-exchange timeouts)
};

namespace spgx
{
// comm.cu
void commUniqueId(void *id128);
void commInit(spgx_ctx *ctx, int rank, int nRanks, const void *id128);
void commDestroy(spgx_ctx *ctx);
}  // namespace spgx

namespace spgx
{
struct EnergyState {
    int kind = 0, S = 0, NP = 0, NC = 0, n = 0;
    std::vector<int32_t> stencils;  // original vertex ids, reference order
    std::vector<double> records;    // NP per stencil, reference order
    // --- device copies (stencils in slot ids; XPBD: both reordered by colour) ---
    DevBuf<int32_t> d_stencils;
    DevBuf<double> d_records;
    // VBD
    DevBuf<uint32_t> d_rowptr, d_entries;
    bool streamed = false;  // all of its incidences live in streams (kernels.h)
    int64_t ownedIncidences = 0;  // incidences of vertices this rank owns
    // XPBD
    std::vector<int32_t> stencilColors;  // per stencil, reference order
    std::vector<int32_t> order;          // position -> stencil id
    std::vector<int32_t> positionOf;     // stencil id -> position
    std::vector<int64_t> colorOffsets;   // nColors + 1
    DevBuf<double> d_lambda;
    DevBuf<double> d_stencilW;  // XPBD: inverse masses per stencil vertex, in the order of d_stencils
};
}  // namespace spgx

namespace spgx
{
void p2pRelease(spgx_group *grp);
}

struct spgx_group {
    spgx_ctx *ctx = nullptr;
    int nVerts = 0;
    std::vector<double> x0, xInitial, mass;  // host, reference vertex order
    std::vector<double> xPending, vPending;  // state held on host until finalize uploads it
    std::vector<spgx::EnergyState> energies;

    int solverKind = -1;
    bool finalized = false;
    // domain decomposition (partition.h): this rank's share of the group
    bool partitioned = false;
    int rank = 0, nRanks = 1;
    std::vector<int32_t> owner;  // per global vertex
    int nOwned = 0, nLocal = 0;  // local slots: [0, nOwned) owned, [nOwned, nLocal) halo
    int localLo = 0, localHi = 0, ownedLo = 0, ownedHi = 0;  // global index ranges spanned
    bool ownedContiguous = true;
    std::vector<int> peers;
    std::vector<int64_t> sendOffsets, recvSlot, recvCount;  // (colour, peer)-major, partition.h
    spgx::DevBuf<int32_t> d_sendSlots;
    spgx::DevBuf<double> d_sendBuf;
    int64_t haloBytesPerSweep = 0;
    bool commWarm = false;
    // peer-memory halo exchange fused into the colour pass (comm.cu, halo.cuh)
    struct PushEntry {
        int32_t slot, peer, colour, pos;  // owned slot, index into `peers`, its colour, position in the (colour, peer) send list
    };
    struct P2PState {
        bool connected = false;
        int mode = 0;                       // SPGX_HALO_*: what the next step() uses
        std::vector<int32_t> sendSlots;     // host copy of d_sendSlots
        std::vector<PushEntry> pushLocal;   // in slice order
        std::vector<int> pushWarps;         // per colour: slices that hold push entries
        spgx::DevBuf<uint32_t> d_sync;      // [0, kMaxPeers) flags raised by the neighbours, [kMaxPeers] epoch base, [+1] push counter
        spgx::DevBuf<uint2> d_push;
        uint32_t *hostError = nullptr, *devError = nullptr;  // mapped pinned word: a wait ran into its timeout
        std::vector<void *> mapped;         // cudaIpcOpenMemHandle results to close
        spgx::VbdHaloDev dev{};
    } p2p;
    // domain-decomposed XPBD (xpbd_partition.h): one exchange per (energy, colour) pass.  Pass ids are energy-major,
    // colour-minor over ALL global colours of every energy (a pass this rank holds no stencil of still owes its event).
    struct XpbdHaloState {
        int nPasses = 0;
        std::vector<int32_t> halo;          // global ids of the halo vertices, ascending (slot nOwned + i)
        std::vector<int64_t> sendOffsets;   // nPasses * nPeers + 1, (pass, peer)-major, into sendSlots
        std::vector<int32_t> sendSlots;     // owned local slots
        std::vector<int32_t> sendVerts;     // the same entries as global vertex ids (resolved against the peers' halo lists)
        std::vector<int64_t> recvOffsets;   // same shape, into recvSlots
        std::vector<int32_t> recvSlots;     // halo local slots, in the sender's order
        std::vector<int64_t> pushOffsets;   // nPasses + 1 into d_push (peer-memory exchange)
        spgx::DevBuf<int32_t> d_recvSlots;
        spgx::DevBuf<double> d_recvBuf;
    } xh;
    bool identityOrder = true;  // device slot == vertex id (every solver but VBD and partitioned XPBD)
    // vertex colouring / ordering (VBD); identity ordering for XPBD
    std::vector<int32_t> vertexColors;
    int colouringRounds = 0;  // > 0: the colouring was built on the device in that many dependency levels (colour.cu)
    std::vector<int32_t> slot2vert, vert2slot;
    std::vector<int64_t> colorOffsets;
    int vbdLanes = 4;                   // threads per vertex in the VBD colour pass
    int vbdOcc = 4;                     // resident-blocks-per-SM hint of the streaming kernel
    int vbdNbrFixed = 1;                // neighbour table at a fixed stride per slice (kernels.h: VbdArgs::nbrFixed)
    int vbdWedge = 1;                   // membrane + bending incidences of one fan triangle joined into one record (streams.cuh: ST_WEDGE)
    int64_t wedgeRecords = 0;           // wedge records of the owned vertices (kernel_info)
    // streaming layout (kernels.h): unique-neighbour tables + one stream per (energy, stream type)
    struct StreamState {
        int type = 0;
        spgx::DevBuf<uint32_t> d_idx;
        spgx::DevBuf<double> d_params;
    };
    std::vector<StreamState> streams;  // indexed by stream type when non-empty
    spgx::DevBuf<uint32_t> d_sliceDesc, d_nbrIds;
    std::vector<uint32_t> sliceDescHost;  // host copy of d_sliceDesc (spgx_energy_update_record patches stream records in place)
    int maxNbr = 0, nSlices = 0, vbdLookahead = 512, vbdPdl = 1, vbdCarveout = -1;
    int64_t vbdL2Persist = 0;  // bytes of L2 set aside for the persisting window over the positions (0: off)
    unsigned streamMask = 0;
    double streamBytes = 0;  // bytes of stream + neighbour tables one full sweep reads (reported)
    std::vector<int32_t> sliceBase;     // per colour: index of its first warp-slice (32 / vbdLanes vertices each)
    double uniqueNeighbourSum = 0;  // sum over vertices of the number of distinct stencil neighbours

    // implicit-Euler family (api_newton.cu, newton.cu): sparse pattern, element Hessians, CG work vectors
    struct NewtonState {
        std::vector<uint32_t> bsrRowptr;  // nVerts + 1 (host copies serve spgx_system_stiffness)
        std::vector<int32_t> bsrCol;
        int64_t nnzb = 0;
        spgx::DevBuf<uint32_t> d_bsrRowptr, d_diagPos;
        spgx::DevBuf<int32_t> d_bsrCol, d_bsrRow;
        spgx::DevBuf<double> d_bsrH;
        spgx::DevBuf<double> d_elemArena;   // the element Hessians of all energies, one allocation
        std::vector<double *> d_elemH;      // per energy: its range of the arena
        spgx::DevBuf<uint32_t> d_asmPtr;    // nnzb + 1: the sources of every block of the sparse Hessian ...
        spgx::DevBuf<uint32_t> d_asmSrc;    // ... as 3x3-block indices into the arena, in summation order
        spgx::DevBuf<double> d_f, d_rhs, d_sol, d_r, d_p, d_z, d_tmp, d_invdiag, d_x0, d_v0;
        spgx::DevBuf<double> d_xi, d_scalar;  // adaptive solvers: the accepted Newton iterate (x or v), a norm
        double *hostScalar = nullptr;         // pinned
        double currentDt = -1.0;              // implicitEulerNewtonRobust.h:21-22 / quasiStaticNewtonRobust.h:24-25
        int consecutiveSuccessfulSteps = 0;
        spgx::DevBuf<double> d_xm1, d_vm1;   // BDF2: the state before the previous substep (bdf2.h:19-20)
        bool bdf2Init = false;
        spgx::DevBuf<double> d_partials;
        spgx::DevBuf<spgx::CgState> d_state;
        spgx::CgState *hostState = nullptr;  // pinned
        bool assembled = false;              // d_bsrH holds the Hessian of some state
        std::vector<int32_t> cgIterations;   // of every linear solve of the last step call
        double elemHessianBytes = 0;
        cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};  // substep start, system assembled, linear system solved
    } newton;
    // positions, inertial positions and masses of a VBD group are views into ONE allocation, so that a single L2
    // access-policy window covers everything the colour pass reads per vertex besides the incidence streams
    spgx::DevBuf<double> d_stateArena;
    spgx::DevBuf<double> d_x, d_v, d_xOld, d_xInertial, d_mass, d_invMass, d_stage, d_g;
    spgx::DevBuf<int32_t> d_slot2vert;
    spgx::DevBuf<double> d_eval;
    // asynchronous state transfer (api_step.cu): staging in vertex order, one pair per direction, so that the upload of the
    // next frame's inputs and the download of the previous frame's result overlap the step that is running
    struct AsyncXfer {
        spgx::DevBuf<double> d_inX, d_inV, d_outX, d_outV;
        cudaEvent_t uploaded = nullptr, inFree = nullptr, outReady = nullptr, downloaded = nullptr;
        bool inUsed = false, outUsed = false;
    } xfer;

    // CUDA graph of one step() (keyed by the step parameters)
    bool graphMode = true;
    cudaGraphExec_t graphExec = nullptr;
    struct GraphKey {
        int solver = -1, nsub = 0, iters = 0;
        double dt = 0, g[3] = {0, 0, 0};
        bool operator==(const GraphKey &o) const
        {
            return solver == o.solver && nsub == o.nsub && iters == o.iters && dt == o.dt && g[0] == o.g[0] &&
                   g[1] == o.g[1] && g[2] == o.g[2];
        }
    } graphKey;
    int64_t graphLaunches = 0;

    spgx_timings timings{};
    cudaEvent_t evStart = nullptr, evStop = nullptr;
    int64_t launchCounter = 0;
    // per-phase events (plain-launch mode only; a captured graph is timed as a whole)
    std::vector<cudaEvent_t> phaseEvents;
    size_t phaseCursor = 0;
    void markPhase(cudaStream_t st)
    {
        if (phaseCursor == phaseEvents.size()) {
            cudaEvent_t ev = nullptr;
            if (cudaEventCreate(&ev) != cudaSuccess) throw spgx::CudaError("cudaEventCreate failed");
            phaseEvents.push_back(ev);
        }
        if (cudaEventRecord(phaseEvents[phaseCursor++], st) != cudaSuccess) throw spgx::CudaError("cudaEventRecord failed");
    }

    ~spgx_group()
    {
        for (auto ev : phaseEvents) cudaEventDestroy(ev);
        if (graphExec) cudaGraphExecDestroy(graphExec);
        if (evStart) cudaEventDestroy(evStart);
        if (evStop) cudaEventDestroy(evStop);
        if (newton.hostState) cudaFreeHost(newton.hostState);
        if (newton.hostScalar) cudaFreeHost(newton.hostScalar);
        for (auto e : newton.ev)
            if (e) cudaEventDestroy(e);
        for (auto e : {xfer.uploaded, xfer.inFree, xfer.outReady, xfer.downloaded})
            if (e) cudaEventDestroy(e);
        spgx::p2pRelease(this);
    }
};

namespace spgx
{
// a wait of the peer-memory halo exchange ran into its timeout (halo.cuh): report it once, as SPGX_ERR_COMM
inline void throwIfHaloTimedOut(spgx_group *grp)
{
    if (grp && grp->p2p.hostError && *grp->p2p.hostError) {
        *grp->p2p.hostError = 0;
        throw CommError("halo exchange: a neighbouring rank did not raise its event counter within the timeout "
                        "(SPGX_HALO_TIMEOUT_S): ranks out of step, or a rank died");
    }
}
int enqueueHaloExchange(spgx_group *grp, int colour, cudaStream_t st);  // comm.cu
}  // namespace spgx
#include "host_build.h"
namespace spgx
{
// colour.cu: colorVertices (graphColoring.cpp:72-84) as a level-by-level evaluation of its dependency DAG on the device
std::vector<int32_t> colorVerticesDevice(spgx_ctx *ctx, int nVerts, const std::vector<FlatStencilsView> &sets, int *rounds);
int enqueueXpbdHaloExchange(spgx_group *grp, int pass, cudaStream_t st);  // comm.cu
// api_newton.cu: incidence tables, sparse pattern and work buffers of a group finalized with SPGX_NEWTON
void newtonBuild(spgx_group *grp, cudaStream_t st);
// peer-memory exchange (comm.cu)
void p2pPrepare(spgx_group *grp, cudaStream_t st);   // after finalize: sync words, error word; closes old mappings
void p2pRelease(spgx_group *grp);
void p2pDisconnect(spgx_group *grp);                 // unmap the neighbours' arrays, keep this rank's own buffers
int64_t p2pExport(spgx_group *grp, void *blob, int64_t capacity);
void p2pConnect(spgx_group *grp, int nBlobs, const void *const *blobs, const int64_t *sizes);
}
