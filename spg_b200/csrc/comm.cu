// Multi-GPU plumbing of the VBD sweep: one process per GPU, NCCL point-to-point over NVLink for the per-colour halo
// exchange (SURVEY.md section 8e / north_star "per-colour boundary-vertex halo exchange over NVLink via NCCL").
//
// NCCL is bound at run time (dlopen of libnccl.so.2: the copy the host process already loaded -- e.g. the one bundled
// with torch when the ranks were launched by torchrun -- else the system one), so that libspgx.so itself has no link
// dependency on it and single-GPU users never touch it.  Only the stable point-to-point subset is used.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "group.h"
#include "halo.cuh"

namespace spgx
{
namespace
{
struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};

NcclApi &nccl()
{
    static NcclApi api;
    if (api.lib) return api;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) throw CommError(std::string("cannot load NCCL (libnccl.so.2): ") + dlerror());
    auto sym = [&](const char *name) {
        void *p = dlsym(h, name);
        if (!p) throw CommError(std::string("NCCL symbol missing: ") + name);
        return p;
    };
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
    api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
    api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
    api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
    api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
    api.lib = h;
    return api;
}

void ncclCheck(ncclResult_t r, const char *what)
{
    if (r != ncclSuccess) throw CommError(std::string(what) + ": " + nccl().GetErrorString(r));
}

// sendBuf[i] = x[slots[i]] over 3-vectors
__global__ void packKernel(const double *__restrict__ x, const int32_t *__restrict__ slots, double *__restrict__ out, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t o = 3 * static_cast<size_t>(slots[i]);
    out[3 * static_cast<size_t>(i)] = x[o];
    out[3 * static_cast<size_t>(i) + 1] = x[o + 1];
    out[3 * static_cast<size_t>(i) + 2] = x[o + 2];
}
// peer-memory exchange outside the colour passes: one thread raises an event / waits for one and advances the epoch
__global__ void haloSignalKernel(const VbdHaloDev h, int event)
{
    __threadfence_system();
    haloRaise(h, haloEpochBase(h) + static_cast<uint32_t>(event) + 1u);
}
__global__ void haloWaitAdvanceKernel(const VbdHaloDev h, int event, uint32_t *epochBase)
{
    const uint32_t value = haloEpochBase(h) + static_cast<uint32_t>(event) + 1u;
    haloWaitAll(h, value);
    *reinterpret_cast<volatile uint32_t *>(epochBase) = value;
}

// XPBD (xpbd_partition.h): after an (energy, colour) pass the owned vertices it moved go into the halo slots of the
// neighbours that keep them without holding the stencil that moved them; the last block to finish its stores raises the
// pass's event on every neighbour (same producer pattern as haloPushSlice, vbd_stream.cuh).  A pass with nothing to send is
// one block that only raises the event.
__global__ void __launch_bounds__(128) xpbdHaloPushKernel(const VbdHaloDev h, const double *__restrict__ x, int begin, int end, int event)
{
    const int i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i < end) {
        const uint2 e = __ldg(h.push + i);
        const double *src = x + 3 * static_cast<size_t>(e.x);
        double *dst = h.peerX[e.y >> 28] + 3 * static_cast<size_t>(e.y & 0x0FFFFFFFu);
        dst[0] = src[0];
        dst[1] = src[1];
        dst[2] = src[2];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t done = atomicAdd(h.pushDone, 1u) + 1u;
        if (done == gridDim.x) {
            *reinterpret_cast<volatile uint32_t *>(h.pushDone) = 0u;
            __threadfence_system();
            haloRaise(h, haloEpochBase(h) + static_cast<uint32_t>(event) + 1u);
        }
    }
}
// x[slots[i]] = buf[i] over 3-vectors (what ncclRecv delivered, in the sender's order)
__global__ void unpackKernel(double *__restrict__ x, const int32_t *__restrict__ slots, const double *__restrict__ buf, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t o = 3 * static_cast<size_t>(slots[i]);
    x[o] = buf[3 * static_cast<size_t>(i)];
    x[o + 1] = buf[3 * static_cast<size_t>(i) + 1];
    x[o + 2] = buf[3 * static_cast<size_t>(i) + 2];
}

constexpr uint32_t kBlobMagic = 0x53504758u;  // "SPGX"
constexpr uint32_t kBlobMagicXpbd = 0x53504759u;
// XPBD blob: header, int32 peers[nPeers] (padded to 8 bytes), int32 halo[nColors] (nColors field = number of halo vertices;
// global ids ascending: halo vertex i sits in slot nLocal - nHalo + i)

struct BlobHeader {
    uint32_t magic;
    int32_t rank, nRanks, nPeers, nColors, reserved;
    int64_t nLocal;
    cudaIpcMemHandle_t x, sync;
};
// followed by: int32 peers[nPeers] (padded to 8 bytes); int64 recvSlot[nColors * nPeers]; int64 recvCount[nColors * nPeers]
size_t paddedPeersBytes(int nPeers)
{
    return (static_cast<size_t>(nPeers) * 4 + 7) / 8 * 8;
}
size_t blobBytes(int nPeers, int nColors)
{
    return sizeof(BlobHeader) + paddedPeersBytes(nPeers) + 2 * sizeof(int64_t) * static_cast<size_t>(nColors) * nPeers;
}
size_t xpbdBlobBytes(int nPeers, int nHalo)
{
    return sizeof(BlobHeader) + paddedPeersBytes(nPeers) + (static_cast<size_t>(nHalo) * 4 + 7) / 8 * 8;
}
}  // namespace

void launchHaloSignal(const VbdHaloDev &h, int event, cudaStream_t stream)
{
    haloSignalKernel<<<1, 1, 0, stream>>>(h, event);
}
void launchHaloWaitAdvance(const VbdHaloDev &h, int event, uint32_t *epochBase, cudaStream_t stream)
{
    haloWaitAdvanceKernel<<<1, 1, 0, stream>>>(h, event, epochBase);
}

void launchXpbdHaloPush(const VbdHaloDev &h, const double *x, int begin, int end, int event, cudaStream_t stream)
{
    const int n = end - begin;
    xpbdHaloPushKernel<<<std::max(1, (n + 127) / 128), 128, 0, stream>>>(h, x, begin, end, event);
}

void p2pDisconnect(spgx_group *grp)
{
    auto &p = grp->p2p;
    for (void *m : p.mapped) cudaIpcCloseMemHandle(m);
    p.mapped.clear();
    if (p.connected && grp->graphExec) {  // the captured launches hold the neighbours' pointers
        cudaGraphExecDestroy(grp->graphExec);
        grp->graphExec = nullptr;
    }
    p.connected = false;
    if (p.mode == SPGX_HALO_P2P) p.mode = grp->partitioned ? SPGX_HALO_NCCL : SPGX_HALO_NONE;
    p.dev = VbdHaloDev{};
}

void p2pRelease(spgx_group *grp)
{
    p2pDisconnect(grp);
    auto &p = grp->p2p;
    if (p.hostError) cudaFreeHost(p.hostError);
    p.hostError = p.devError = nullptr;
    p.d_sync.release();
    p.d_push.release();
}

// after finalize (the position array has just been allocated): fresh sync words, fresh error word
void p2pPrepare(spgx_group *grp, cudaStream_t st)
{
    p2pRelease(grp);
    auto &p = grp->p2p;
    p.mode = grp->partitioned ? SPGX_HALO_NCCL : SPGX_HALO_NONE;
    if (!grp->partitioned || grp->peers.empty()) return;
    if (static_cast<int>(grp->peers.size()) > kMaxPeers) return;  // NCCL exchange only
    p.d_sync.alloc(kMaxPeers + 8);
    SPGX_CUDA(cudaMemsetAsync(p.d_sync.ptr, 0, p.d_sync.count * sizeof(uint32_t), st));
    SPGX_CUDA(cudaHostAlloc(reinterpret_cast<void **>(&p.hostError), sizeof(uint32_t), cudaHostAllocMapped));
    *p.hostError = 0;
    SPGX_CUDA(cudaHostGetDevicePointer(reinterpret_cast<void **>(&p.devError), p.hostError, 0));
    SPGX_CUDA(cudaStreamSynchronize(st));
}

namespace
{
int64_t p2pExportXpbd(spgx_group *grp, void *blob, int64_t capacity)
{
    const int nPeers = static_cast<int>(grp->peers.size());
    const int nHalo = static_cast<int>(grp->xh.halo.size());
    const size_t bytes = xpbdBlobBytes(nPeers, nHalo);
    if (!blob) return static_cast<int64_t>(bytes);
    if (capacity < static_cast<int64_t>(bytes)) throw std::runtime_error("p2p export: blob too small");
    std::memset(blob, 0, bytes);
    BlobHeader h{};
    h.magic = kBlobMagicXpbd;
    h.rank = grp->rank;
    h.nRanks = grp->nRanks;
    h.nPeers = nPeers;
    h.nColors = nHalo;
    h.reserved = grp->xh.nPasses;
    h.nLocal = grp->nLocal;
    if (nPeers > 0) {
        if (!grp->p2p.d_sync.ptr) throw StateError("p2p export: more neighbours than the peer-memory exchange supports");
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        SPGX_CUDA(cudaIpcGetMemHandle(&h.x, grp->d_x.ptr));
        SPGX_CUDA(cudaIpcGetMemHandle(&h.sync, grp->p2p.d_sync.ptr));
    }
    char *out = static_cast<char *>(blob);
    std::memcpy(out, &h, sizeof(h));
    out += sizeof(h);
    std::memcpy(out, grp->peers.data(), sizeof(int32_t) * nPeers);
    out += paddedPeersBytes(nPeers);
    std::memcpy(out, grp->xh.halo.data(), sizeof(int32_t) * nHalo);
    return static_cast<int64_t>(bytes);
}

// XPBD: the push list is pass-major; the slot of a sent vertex on the neighbour is its rank in the neighbour's halo list
void p2pConnectXpbd(spgx_group *grp, int nBlobs, const void *const *blobs, const int64_t *sizes)
{
    if (nBlobs != grp->nRanks) throw std::runtime_error("p2p connect: one blob per rank expected");
    auto &p = grp->p2p;
    auto &xh = grp->xh;
    const int nPeers = static_cast<int>(grp->peers.size());
    if (nPeers > kMaxPeers) throw StateError("p2p connect: more neighbours than the peer-memory exchange supports");
    SPGX_CUDA(cudaSetDevice(grp->ctx->device));
    for (void *m : p.mapped) cudaIpcCloseMemHandle(m);
    p.mapped.clear();
    p.connected = false;
    VbdHaloDev dev{};
    dev.nPeers = nPeers;
    std::vector<const int32_t *> peerHalo(nPeers);
    std::vector<int64_t> peerHaloBase(nPeers), peerNHalo(nPeers);
    for (int q = 0; q < nPeers; ++q) {
        const int r = grp->peers[q];
        const char *b = static_cast<const char *>(blobs[r]);
        BlobHeader h;
        if (!b || sizes[r] < static_cast<int64_t>(sizeof(h))) throw std::runtime_error("p2p connect: blob missing or truncated");
        std::memcpy(&h, b, sizeof(h));
        if (h.magic != kBlobMagicXpbd || h.rank != r || h.nRanks != grp->nRanks || h.reserved != xh.nPasses)
            throw std::runtime_error("p2p connect: blob does not match this group (solver / rank / ranks / passes)");
        if (sizes[r] < static_cast<int64_t>(xpbdBlobBytes(h.nPeers, h.nColors))) throw std::runtime_error("p2p connect: blob truncated");
        const int32_t *peersOfR = reinterpret_cast<const int32_t *>(b + sizeof(h));
        int idxAtPeer = -1;
        for (int i = 0; i < h.nPeers; ++i)
            if (peersOfR[i] == grp->rank) idxAtPeer = i;
        if (idxAtPeer < 0) throw std::runtime_error("p2p connect: neighbour does not list this rank as its neighbour");
        if (h.nLocal >= (1ll << 28)) throw std::runtime_error("p2p connect: neighbour holds too many slots for the packed push entry");
        peerHalo[q] = reinterpret_cast<const int32_t *>(b + sizeof(h) + paddedPeersBytes(h.nPeers));
        peerNHalo[q] = h.nColors;
        peerHaloBase[q] = h.nLocal - h.nColors;
        void *px = nullptr, *ps = nullptr;
        SPGX_CUDA(cudaIpcOpenMemHandle(&px, h.x, cudaIpcMemLazyEnablePeerAccess));
        p.mapped.push_back(px);
        SPGX_CUDA(cudaIpcOpenMemHandle(&ps, h.sync, cudaIpcMemLazyEnablePeerAccess));
        p.mapped.push_back(ps);
        dev.peerX[q] = static_cast<double *>(px);
        dev.peerFlag[q] = static_cast<uint32_t *>(ps) + idxAtPeer;
    }
    std::vector<uint2> push;
    xh.pushOffsets.assign(static_cast<size_t>(xh.nPasses) + 1, 0);
    for (int pass = 0; pass < xh.nPasses; ++pass) {
        for (int q = 0; q < nPeers; ++q) {
            const size_t cell = static_cast<size_t>(pass) * nPeers + q;
            for (int64_t i = xh.sendOffsets[cell]; i < xh.sendOffsets[cell + 1]; ++i) {
                const int32_t v = xh.sendVerts[i];
                const int32_t *lo = peerHalo[q], *hi = peerHalo[q] + peerNHalo[q];
                const int32_t *it = std::lower_bound(lo, hi, v);
                if (it == hi || *it != v) throw std::runtime_error("p2p connect: a sent vertex is not in the neighbour's halo");
                const int64_t remote = peerHaloBase[q] + (it - lo);
                push.push_back(make_uint2(static_cast<uint32_t>(xh.sendSlots[i]), (static_cast<uint32_t>(q) << 28) | static_cast<uint32_t>(remote)));
            }
        }
        xh.pushOffsets[pass + 1] = static_cast<int64_t>(push.size());
    }
    cudaStream_t st = grp->ctx->stream;
    p.d_push.alloc(std::max<size_t>(push.size(), 1));
    if (!push.empty())
        SPGX_CUDA(cudaMemcpyAsync(p.d_push.ptr, push.data(), push.size() * sizeof(uint2), cudaMemcpyHostToDevice, st));
    SPGX_CUDA(cudaStreamSynchronize(st));
    dev.myFlag = p.d_sync.ptr;
    dev.epochBase = p.d_sync.ptr + kMaxPeers;
    dev.pushDone = p.d_sync.ptr + kMaxPeers + 1;
    dev.push = p.d_push.ptr;
    dev.error = p.devError;
    double timeoutS = 10.0;
    if (const char *env = getenv("SPGX_HALO_TIMEOUT_S")) timeoutS = std::max(0.001, atof(env));
    dev.timeoutNs = static_cast<unsigned long long>(timeoutS * 1e9);
    p.dev = dev;
    p.connected = true;
    p.mode = SPGX_HALO_P2P;
    if (grp->graphExec) {
        cudaGraphExecDestroy(grp->graphExec);
        grp->graphExec = nullptr;
    }
}
}  // namespace

int64_t p2pExport(spgx_group *grp, void *blob, int64_t capacity)
{
    if (!grp->finalized || (grp->solverKind != SPGX_VBD && grp->solverKind != SPGX_XPBD))
        throw StateError("p2p export needs a group finalized for VBD or XPBD");
    if (!grp->partitioned) throw StateError("p2p export needs a partitioned group (spgx_group_set_partition)");
    if (grp->solverKind == SPGX_XPBD) return p2pExportXpbd(grp, blob, capacity);
    const int nPeers = static_cast<int>(grp->peers.size());
    const int nColors = static_cast<int>(grp->colorOffsets.size()) - 1;
    const size_t bytes = blobBytes(nPeers, nColors);
    if (!blob) return static_cast<int64_t>(bytes);
    if (capacity < static_cast<int64_t>(bytes)) throw std::runtime_error("p2p export: blob too small");
    std::memset(blob, 0, bytes);
    BlobHeader h{};
    h.magic = kBlobMagic;
    h.rank = grp->rank;
    h.nRanks = grp->nRanks;
    h.nPeers = nPeers;
    h.nColors = nColors;
    h.nLocal = grp->nLocal;
    if (nPeers > 0) {
        if (!grp->p2p.d_sync.ptr) throw StateError("p2p export: more neighbours than the peer-memory exchange supports");
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        SPGX_CUDA(cudaIpcGetMemHandle(&h.x, grp->d_x.ptr));
        SPGX_CUDA(cudaIpcGetMemHandle(&h.sync, grp->p2p.d_sync.ptr));
    }
    char *out = static_cast<char *>(blob);
    std::memcpy(out, &h, sizeof(h));
    out += sizeof(h);
    std::memcpy(out, grp->peers.data(), sizeof(int32_t) * nPeers);
    out += paddedPeersBytes(nPeers);
    std::memcpy(out, grp->recvSlot.data(), sizeof(int64_t) * grp->recvSlot.size());
    out += sizeof(int64_t) * static_cast<size_t>(nColors) * nPeers;
    std::memcpy(out, grp->recvCount.data(), sizeof(int64_t) * grp->recvCount.size());
    return static_cast<int64_t>(bytes);
}

// blobs[r] = what rank r exported.  Opens the neighbours' arrays and turns the local push list into device entries
// (local slot, neighbour << 28 | slot on the neighbour): what peer p receives from this rank after colour c lands in
// its halo range recvSlot[c, me], in the order of this rank's (c, p) send list (partition.h).
void p2pConnect(spgx_group *grp, int nBlobs, const void *const *blobs, const int64_t *sizes)
{
    if (grp->finalized && grp->solverKind == SPGX_XPBD && grp->partitioned) {
        p2pConnectXpbd(grp, nBlobs, blobs, sizes);
        return;
    }
    if (!grp->finalized || grp->solverKind != SPGX_VBD || !grp->partitioned)
        throw StateError("p2p connect needs a partitioned group finalized for VBD or XPBD");
    if (nBlobs != grp->nRanks) throw std::runtime_error("p2p connect: one blob per rank expected");
    if (grp->streamMask == 0) throw StateError("p2p connect: the group's energies have no streaming colour-pass kernel");
    auto &p = grp->p2p;
    const int nPeers = static_cast<int>(grp->peers.size());
    const int nColors = static_cast<int>(grp->colorOffsets.size()) - 1;
    if (nPeers > kMaxPeers) throw StateError("p2p connect: more neighbours than the peer-memory exchange supports");
    SPGX_CUDA(cudaSetDevice(grp->ctx->device));
    for (void *m : p.mapped) cudaIpcCloseMemHandle(m);
    p.mapped.clear();
    p.connected = false;
    VbdHaloDev dev{};
    dev.nPeers = nPeers;
    std::vector<const int64_t *> peerRecvSlot(nPeers), peerRecvCount(nPeers);
    std::vector<int> idxAtPeer(nPeers, -1), peerNPeers(nPeers, 0);
    for (int q = 0; q < nPeers; ++q) {
        const int r = grp->peers[q];
        const char *b = static_cast<const char *>(blobs[r]);
        BlobHeader h;
        if (!b || sizes[r] < static_cast<int64_t>(sizeof(h))) throw std::runtime_error("p2p connect: blob missing or truncated");
        std::memcpy(&h, b, sizeof(h));
        if (h.magic != kBlobMagic || h.rank != r || h.nRanks != grp->nRanks || h.nColors != nColors)
            throw std::runtime_error("p2p connect: blob does not match this group (rank / ranks / colours)");
        if (sizes[r] < static_cast<int64_t>(blobBytes(h.nPeers, h.nColors))) throw std::runtime_error("p2p connect: blob truncated");
        const int32_t *peersOfR = reinterpret_cast<const int32_t *>(b + sizeof(h));
        for (int i = 0; i < h.nPeers; ++i)
            if (peersOfR[i] == grp->rank) idxAtPeer[q] = i;
        if (idxAtPeer[q] < 0) throw std::runtime_error("p2p connect: neighbour does not list this rank as its neighbour");
        peerNPeers[q] = h.nPeers;
        peerRecvSlot[q] = reinterpret_cast<const int64_t *>(b + sizeof(h) + paddedPeersBytes(h.nPeers));
        peerRecvCount[q] = peerRecvSlot[q] + static_cast<size_t>(h.nColors) * h.nPeers;
        for (int c = 0; c < nColors; ++c) {
            const int64_t mine = grp->sendOffsets[static_cast<size_t>(c) * nPeers + q + 1] - grp->sendOffsets[static_cast<size_t>(c) * nPeers + q];
            if (mine != peerRecvCount[q][static_cast<size_t>(c) * h.nPeers + idxAtPeer[q]])
                throw std::runtime_error("p2p connect: send list and the neighbour's halo range disagree");
        }
        if (h.nLocal >= (1ll << 28)) throw std::runtime_error("p2p connect: neighbour holds too many slots for the packed push entry");
        void *px = nullptr, *ps = nullptr;
        SPGX_CUDA(cudaIpcOpenMemHandle(&px, h.x, cudaIpcMemLazyEnablePeerAccess));
        p.mapped.push_back(px);
        SPGX_CUDA(cudaIpcOpenMemHandle(&ps, h.sync, cudaIpcMemLazyEnablePeerAccess));
        p.mapped.push_back(ps);
        dev.peerX[q] = static_cast<double *>(px);
        dev.peerFlag[q] = static_cast<uint32_t *>(ps) + idxAtPeer[q];
    }
    std::vector<uint2> push(p.pushLocal.size());
    for (size_t i = 0; i < push.size(); ++i) {
        const auto &e = p.pushLocal[i];
        const int64_t remote = peerRecvSlot[e.peer][static_cast<size_t>(e.colour) * peerNPeers[e.peer] + idxAtPeer[e.peer]] + e.pos;
        push[i] = make_uint2(static_cast<uint32_t>(e.slot), (static_cast<uint32_t>(e.peer) << 28) | static_cast<uint32_t>(remote));
    }
    cudaStream_t st = grp->ctx->stream;
    p.d_push.alloc(std::max<size_t>(push.size(), 1));
    if (!push.empty())
        SPGX_CUDA(cudaMemcpyAsync(p.d_push.ptr, push.data(), push.size() * sizeof(uint2), cudaMemcpyHostToDevice, st));
    SPGX_CUDA(cudaStreamSynchronize(st));
    dev.myFlag = p.d_sync.ptr;
    dev.epochBase = p.d_sync.ptr + kMaxPeers;
    dev.pushDone = p.d_sync.ptr + kMaxPeers + 1;
    dev.push = p.d_push.ptr;
    dev.error = p.devError;
    double timeoutS = 10.0;
    if (const char *env = getenv("SPGX_HALO_TIMEOUT_S")) timeoutS = std::max(0.001, atof(env));
    dev.timeoutNs = static_cast<unsigned long long>(timeoutS * 1e9);
    dev.debug = getenv("SPGX_DEBUG_HALO") ? atoi(getenv("SPGX_DEBUG_HALO")) : 0;
    p.dev = dev;
    p.connected = true;
    p.mode = SPGX_HALO_P2P;
    if (grp->graphExec) {
        cudaGraphExecDestroy(grp->graphExec);
        grp->graphExec = nullptr;
    }
}

void commUniqueId(void *id128)
{
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    ncclCheck(nccl().GetUniqueId(&id), "ncclGetUniqueId");
    std::memcpy(id128, &id, sizeof(id));
}

void commInit(spgx_ctx *ctx, int rank, int nRanks, const void *id128)
{
    if (ctx->comm) throw StateError("the context already has a communicator");
    ncclUniqueId id;
    std::memcpy(&id, id128, sizeof(id));
    ncclComm_t comm = nullptr;
    ncclCheck(nccl().CommInitRank(&comm, nRanks, id, rank), "ncclCommInitRank");
    ctx->comm = comm;
    ctx->rank = rank;
    ctx->nRanks = nRanks;
}

void commDestroy(spgx_ctx *ctx)
{
    if (ctx->comm) nccl().CommDestroy(static_cast<ncclComm_t>(ctx->comm));
    ctx->comm = nullptr;
}

// after colour pass `colour`: pack the owned boundary vertices of that colour, then one grouped send / recv per peer.
// What arrives lands directly in the contiguous halo range of (peer, colour): no unpack.  Returns kernels launched.
int enqueueHaloExchange(spgx_group *grp, int colour, cudaStream_t st)
{
    const int nPeers = static_cast<int>(grp->peers.size());
    if (nPeers == 0) return 0;
    if (!grp->ctx->comm) throw StateError("partitioned group stepped without a communicator (spgx_ctx_init_comm)");
    auto comm = static_cast<ncclComm_t>(grp->ctx->comm);
    const size_t q0 = static_cast<size_t>(colour) * nPeers;
    const int64_t s0 = grp->sendOffsets[q0], s1 = grp->sendOffsets[q0 + nPeers];
    int launches = 0;
    if (s1 > s0) {
        const int n = static_cast<int>(s1 - s0);
        packKernel<<<(n + 127) / 128, 128, 0, st>>>(grp->d_x.ptr, grp->d_sendSlots.ptr + s0, grp->d_sendBuf.ptr + 3 * s0, n);
        ++launches;
    }
    bool any = false;
    for (int p = 0; p < nPeers; ++p)
        if (grp->sendOffsets[q0 + p + 1] > grp->sendOffsets[q0 + p] || grp->recvCount[q0 + p] > 0) any = true;
    if (!any) return launches;
    ncclCheck(nccl().GroupStart(), "ncclGroupStart");
    for (int p = 0; p < nPeers; ++p) {
        const int64_t a = grp->sendOffsets[q0 + p], b = grp->sendOffsets[q0 + p + 1];
        if (b > a)
            ncclCheck(nccl().Send(grp->d_sendBuf.ptr + 3 * a, static_cast<size_t>(3 * (b - a)), ncclDouble, grp->peers[p], comm, st),
                      "ncclSend");
        const int64_t rc = grp->recvCount[q0 + p];
        if (rc > 0)
            ncclCheck(nccl().Recv(grp->d_x.ptr + 3 * grp->recvSlot[q0 + p], static_cast<size_t>(3 * rc), ncclDouble, grp->peers[p],
                                  comm, st),
                      "ncclRecv");
    }
    ncclCheck(nccl().GroupEnd(), "ncclGroupEnd");
    return launches + 1;
}

// XPBD, after (energy, colour) pass `pass`: pack what the pass moved for each neighbour, one grouped send / recv, scatter
// what arrived into the halo slots.  Returns kernels launched.
int enqueueXpbdHaloExchange(spgx_group *grp, int pass, cudaStream_t st)
{
    const int nPeers = static_cast<int>(grp->peers.size());
    if (nPeers == 0) return 0;
    if (!grp->ctx->comm) throw StateError("partitioned group stepped without a communicator (spgx_ctx_init_comm) or peer connection");
    auto comm = static_cast<ncclComm_t>(grp->ctx->comm);
    const auto &xh = grp->xh;
    const size_t q0 = static_cast<size_t>(pass) * nPeers;
    const int64_t s0 = xh.sendOffsets[q0], s1 = xh.sendOffsets[q0 + nPeers];
    const int64_t r0 = xh.recvOffsets[q0], r1 = xh.recvOffsets[q0 + nPeers];
    if (s1 == s0 && r1 == r0) return 0;
    int launches = 0;
    if (s1 > s0) {
        const int n = static_cast<int>(s1 - s0);
        packKernel<<<(n + 127) / 128, 128, 0, st>>>(grp->d_x.ptr, grp->d_sendSlots.ptr + s0, grp->d_sendBuf.ptr + 3 * s0, n);
        ++launches;
    }
    ncclCheck(nccl().GroupStart(), "ncclGroupStart");
    for (int p = 0; p < nPeers; ++p) {
        const int64_t a = xh.sendOffsets[q0 + p], b = xh.sendOffsets[q0 + p + 1];
        if (b > a)
            ncclCheck(nccl().Send(grp->d_sendBuf.ptr + 3 * a, static_cast<size_t>(3 * (b - a)), ncclDouble, grp->peers[p], comm, st), "ncclSend");
        const int64_t c = xh.recvOffsets[q0 + p], d = xh.recvOffsets[q0 + p + 1];
        if (d > c)
            ncclCheck(nccl().Recv(grp->xh.d_recvBuf.ptr + 3 * c, static_cast<size_t>(3 * (d - c)), ncclDouble, grp->peers[p], comm, st), "ncclRecv");
    }
    ncclCheck(nccl().GroupEnd(), "ncclGroupEnd");
    ++launches;
    if (r1 > r0) {
        const int n = static_cast<int>(r1 - r0);
        unpackKernel<<<(n + 127) / 128, 128, 0, st>>>(grp->d_x.ptr, grp->xh.d_recvSlots.ptr + r0, grp->xh.d_recvBuf.ptr + 3 * r0, n);
        ++launches;
    }
    return launches;
}
}  // namespace spgx
