// Multi-GPU plumbing of the VBD sweep: one process per GPU, NCCL point-to-point over NVLink for the per-colour halo
// exchange (SURVEY.md section 8e / north_star "per-colour boundary-vertex halo exchange over NVLink via NCCL").
//
// NCCL is bound at run time (dlopen of libnccl.so.2: the copy the host process already loaded -- e.g. the one bundled
// with torch when the ranks were launched by torchrun -- else the system one), so that libspgx.so itself has no link
// dependency on it and single-GPU users never touch it.  Only the stable point-to-point subset is used.
#include <dlfcn.h>
#include <nccl.h>

#include <cstring>

#include "group.h"

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
}  // namespace

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
}  // namespace spgx
