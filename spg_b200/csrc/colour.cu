// Vertex colouring built on the device, bit-identical to the reference's sequential greedy colouring
// (spg/source/utils/graphColoring.cpp:21-49 through colorVertices :72-84).
//
// Graph::greedyColoring visits the nodes in index order and gives node n the smallest colour not used by an ALREADY
// COLOURED neighbour, i.e. by a neighbour with a smaller index.  colour(n) is therefore a pure function of the colours of
// n's lower-indexed neighbours: a dependency DAG.  Evaluating that DAG level by level -- a node is coloured as soon as all of
// its lower-indexed neighbours are -- gives exactly the sequential result, with every level a parallel kernel.  (This is
// Jones-Plassmann with the node index as the priority; a random priority would be faster to converge and would NOT reproduce
// the reference's colour classes.)  The reference's own implementation is O(N^2) (40 s at 1 M vertices, infeasible at 16 M,
// SURVEY 8a3); host_build.h does it in O(N deg) on one core (0.4 s at 1 M, 6.6 s at 16 M); this one takes as many rounds as
// the longest increasing-index path (a few thousand on a row-major grid) of a few microseconds each.
//
// Adjacency is never materialised: two vertices are neighbours iff they share a stencil, so a node walks
// node -> incident stencils -> their vertices.  The node -> stencil incidence is built here on the device (count, scan, fill;
// the order inside a row is irrelevant to a set-valued first-fit).  Readiness is tracked with a counter per node of its
// (stencil, lower-indexed vertex) pairs; a node that gets coloured decrements the counters of its higher-indexed neighbours
// with the same multiplicity, and the one that takes a counter to zero appends that node to the next frontier.
#include <cooperative_groups.h>

#include <algorithm>
#include <mutex>
#include <vector>

#include "group.h"
#include "host_build.h"

namespace spgx
{
namespace
{
constexpr int kMaxSets = kMaxEnergies;
struct ColourSets {
    int nSets;
    const int32_t *entries[kMaxSets];  // vertex ids, S per stencil
    int S[kMaxSets];
    long long entryBase[kMaxSets + 1];  // first flat entry (over all sets) of each set
    int edgeBase[kMaxSets + 1];         // first global stencil id of each set
};

__device__ __forceinline__ int setOfEntry(const ColourSets &cs, long long flat)
{
    int k = 0;
    while (k + 1 < cs.nSets && flat >= cs.entryBase[k + 1]) ++k;
    return k;
}
__device__ __forceinline__ int setOfEdge(const ColourSets &cs, int edge)
{
    int k = 0;
    while (k + 1 < cs.nSets && edge >= cs.edgeBase[k + 1]) ++k;
    return k;
}

__global__ void degreeKernel(const ColourSets cs, long long nEntries, uint32_t *deg)
{
    const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= nEntries) return;
    const int k = setOfEntry(cs, i);
    atomicAdd(deg + cs.entries[k][i - cs.entryBase[k]], 1u);
}
__global__ void fillKernel(const ColourSets cs, long long nEntries, const uint32_t *nodePtr, uint32_t *cursor, int32_t *nodeEdges)
{
    const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= nEntries) return;
    const int k = setOfEntry(cs, i);
    const long long local = i - cs.entryBase[k];
    const int node = cs.entries[k][local];
    const uint32_t pos = atomicAdd(cursor + node, 1u);
    nodeEdges[nodePtr[node] + pos] = cs.edgeBase[k] + static_cast<int>(local / cs.S[k]);
}

// number of (incident stencil, lower-indexed vertex of it) pairs per node; nodes with none start the first frontier
__global__ void dependencyKernel(const ColourSets cs, int nNodes, const uint32_t *nodePtr, const int32_t *nodeEdges, uint32_t *cnt,
                                 int32_t *frontier, uint32_t *frontierSize)
{
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nNodes) return;
    uint32_t c = 0;
    for (uint32_t q = nodePtr[v]; q < nodePtr[v + 1]; ++q) {
        const int e = nodeEdges[q], k = setOfEdge(cs, e);
        const int32_t *st = cs.entries[k] + static_cast<long long>(e - cs.edgeBase[k]) * cs.S[k];
        for (int l = 0; l < cs.S[k]; ++l) c += st[l] < v;
    }
    cnt[v] = c;
    if (c == 0) frontier[atomicAdd(frontierSize, 1u)] = v;
}

__device__ __forceinline__ void colourNode(const ColourSets &cs, const uint32_t *nodePtr, const int32_t *nodeEdges, int v, int32_t *colour,
                                           uint32_t *cnt, int32_t *next, uint32_t *nextSize);

// one level of the DAG: every node of the frontier takes the smallest colour none of its lower-indexed neighbours has
// (their colours are final: they were written in earlier launches), then releases its higher-indexed neighbours
__global__ void colourLevelKernel(const ColourSets cs, const uint32_t *nodePtr, const int32_t *nodeEdges, const int32_t *frontier,
                                  uint32_t frontierSize, int32_t *colour, uint32_t *cnt, int32_t *next, uint32_t *nextSize)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= frontierSize) return;
    colourNode(cs, nodePtr, nodeEdges, frontier[t], colour, cnt, next, nextSize);
}

// the body of one node of a level (shared by the two kernels below)
__device__ __forceinline__ void colourNode(const ColourSets &cs, const uint32_t *nodePtr, const int32_t *nodeEdges, int v, int32_t *colour,
                                           uint32_t *cnt, int32_t *next, uint32_t *nextSize)
{
    const uint32_t qb = nodePtr[v], qe = nodePtr[v + 1];
    int chosen = -1;
    for (int base = 0; chosen < 0; base += 64) {
        unsigned long long used = 0ull;
        for (uint32_t q = qb; q < qe; ++q) {
            const int e = nodeEdges[q], k = setOfEdge(cs, e);
            const int32_t *st = cs.entries[k] + static_cast<long long>(e - cs.edgeBase[k]) * cs.S[k];
            for (int l = 0; l < cs.S[k]; ++l) {
                const int u = st[l];
                if (u < v) {
                    const int c = __ldcg(colour + u) - base;  // written by another thread in an earlier level: read at L2
                    if (c >= 0 && c < 64) used |= 1ull << c;
                }
            }
        }
        if (~used) chosen = base + __ffsll(static_cast<long long>(~used)) - 1;
    }
    __stcg(colour + v, chosen);
    for (uint32_t q = qb; q < qe; ++q) {
        const int e = nodeEdges[q], k = setOfEdge(cs, e);
        const int32_t *st = cs.entries[k] + static_cast<long long>(e - cs.edgeBase[k]) * cs.S[k];
        for (int l = 0; l < cs.S[k]; ++l) {
            const int w = st[l];
            if (w > v && atomicSub(cnt + w, 1u) == 1u) next[atomicAdd(nextSize, 1u)] = w;
        }
    }
}

// one node by one WARP: the lanes stride over the node's (incident stencil, vertex) pairs, so a node costs three dependent
// memory round trips (incidence -> stencil -> colour) instead of that many per pair
__device__ __forceinline__ void colourNodeWarp(const ColourSets &cs, const uint32_t *nodePtr, const int32_t *nodeEdges, int v, int lane,
                                               int32_t *colour, uint32_t *cnt, int32_t *next, uint32_t *nextSize)
{
    const uint32_t qb = nodePtr[v], qe = nodePtr[v + 1];
    int chosen = -1;
    for (int base = 0; chosen < 0; base += 64) {
        unsigned long long used = 0ull;
        for (uint32_t q = qb + lane; q < qe; q += 32) {
            const int e = nodeEdges[q], k = setOfEdge(cs, e);
            const int32_t *st = cs.entries[k] + static_cast<long long>(e - cs.edgeBase[k]) * cs.S[k];
            for (int l = 0; l < cs.S[k]; ++l) {
                const int u = st[l];
                if (u < v) {
                    const int c = __ldcg(colour + u) - base;
                    if (c >= 0 && c < 64) used |= 1ull << c;
                }
            }
        }
        const unsigned lo = __reduce_or_sync(0xffffffffu, static_cast<unsigned>(used));
        const unsigned hi = __reduce_or_sync(0xffffffffu, static_cast<unsigned>(used >> 32));
        const unsigned long long all = (static_cast<unsigned long long>(hi) << 32) | lo;
        if (~all) chosen = base + __ffsll(static_cast<long long>(~all)) - 1;
    }
    if (lane == 0) __stcg(colour + v, chosen);
    for (uint32_t q = qb + lane; q < qe; q += 32) {
        const int e = nodeEdges[q], k = setOfEdge(cs, e);
        const int32_t *st = cs.entries[k] + static_cast<long long>(e - cs.edgeBase[k]) * cs.S[k];
        for (int l = 0; l < cs.S[k]; ++l) {
            const int w = st[l];
            if (w > v && atomicSub(cnt + w, 1u) == 1u) next[atomicAdd(nextSize, 1u)] = w;
        }
    }
}

// EVERY level in one cooperative launch: a grid barrier between levels instead of a launch, a copy and a host
// synchronisation, a warp per node.  Three frontier sizes rotate: level L reads sizes[L % 3], appends through
// sizes[(L + 1) % 3] and clears sizes[(L + 2) % 3] (last read at the start of level L - 1).  state: [levels, nodes coloured]
constexpr int kLevelThreads = 512;
__global__ void __launch_bounds__(kLevelThreads) colourAllLevelsKernel(const ColourSets cs, const uint32_t *nodePtr, const int32_t *nodeEdges,
                                                                       int32_t *frontierA, int32_t *frontierB, uint32_t *sizes,
                                                                       int32_t *colour, uint32_t *cnt, uint32_t *state)
{
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    const int lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nWarps = (gridDim.x * blockDim.x) >> 5;
    uint32_t level = 0, coloured = 0;
    for (;;) {
        const uint32_t size = *reinterpret_cast<volatile uint32_t *>(sizes + level % 3);
        if (size == 0) break;
        if (blockIdx.x == 0 && threadIdx.x == 0) *reinterpret_cast<volatile uint32_t *>(sizes + (level + 2) % 3) = 0u;
        const int32_t *frontier = (level & 1u) ? frontierB : frontierA;
        int32_t *next = (level & 1u) ? frontierA : frontierB;
        for (uint32_t t = warp; t < size; t += nWarps)
            colourNodeWarp(cs, nodePtr, nodeEdges, __ldcg(frontier + t), lane, colour, cnt, next, sizes + (level + 1) % 3);
        coloured += size;
        ++level;
        grid.sync();
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        state[0] = level;
        state[1] = coloured;
    }
}
}  // namespace

// colorVertices (graphColoring.cpp:72-84) on the device of `ctx`; *rounds (optional): levels of the dependency DAG
std::vector<int32_t> colorVerticesDevice(spgx_ctx *ctx, int nVerts, const std::vector<FlatStencilsView> &sets, int *rounds)
{
    std::vector<int32_t> colours(static_cast<size_t>(nVerts), -1);
    if (rounds) *rounds = 0;
    if (nVerts == 0) return colours;
    if (static_cast<int>(sets.size()) > kMaxSets) throw std::runtime_error("too many stencil sets for the device colouring");
    SPGX_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    ColourSets cs{};
    cs.nSets = static_cast<int>(sets.size());
    std::vector<DevBuf<int32_t>> d_entries(sets.size());
    long long nEntries = 0;
    long long nEdges = 0;
    for (size_t k = 0; k < sets.size(); ++k) {
        const auto &fs = sets[k];
        for (int64_t i = 0; i < fs.nEntries; ++i)
            if (fs.entries[i] < 0 || fs.entries[i] >= nVerts) throw std::runtime_error("Trying to add edge to graph with invalid node indices");
        d_entries[k].alloc(static_cast<size_t>(std::max<int64_t>(fs.nEntries, 1)));
        if (fs.nEntries)
            SPGX_CUDA(cudaMemcpyAsync(d_entries[k].ptr, fs.entries, static_cast<size_t>(fs.nEntries) * sizeof(int32_t), cudaMemcpyHostToDevice, st));
        cs.entries[k] = d_entries[k].ptr;
        cs.S[k] = fs.stencilSize;
        cs.entryBase[k] = nEntries;
        cs.edgeBase[k] = static_cast<int>(nEdges);
        nEntries += fs.nEntries;
        nEdges += fs.nEntries / fs.stencilSize;
    }
    cs.entryBase[sets.size()] = nEntries;
    cs.edgeBase[sets.size()] = static_cast<int>(nEdges);
    if (nEntries >= (1ll << 32) || nEdges >= (1ll << 31)) throw std::runtime_error("too many stencil entries for the device colouring");
    const int threads = 256;
    const auto blocks = [&](long long n) { return static_cast<unsigned>((n + threads - 1) / threads); };

    // node -> incident stencils: count, scan (host: one pass over nVerts integers), fill
    DevBuf<uint32_t> d_deg, d_nodePtr, d_cnt, d_sizes;
    DevBuf<int32_t> d_nodeEdges, d_colour, d_frontier[2];
    d_deg.alloc(static_cast<size_t>(nVerts));
    SPGX_CUDA(cudaMemsetAsync(d_deg.ptr, 0, static_cast<size_t>(nVerts) * sizeof(uint32_t), st));
    if (nEntries) degreeKernel<<<blocks(nEntries), threads, 0, st>>>(cs, nEntries, d_deg.ptr);
    std::vector<uint32_t> nodePtr(static_cast<size_t>(nVerts) + 1, 0);
    SPGX_CUDA(cudaMemcpyAsync(nodePtr.data() + 1, d_deg.ptr, static_cast<size_t>(nVerts) * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    SPGX_CUDA(cudaStreamSynchronize(st));
    for (int v = 0; v < nVerts; ++v) nodePtr[v + 1] += nodePtr[v];
    d_nodePtr.upload(nodePtr, st);
    d_nodeEdges.alloc(static_cast<size_t>(std::max<long long>(nEntries, 1)));
    SPGX_CUDA(cudaMemsetAsync(d_deg.ptr, 0, static_cast<size_t>(nVerts) * sizeof(uint32_t), st));  // reused as the fill cursor
    if (nEntries) fillKernel<<<blocks(nEntries), threads, 0, st>>>(cs, nEntries, d_nodePtr.ptr, d_deg.ptr, d_nodeEdges.ptr);

    d_cnt.alloc(static_cast<size_t>(nVerts));
    d_colour.alloc(static_cast<size_t>(nVerts));
    d_frontier[0].alloc(static_cast<size_t>(nVerts));
    d_frontier[1].alloc(static_cast<size_t>(nVerts));
    d_sizes.alloc(3);
    SPGX_CUDA(cudaMemsetAsync(d_sizes.ptr, 0, 3 * sizeof(uint32_t), st));
    SPGX_CUDA(cudaMemsetAsync(d_colour.ptr, 0xFF, static_cast<size_t>(nVerts) * sizeof(int32_t), st));
    dependencyKernel<<<blocks(nVerts), threads, 0, st>>>(cs, nVerts, d_nodePtr.ptr, d_nodeEdges.ptr, d_cnt.ptr, d_frontier[0].ptr, d_sizes.ptr);

    // levels: one cooperative launch walks all of them (grid barrier in between); where the device cannot co-schedule the grid
    // one launch per level from the host (colourLevelKernel)
    DevBuf<uint32_t> d_state;
    d_state.alloc(2);
    SPGX_CUDA(cudaMemsetAsync(d_state.ptr, 0, 2 * sizeof(uint32_t), st));
    uint32_t *host = nullptr;  // pinned
    SPGX_CUDA(cudaHostAlloc(reinterpret_cast<void **>(&host), 4 * sizeof(uint32_t), cudaHostAllocDefault));
    try {
        int coop = 0, perSm = 0;
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, ctx->device);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, colourAllLevelsKernel, kLevelThreads, 0);
        bool doneOnDevice = false;
        if (coop && perSm >= 1) {
            // enough warps for a diagonal wave of a 16 M-vertex grid, few enough blocks for a cheap barrier
            const int gridBlocks = std::min(ctx->prop.multiProcessorCount, 96);
            int32_t *fa = d_frontier[0].ptr, *fb = d_frontier[1].ptr;
            uint32_t *sizesPtr = d_sizes.ptr, *cntPtr = d_cnt.ptr, *statePtr = d_state.ptr;
            int32_t *colourPtr = d_colour.ptr;
            const uint32_t *np = d_nodePtr.ptr;
            const int32_t *ne = d_nodeEdges.ptr;
            void *kargs[] = {&cs, &np, &ne, &fa, &fb, &sizesPtr, &colourPtr, &cntPtr, &statePtr};
            if (cudaLaunchCooperativeKernel(reinterpret_cast<void *>(colourAllLevelsKernel), dim3(gridBlocks), dim3(kLevelThreads), kargs, 0, st) ==
                cudaSuccess) {
                SPGX_CUDA(cudaMemcpyAsync(host, d_state.ptr, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
                SPGX_CUDA(cudaStreamSynchronize(st));
                doneOnDevice = true;
            } else {
                cudaGetLastError();
            }
        }
        if (!doneOnDevice) {
            host[0] = host[1] = 0;
            for (uint32_t level = 0;; ++level) {
                SPGX_CUDA(cudaMemcpyAsync(host + 2, d_sizes.ptr + level % 3, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
                SPGX_CUDA(cudaStreamSynchronize(st));
                const uint32_t size = host[2];
                if (size == 0) break;
                SPGX_CUDA(cudaMemsetAsync(d_sizes.ptr + (level + 1) % 3, 0, sizeof(uint32_t), st));
                colourLevelKernel<<<blocks(size), threads, 0, st>>>(cs, d_nodePtr.ptr, d_nodeEdges.ptr, d_frontier[level & 1u].ptr, size, d_colour.ptr,
                                                                   d_cnt.ptr, d_frontier[1 - (level & 1u)].ptr, d_sizes.ptr + (level + 1) % 3);
                host[0] = level + 1;
                host[1] += size;
            }
        }
        SPGX_CUDA(cudaGetLastError());
        if (static_cast<long long>(host[1]) != nVerts) throw std::runtime_error("device colouring: the dependency levels did not reach every vertex");
        if (rounds) *rounds = static_cast<int>(host[0]);
        SPGX_CUDA(cudaMemcpyAsync(colours.data(), d_colour.ptr, static_cast<size_t>(nVerts) * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        SPGX_CUDA(cudaStreamSynchronize(st));
    } catch (...) {
        cudaFreeHost(host);
        throw;
    }
    cudaFreeHost(host);
    return colours;
}
}  // namespace spgx

using namespace spgx;

extern "C" int spgx_color_vertices_device(spgx_ctx *ctx, int nVerts, int nSets, const int32_t *const *entries, const int64_t *nEntries,
                                          const int32_t *stencilSizes, int32_t *colors, int *rounds)
{
    return guarded([&] {
        if (!ctx) throw StateError("null context");
        std::vector<FlatStencilsView> sets;
        for (int s = 0; s < nSets; ++s) sets.push_back({entries[s], nEntries[s], stencilSizes[s]});
        const auto c = colorVerticesDevice(ctx, nVerts, sets, rounds);
        std::copy(c.begin(), c.end(), colors);
    });
}

// ---- the same colouring for callers that may not allocate -------------------------------------------------------------------
// spgx_color_vertices_noalloc: no counterpart in the reference; added for callers that
// cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Everything
// colorVerticesDevice allocates per call -- incidence counts, row pointers, incidence lists, readiness counters, the two
// frontiers, the frontier sizes and the result state -- lives in ONE file-scope device array of a fixed, modest size
// (kNoAllocMaxVerts vertices, kNoAllocMaxEntries stencil entries: about 13 MiB, held by every GPU that loads the library
// until exit).  Inputs and the output are the caller's DEVICE buffers; the row-pointer scan runs on the device instead of the
// host.  Every call shares the array and finds the previous call's contents in it: what has to start at zero is cleared at
// the start of each call, and two calls that overlap on the GPU (two streams, two threads, two contexts on one device) race.
namespace spgx
{
namespace
{
constexpr int kNoAllocMaxVerts = SPGX_NOALLOC_MAX_VERTS;            // include/spgx.h
constexpr long long kNoAllocMaxEntries = SPGX_NOALLOC_MAX_ENTRIES;

struct NoAllocScratch {
    uint32_t sizes[4];   // three rotating frontier sizes (+ pad)
    uint32_t state[4];   // levels, nodes coloured, out-of-range entries seen, pad
    uint32_t pad[56];    // the arrays below start on a 256-byte boundary
    uint32_t deg[kNoAllocMaxVerts];           // incidence counts, then the fill cursor
    uint32_t nodePtr[kNoAllocMaxVerts + 64];  // nVerts + 1 used
    uint32_t cnt[kNoAllocMaxVerts];
    int32_t frontier[2][kNoAllocMaxVerts];
    int32_t nodeEdges[kNoAllocMaxEntries];
};
}  // namespace
}  // namespace spgx

// In host code this name is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress (below).
static __device__ __align__(256) unsigned char g_colourNoAllocScratch[sizeof(spgx::NoAllocScratch)];

namespace spgx
{
namespace
{
// an entry outside [0, nVerts) is counted, not followed (the allocating entry point validates on the host instead)
__global__ void rangeCheckKernel(const ColourSets cs, long long nEntries, int nVerts, uint32_t *bad)
{
    const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= nEntries) return;
    const int k = setOfEntry(cs, i);
    const int v = cs.entries[k][i - cs.entryBase[k]];
    if (v < 0 || v >= nVerts) atomicAdd(bad, 1u);
}

// nodePtr[0] = 0, nodePtr[v + 1] = deg[0] + ... + deg[v]: one block, a contiguous run of nodes per thread, the per-thread
// totals scanned through shared memory (the allocating entry point scans on the host)
constexpr int kScanThreads = 1024;
__global__ void __launch_bounds__(kScanThreads) rowPointerKernel(const uint32_t *deg, uint32_t *nodePtr, int n)
{
    __shared__ uint32_t warpTotals[kScanThreads / 32];
    const int per = (n + kScanThreads - 1) / kScanThreads;
    const int begin = min(n, static_cast<int>(threadIdx.x) * per), end = min(n, begin + per);
    uint32_t sum = 0;
    for (int v = begin; v < end; ++v) sum += deg[v];
    uint32_t incl = sum;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += up;
    }
    if (lane == 31) warpTotals[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = warpTotals[lane];
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, w, d);
            if (lane >= d) w += up;
        }
        warpTotals[lane] = w;
    }
    __syncthreads();
    uint32_t run = incl - sum + (warp ? warpTotals[warp - 1] : 0u);
    if (threadIdx.x == 0) nodePtr[0] = 0u;
    for (int v = begin; v < end; ++v) {
        run += deg[v];
        nodePtr[v + 1] = run;
    }
}

// the device address of the scratch array, taken once per device with that device current
NoAllocScratch *noAllocScratch(int device)
{
    constexpr int kMaxDevices = 64;
    static std::mutex mutex;
    static NoAllocScratch *address[kMaxDevices] = {};
    if (device < 0 || device >= kMaxDevices) throw std::runtime_error("spgx_color_vertices_noalloc: device ordinal out of range");
    std::lock_guard<std::mutex> lock(mutex);
    if (!address[device]) {
        SPGX_CUDA(cudaSetDevice(device));
        void *p = nullptr;
        SPGX_CUDA(cudaGetSymbolAddress(&p, g_colourNoAllocScratch));
        address[device] = static_cast<NoAllocScratch *>(p);
    }
    return address[device];
}
}  // namespace
}  // namespace spgx

// d_entries: nSets DEVICE arrays of vertex ids (the array of pointers itself is host memory); d_colors: nVerts ints on the
// device; *rounds (host, may be NULL): levels of the dependency DAG.  Runs on the context's stream and returns after it.
extern "C" int spgx_color_vertices_noalloc(spgx_ctx *ctx, int nVerts, int nSets, const int32_t *const *d_entries, const int64_t *nEntries,
                                           const int32_t *stencilSizes, int32_t *d_colors, int *rounds)
{
    return guarded([&] {
        if (!ctx) throw StateError("null context");
        if (rounds) *rounds = 0;
        if (nVerts < 0 || nSets < 0) throw std::runtime_error("spgx_color_vertices_noalloc: negative size");
        if (nVerts > kNoAllocMaxVerts)
            throw std::runtime_error("spgx_color_vertices_noalloc: " + std::to_string(nVerts) + " vertices exceed the static scratch (" +
                                     std::to_string(kNoAllocMaxVerts) + "); use spgx_color_vertices_device");
        if (nSets > kMaxSets) throw std::runtime_error("too many stencil sets for the device colouring");
        ColourSets cs{};
        cs.nSets = nSets;
        long long total = 0, nEdges = 0;
        for (int k = 0; k < nSets; ++k) {
            if (nEntries[k] < 0 || stencilSizes[k] <= 0 || nEntries[k] % stencilSizes[k]) throw std::runtime_error("spgx_color_vertices_noalloc: bad stencil set");
            cs.entries[k] = d_entries[k];
            cs.S[k] = stencilSizes[k];
            cs.entryBase[k] = total;
            cs.edgeBase[k] = static_cast<int>(nEdges);
            total += nEntries[k];
            nEdges += nEntries[k] / stencilSizes[k];
            if (total > kNoAllocMaxEntries)
                throw std::runtime_error("spgx_color_vertices_noalloc: the stencil entries exceed the static scratch (" +
                                         std::to_string(kNoAllocMaxEntries) + "); use spgx_color_vertices_device");
        }
        cs.entryBase[nSets] = total;
        cs.edgeBase[nSets] = static_cast<int>(nEdges);
        if (nVerts == 0) return;
        if (!d_colors) throw std::runtime_error("spgx_color_vertices_noalloc: null output");
        NoAllocScratch *sc = noAllocScratch(ctx->device);  // the device address, never the host-side name of the array
        SPGX_CUDA(cudaSetDevice(ctx->device));
        cudaStream_t st = ctx->stream;
        const int threads = 256;
        const auto blocks = [&](long long n) { return static_cast<unsigned>((n + threads - 1) / threads); };
        const size_t vBytes = static_cast<size_t>(nVerts) * sizeof(uint32_t);

        // the array keeps the previous call's contents: sizes, state and the incidence counts start at zero
        SPGX_CUDA(cudaMemsetAsync(sc->sizes, 0, sizeof(sc->sizes) + sizeof(sc->state), st));
        SPGX_CUDA(cudaMemsetAsync(sc->deg, 0, vBytes, st));
        SPGX_CUDA(cudaMemsetAsync(d_colors, 0xFF, vBytes, st));
        uint32_t host[4] = {0, 0, 0, 0};
        if (total) {
            rangeCheckKernel<<<blocks(total), threads, 0, st>>>(cs, total, nVerts, sc->state + 2);
            SPGX_CUDA(cudaMemcpyAsync(host, sc->state, sizeof(host), cudaMemcpyDeviceToHost, st));
            SPGX_CUDA(cudaStreamSynchronize(st));
            if (host[2]) throw std::runtime_error("Trying to add edge to graph with invalid node indices");
            degreeKernel<<<blocks(total), threads, 0, st>>>(cs, total, sc->deg);
        }
        rowPointerKernel<<<1, kScanThreads, 0, st>>>(sc->deg, sc->nodePtr, nVerts);
        SPGX_CUDA(cudaMemsetAsync(sc->deg, 0, vBytes, st));  // reused as the fill cursor
        if (total) fillKernel<<<blocks(total), threads, 0, st>>>(cs, total, sc->nodePtr, sc->deg, sc->nodeEdges);
        dependencyKernel<<<blocks(nVerts), threads, 0, st>>>(cs, nVerts, sc->nodePtr, sc->nodeEdges, sc->cnt, sc->frontier[0], sc->sizes);

        int coop = 0, perSm = 0;
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, ctx->device);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, colourAllLevelsKernel, kLevelThreads, 0);
        bool doneOnDevice = false;
        if (coop && perSm >= 1) {
            const int gridBlocks = std::min(ctx->prop.multiProcessorCount, 96);
            int32_t *fa = sc->frontier[0], *fb = sc->frontier[1], *colourPtr = d_colors;
            uint32_t *sizesPtr = sc->sizes, *cntPtr = sc->cnt, *statePtr = sc->state;
            const uint32_t *np = sc->nodePtr;
            const int32_t *ne = sc->nodeEdges;
            void *kargs[] = {&cs, &np, &ne, &fa, &fb, &sizesPtr, &colourPtr, &cntPtr, &statePtr};
            if (cudaLaunchCooperativeKernel(reinterpret_cast<void *>(colourAllLevelsKernel), dim3(gridBlocks), dim3(kLevelThreads), kargs, 0, st) ==
                cudaSuccess) {
                SPGX_CUDA(cudaMemcpyAsync(host, sc->state, sizeof(host), cudaMemcpyDeviceToHost, st));
                SPGX_CUDA(cudaStreamSynchronize(st));
                doneOnDevice = true;
            } else {
                cudaGetLastError();
            }
        }
        if (!doneOnDevice) {  // one launch per level, the frontier size read back in between
            host[0] = host[1] = 0;
            for (uint32_t level = 0;; ++level) {
                uint32_t size = 0;
                SPGX_CUDA(cudaMemcpyAsync(&size, sc->sizes + level % 3, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
                SPGX_CUDA(cudaStreamSynchronize(st));
                if (size == 0) break;
                SPGX_CUDA(cudaMemsetAsync(sc->sizes + (level + 1) % 3, 0, sizeof(uint32_t), st));
                colourLevelKernel<<<blocks(size), threads, 0, st>>>(cs, sc->nodePtr, sc->nodeEdges, sc->frontier[level & 1u], size, d_colors, sc->cnt,
                                                                   sc->frontier[1 - (level & 1u)], sc->sizes + (level + 1) % 3);
                host[0] = level + 1;
                host[1] += size;
            }
            SPGX_CUDA(cudaStreamSynchronize(st));
        }
        SPGX_CUDA(cudaGetLastError());
        if (static_cast<long long>(host[1]) != nVerts) throw std::runtime_error("device colouring: the dependency levels did not reach every vertex");
        if (rounds) *rounds = static_cast<int>(host[0]);
    });
}
