// Persistent form of the streaming VBD colour pass (vbd_stream.cuh; reference: spg/source/sim/solver/vbd.cpp:99-157).
//
// vbdStreamKernel gives every warp ONE slice, so each warp pays the dependent chain  descriptor -> neighbour ids ->
// neighbour positions -> first record  before its first evaluation and  mass / inertial position -> solve  after its
// last one: 25-35 % of the resident-warp time of a cloth pass (profiles/r1d_*), hidden only by the other three warps of
// the scheduler.  Here a warp walks the slices w, w + W, w + 2W, ... of the colour (W = warps of the grid, sized to the
// resident capacity) and fetches everything slice i + W needs WHILE it evaluates slice i, without holding registers:
//   A (top of slice i)     cp.async  the two descriptors of slice i + W                       -> shared
//   B (after 2 steps)      cp.async  its neighbour-id rows; TMA bulk prefetch of its streams  -> shared / L2
//   C (after 4 steps)      cp.async  neighbour positions into the OTHER tile buffer, and the owner's x, inertial
//                          position and mass into the other own-state buffer                  -> shared
//   D (before the solve)   the first incidence record of slice i + W                          -> registers
// so the only exposed latencies left are those of a warp's first slice (which programmatic dependent launch overlaps
// with the previous pass).  Evaluation order per vertex is unchanged: results are bit-identical to vbdStreamKernel.
#pragma once
#include "vbd_stream.cuh"

namespace spgx
{
__device__ __forceinline__ void cpAsync4(void *smemDst, const void *gmemSrc)
{
    const unsigned s = static_cast<unsigned>(__cvta_generic_to_shared(smemDst));
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(gmemSrc));
}
__device__ __forceinline__ void cpAsync16(void *smemDst, const void *gmemSrc)
{
    const unsigned s = static_cast<unsigned>(__cvta_generic_to_shared(smemDst));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmemSrc));
}

// per-warp shared memory: two neighbour tiles, two own-state blocks (x, inertial x, mass per vertex), one id block, one
// descriptor pair
struct PersistLayout {
    int tileDoubles, ownDoubles, idWords;
    __host__ __device__ PersistLayout(int maxNbr, int VW)
        : tileDoubles(maxNbr * 3 * VW), ownDoubles(8 * VW), idWords(maxNbr * VW)
    {
    }
    __host__ __device__ size_t bytesPerWarp() const
    {
        return (2 * static_cast<size_t>(tileDoubles) + 2 * ownDoubles) * 8 + (static_cast<size_t>(idWords) * 4 + 15) / 16 * 16 + 64;
    }
};

template <unsigned SMASK, unsigned KMASK, int G, int OCC>
__global__ void __launch_bounds__(kStreamBlockThreads, OCC)
    vbdStreamPersistKernel(const __grid_constant__ VbdArgs a, int end, int sliceBase, int nSlicesOfColour, double invdtSquared)
{
    constexpr int VW = 32 / G;  // vertices per warp
    constexpr int kWarps = kStreamBlockThreads / 32;
    extern __shared__ __align__(16) double smemRaw[];
    const int warpInBlock = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int W = static_cast<int>(gridDim.x) * kWarps;  // slices in flight = stride of a warp
    const int w = blockIdx.x * kWarps + warpInBlock;
    asm volatile("griddepcontrol.launch_dependents;");
    if (w >= nSlicesOfColour) return;
    const int v = lane % VW, sub = lane / VW;

    const PersistLayout L(a.maxNbr, VW);
    char *base = reinterpret_cast<char *>(smemRaw) + static_cast<size_t>(warpInBlock) * L.bytesPerWarp();
    double *const tile0 = reinterpret_cast<double *>(base);
    double *const own0 = tile0 + 2 * L.tileDoubles;
    auto tileOf = [&](int buf) { return tile0 + buf * L.tileDoubles; };  // (pointer arithmetic, not an indexed array:
    auto ownOf = [&](int buf) { return own0 + buf * L.ownDoubles; };     //  an array would live in local memory)
    uint32_t *ids = reinterpret_cast<uint32_t *>(own0 + 2 * L.ownDoubles);
    uint32_t *descS = ids + (L.idWords + 3) / 4 * 4;  // 16 words: descriptor of the slice and of the one after it

    // ---- fetch stages (all global -> shared, no registers held across them) ----
    auto fetchDesc = [&](int slice) {
        if (lane < 4) cpAsync16(descS + 4 * lane, a.sliceDesc + static_cast<size_t>(slice) * kSliceDescWords + 4 * lane);
        cpAsyncCommit();
    };
    auto readDesc = [&](SliceDesc &d, SliceDesc &n) {
        const uint4 *p = reinterpret_cast<const uint4 *>(descS);
        d.lo = p[0], d.hi = p[1], n.lo = p[2], n.hi = p[3];
    };
    auto fetchIds = [&](const SliceDesc &d, const SliceDesc &n) {
        const uint32_t nb = d.lo.y, ne = n.lo.y;
        for (uint32_t k = nb + sub; k < ne; k += G) cpAsync4(ids + (k - nb) * VW + v, a.nbrIds + static_cast<size_t>(k) * VW + v);
        cpAsyncCommit();
        if (lane == 0) {  // the slice's incidence streams: HBM -> L2, one slice-time ahead of their use
            prefetchStreamIfPresent<SMASK, ST_BEND_EDGE>(a, d, n);
            prefetchStreamIfPresent<SMASK, ST_BEND_OPP>(a, d, n);
            prefetchStreamIfPresent<SMASK, ST_MEMBRANE_STVK>(a, d, n);
            prefetchStreamIfPresent<SMASK, ST_TET_SNH>(a, d, n);
            prefetchStreamIfPresent<SMASK, ST_TET_STVK>(a, d, n);
        }
    };
    auto gather = [&](const SliceDesc &d, const SliceDesc &n, int buf) {
        if (a.halo.nPeers && (d.hi.w & 2u)) {  // multi-GPU: this slice reads halo slots (halo.cuh)
            if (lane == 0) haloWaitAll(a.halo, haloEpochBase(a.halo) + static_cast<uint32_t>(a.halo.event));
            __syncwarp();
        }
        const uint32_t rows = n.lo.y - d.lo.y;
        double *P = tileOf(buf) + v;
        for (uint32_t k = sub; k < rows; k += G) {
            const double *src = a.x + 3 * static_cast<size_t>(ids[k * VW + v]);
            double *dst = P + k * (3 * VW);
            cpAsync8(dst, src);
            cpAsync8(dst + VW, src + 1);
            cpAsync8(dst + 2 * VW, src + 2);
        }
        const int slot = static_cast<int>(d.lo.x) + v;
        if (slot < end) {  // own state: components 0-2 x, 3-5 inertial x, 6 mass, spread over the vertex's G lanes
            const size_t o = 3 * static_cast<size_t>(slot);
            double *O = ownOf(buf) + v;
            for (int c = sub; c < 7; c += G) {
                const double *src = c < 3 ? a.x + o + c : (c < 6 ? a.xInertial + o + (c - 3) : a.mass + slot);
                cpAsync8(O + c * VW, src);
            }
        }
        cpAsyncCommit();
    };

    // ---- first slice of the warp: the same stages, serialised; A and B do not depend on the previous pass ----
    int slice = sliceBase + w;
    SliceDesc d, n;
    fetchDesc(slice);
    cpAsyncWaitAll();
    __syncwarp();
    readDesc(d, n);
    fetchIds(d, n);
    cpAsyncWaitAll();
    __syncwarp();
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (a.halo.nPeers && a.halo.pushWarps == 0 && slice == sliceBase && lane == 0) {
        // no boundary vertex of this colour here: the pass still owes its neighbours the event (vbd_stream.cuh)
        __threadfence_system();
        haloRaise(a.halo, haloEpochBase(a.halo) + static_cast<uint32_t>(a.halo.event) + 1u);
    }
    gather(d, n, 0);
    StreamRec<SMASK> r0, r1;
    r0.idx = r1.idx = 0xFFFFFFFFu;
    {
        const StreamPipe<SMASK, VW> first(d, n);
        if (first.total > 0) first.template loadFrom<0>(a, r0, 0, lane);
    }
    cpAsyncWaitAll();
    __syncwarp();

    int buf = 0;
    while (true) {
        const int slot = static_cast<int>(d.lo.x) + v;
        const bool valid = slot < end;
        const StreamPipe<SMASK, VW> pipe(d, n);
        const int next = slice + W;
        const bool hasNext = next - sliceBase < nSlicesOfColour;
        __syncwarp();                  // every lane has read the descriptors this slice came with
        if (hasNext) fetchDesc(next);  // A
        const double *P = tileOf(buf) + v;
        const double *O = ownOf(buf) + v;
        V3d xo(0.0, 0.0, 0.0);
        if (valid) xo = V3d(O[0], O[VW], O[2 * VW]);

        V3d f(0.0, 0.0, 0.0);
        Sym3 H{0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        // the descriptors of the next slice stay in shared memory (valid from stage B until the next stage A) and are
        // re-read where needed instead of living in 16 registers across the evaluators
        int stage = hasNext ? 0 : 3;
        auto advance = [&]() {
            cpAsyncWaitAll();
            __syncwarp();
            SliceDesc dn, nn;
            readDesc(dn, nn);
            if (stage == 0)
                fetchIds(dn, nn);  // B
            else
                gather(dn, nn, buf ^ 1);  // C
            ++stage;
        };
        // incidence streams of slice i (same two-deep record pipeline as vbdStreamKernel)
        for (uint32_t s = 0; s < pipe.total; s += 2) {
            if (s >= 2 && stage < 2) advance();
            int l = orderAfter(lane, r0.idx);
            if (s + 1 < pipe.total) pipe.template loadFrom<0>(a, r1, s + 1, l);
            pipe.template evalFrom<0>(r0, s, P + (l - lane), xo, f, H);
            if (s + 1 >= pipe.total) break;
            l = orderAfter(lane, r1.idx);
            if (s + 2 < pipe.total) pipe.template loadFrom<0>(a, r0, s + 2, l);
            pipe.template evalFrom<0>(r1, s + 1, P + (l - lane), xo, f, H);
        }
        while (stage < 2) advance();  // short slices: whatever is left of B / C
        if ((d.hi.w & 1u) && valid) accumulateCsrEnergies<KMASK>(a, slot, sub, G, f, H);
        // D: the first record of the next slice, requested ahead of this slice's reduction and solve
        r0.idx = r1.idx = 0xFFFFFFFFu;
        if (hasNext) {
            SliceDesc dn, nn;
            readDesc(dn, nn);
            const StreamPipe<SMASK, VW> np(dn, nn);
            if (np.total > 0) np.template loadFrom<0>(a, r0, 0, lane);
        } else if (a.lookahead > 0 && lane == 0) {
            // last slice of this warp: warm L2 with what the warp that takes over its position in the NEXT colour pass
            // (slices are numbered across colours; past the last colour the next sweep starts over) reads first
            int la = slice + W;
            if (la >= a.nSlices) la %= a.nSlices;
            SliceDesc pd, pn;
            pd.load(a.sliceDesc, la);
            pn.load(a.sliceDesc, la + 1);
            bulkPrefetchL2(a.nbrIds + static_cast<size_t>(pd.lo.y) * VW, (pn.lo.y - pd.lo.y) * (VW * 4u));
            prefetchStreamIfPresent<SMASK, ST_BEND_EDGE>(a, pd, pn);
            prefetchStreamIfPresent<SMASK, ST_BEND_OPP>(a, pd, pn);
            prefetchStreamIfPresent<SMASK, ST_MEMBRANE_STVK>(a, pd, pn);
            prefetchStreamIfPresent<SMASK, ST_TET_SNH>(a, pd, pn);
            prefetchStreamIfPresent<SMASK, ST_TET_STVK>(a, pd, pn);
        }
        double mass = 0.0;
        V3d xi(0.0, 0.0, 0.0);
        if (valid && sub == 0) {
            xi = V3d(O[3 * VW], O[4 * VW], O[5 * VW]);
            mass = O[6 * VW];
        }
        const V3d xn = reduceSolveWriteWith<G>(a, slot, valid, sub, xo, f, H, invdtSquared, mass, xi);
        if (a.halo.nPeers) {
            const uint32_t pb = d.hi.w >> 2, pe = n.hi.w >> 2;
            if (pe > pb) haloPushSlice(a.halo, d.lo.x, pb, pe, lane, xn);
        }
        if (!hasNext) break;
        cpAsyncWaitAll();
        __syncwarp();
        readDesc(d, n);
        slice = next;
        buf ^= 1;
    }
}

template <unsigned SMASK, unsigned KMASK, int G, int OCC>
void launchStreamPersistGO(const VbdArgs &args, int begin, int end, int sliceBase, double invdtSquared, cudaStream_t stream)
{
    constexpr int VW = 32 / G;
    constexpr int warpsPerBlock = kStreamBlockThreads / 32;
    const int n = end - begin;
    if (n <= 0) return;
    const int slices = (n + VW - 1) / VW;
    const PersistLayout L(args.maxNbr, VW);
    const size_t smem = warpsPerBlock * L.bytesPerWarp();
    if (smem > 227 * 1024) return;  // rejected at finalize
    auto kernel = vbdStreamPersistKernel<SMASK, KMASK, G, OCC>;
    static size_t configured = 0;  // per instantiation
    static int blocksPerSm = 0, smCount = 0;
    if (smem > configured) {
        cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        int pct = args.carveout;  // as in launchStreamGO: shared memory for the resident blocks, the rest stays L1
        if (pct < 0) pct = static_cast<int>(std::min<size_t>(100, (OCC * (smem + 1024) * 100 + 228 * 1024 - 1) / (228 * 1024) + 2));
        cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
        configured = smem;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocksPerSm, kernel, kStreamBlockThreads, smem);
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&smCount, cudaDevAttrMultiProcessorCount, dev);
        if (blocksPerSm < 1) blocksPerSm = 1;
    }
    // one resident wave: every warp walks its slices with stride W
    const int capacity = smCount * blocksPerSm;
    const int blocksNeeded = (slices + warpsPerBlock - 1) / warpsPerBlock;
    const int blocks = blocksNeeded < capacity ? blocksNeeded : capacity;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(blocks);
    cfg.blockDim = dim3(kStreamBlockThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = args.pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, args, end, sliceBase, slices, invdtSquared);
}
}  // namespace spgx
