// Streaming Vertex Block Descent colour pass (reference: spg/source/sim/solver/vbd.cpp:99-157).
//
// A group of G lanes owns one vertex of the colour (VW = 32 / G vertices per warp).  Per warp:
//   1. the unique-neighbour positions of its VW vertices are gathered ONCE into a shared-memory tile with
//      cp.async (LDGSTS: no registers held while the L2 gathers are in flight), laid out [neighbour][xyz][vertex]
//      so that every later read is bank-conflict free;
//   2. each incidence stream (streams.cuh) of the slice is read fully coalesced (one 128 B / 256 B line per field
//      per warp), two steps ahead in registers; the canonical evaluator picks its neighbour positions out of the
//      tile by 8-bit local index and accumulates dE/dx and the 3x3 block d2E/dx2;
//   3. energies without a stream (anchors, and any kind without a hand-derived block) run through the generic
//      CSR path with second-order duals seeded on the owning vertex;
//   4. the G partial sums are reduced with warp shuffles; lane group 0 adds the inertia term, applies the
//      det(H) > 1e-10 gate (vbd.cpp:139) and the cofactor solve, and writes the vertex.
#pragma once
#include <algorithm>
#include "halo.cuh"
#include "kernels.h"
#include "streams.cuh"

namespace spgx
{
using D2 = Dual<3, true>;

__device__ __forceinline__ void cpAsync8(double *smemDst, const double *gmemSrc)
{
    const unsigned s = static_cast<unsigned>(__cvta_generic_to_shared(smemDst));
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s), "l"(gmemSrc));
}
__device__ __forceinline__ void cpAsync4(uint32_t *smemDst, const uint32_t *gmemSrc)
{
    const unsigned s = static_cast<unsigned>(__cvta_generic_to_shared(smemDst));
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(gmemSrc));
}
__device__ __forceinline__ void cpAsyncCommit()
{
    asm volatile("cp.async.commit_group;");
}
__device__ __forceinline__ void cpAsyncWaitAll()
{
    asm volatile("cp.async.wait_all;" ::: "memory");
}
// every group but the newest one has landed
__device__ __forceinline__ void cpAsyncWaitAllButOne()
{
    asm volatile("cp.async.wait_group 1;" ::: "memory");
}

// generic CSR path: lanes of the vertex stride over its (stencil, local) entries of one energy
template <int KIND>
__device__ __forceinline__ void accumulateKind(const VbdEnergyDev &en, const double *__restrict__ x, int slot, int sub,
                                               int G, V3d &g, Sym3 &H)
{
    using EK = EnergyKind<KIND>;
    constexpr int S = EK::S;
    const uint32_t rb = en.rowptr[slot], re = en.rowptr[slot + 1];
    for (uint32_t q = rb + sub; q < re; q += G) {
        const uint32_t ent = en.entries[q];
        const uint32_t st = ent >> 2;
        const int local = ent & 3u;
        const int32_t *ids = en.stencils + static_cast<size_t>(st) * S;
        const double *rec = en.records + static_cast<size_t>(st) * EK::NP;
        if constexpr (HasHandDerived<EK>::value) {
            V3d xs[S];
#pragma unroll
            for (int s = 0; s < S; ++s) {
                const size_t o = 3 * static_cast<size_t>(ids[s]);
                xs[s] = V3d(x[o], x[o + 1], x[o + 2]);
            }
            EK::vertexGradHess(xs, rec, local, g, H);
        } else {
            V3<D2> xs[S];
#pragma unroll
            for (int s = 0; s < S; ++s) {
                const size_t o = 3 * static_cast<size_t>(ids[s]);
                const bool own = (s == local);
                xs[s] = V3<D2>(D2(x[o], own ? 0 : -1), D2(x[o + 1], own ? 1 : -1), D2(x[o + 2], own ? 2 : -1));
            }
            const D2 E = stencilEnergy<KIND, D2>(xs, rec);
            g.x += E.g[0];
            g.y += E.g[1];
            g.z += E.g[2];
            H.xx += E.h[0];
            H.xy += E.h[1];
            H.xz += E.h[2];
            H.yy += E.h[3];
            H.yz += E.h[4];
            H.zz += E.h[5];
        }
    }
}

template <unsigned KMASK, int KIND>
__device__ __forceinline__ void accumulateKindIfPresent(const VbdEnergyDev &en, const double *__restrict__ x, int slot,
                                                        int sub, int G, V3d &g, Sym3 &H)
{
    if constexpr ((KMASK >> KIND) & 1u) {
        if (en.kind == KIND) accumulateKind<KIND>(en, x, slot, sub, G, g, H);
    }
}

template <unsigned KMASK>
__device__ __forceinline__ void accumulateCsrEnergies(const VbdArgs &a, int slot, int sub, int G, V3d &g, Sym3 &H)
{
    for (int e = 0; e < a.nEnergies; ++e) {
        const VbdEnergyDev &en = a.e[e];
        if (en.streamed) continue;
#define SPGX_KIND(NAME, ID, BYTES) accumulateKindIfPresent<KMASK, NAME>(en, a.x, slot, sub, G, g, H);
#include "kinds.def"
#undef SPGX_KIND
    }
}

// butterfly over the sub-lane bits (lanes of one vertex sit VW apart), inertia term, gate, solve, write
// returns the position the vertex ends the pass with (meaningful on its sub == 0 lane)
// mass, xi: the vertex's mass and inertial position (read by the caller; used on the valid sub == 0 lane only)
template <int G>
__device__ __forceinline__ V3d reduceSolveWriteWith(const VbdArgs &a, int slot, bool valid, int sub, const V3d &xo, V3d f,
                                                    Sym3 H, double invdtSquared, double mass, const V3d &xi)
{
    V3d xn = xo;
    constexpr int VW = 32 / G;
#pragma unroll
    for (int o = VW; o < 32; o <<= 1) {
        f.x += __shfl_xor_sync(0xffffffffu, f.x, o);
        f.y += __shfl_xor_sync(0xffffffffu, f.y, o);
        f.z += __shfl_xor_sync(0xffffffffu, f.z, o);
        H.xx += __shfl_xor_sync(0xffffffffu, H.xx, o);
        H.xy += __shfl_xor_sync(0xffffffffu, H.xy, o);
        H.xz += __shfl_xor_sync(0xffffffffu, H.xz, o);
        H.yy += __shfl_xor_sync(0xffffffffu, H.yy, o);
        H.yz += __shfl_xor_sync(0xffffffffu, H.yz, o);
        H.zz += __shfl_xor_sync(0xffffffffu, H.zz, o);
    }
    if (valid && sub == 0) {
        const size_t o = 3 * static_cast<size_t>(slot);
        const double mOverdtSquared = mass * invdtSquared;
        // f holds the accumulated energy gradient: force = -gradient; inertia term of vbd.cpp:109-112, :123
        f.x = -f.x - mOverdtSquared * (xo.x - xi.x);
        f.y = -f.y - mOverdtSquared * (xo.y - xi.y);
        f.z = -f.z - mOverdtSquared * (xo.z - xi.z);
        H.xx += mOverdtSquared;
        H.yy += mOverdtSquared;
        H.zz += mOverdtSquared;
        const double d = det(H);
        if (d > 1e-10) {  // vbd.cpp:139
            const V3d dx = solveSym3(H, d, f);
            xn = V3d(xo.x + dx.x, xo.y + dx.y, xo.z + dx.z);
            a.x[o] = xn.x;
            a.x[o + 1] = xn.y;
            a.x[o + 2] = xn.z;
        }
    }
    return xn;
}

template <int G>
__device__ __forceinline__ V3d reduceSolveWrite(const VbdArgs &a, int slot, bool valid, int sub, const V3d &xo, const V3d &f,
                                                const Sym3 &H, double invdtSquared)
{
    double mass = 0.0;
    V3d xi(0.0, 0.0, 0.0);
    if (valid && sub == 0) {
        const size_t o = 3 * static_cast<size_t>(slot);
        mass = a.mass[slot];
        xi = V3d(a.xInertial[o], a.xInertial[o + 1], a.xInertial[o + 2]);
    }
    return reduceSolveWriteWith<G>(a, slot, valid, sub, xo, f, H, invdtSquared, mass, xi);
}

// Peer-memory halo push (kernels.h: VbdHaloDev): the slice's boundary vertices go straight into the neighbours' halo
// slots; the last warp of the pass to finish its stores raises the pass's event on every neighbour.
__device__ __forceinline__ void haloPushSlice(const VbdHaloDev &h, uint32_t firstSlot, uint32_t pb, uint32_t pe, int lane,
                                              const V3d &xn)
{
    for (uint32_t i = pb; i < pe; i += 32) {
        const bool has = i + lane < pe;
        uint2 e = make_uint2(firstSlot, 0u);
        if (has) e = __ldg(h.push + i + lane);
        const int src = static_cast<int>(e.x - firstSlot);  // the vertex's sub == 0 lane
        const double px = __shfl_sync(0xffffffffu, xn.x, src);
        const double py = __shfl_sync(0xffffffffu, xn.y, src);
        const double pz = __shfl_sync(0xffffffffu, xn.z, src);
        if (has && !(h.debug & 4)) {
            double *dst = h.peerX[e.y >> 28] + 3 * static_cast<size_t>(e.y & 0x0FFFFFFFu);
            dst[0] = px;
            dst[1] = py;
            dst[2] = pz;
        }
    }
    if (!(h.debug & 2)) __threadfence_system();
    __syncwarp();
    if (lane == 0) {
        const uint32_t done = atomicAdd(h.pushDone, 1u) + 1u;
        if (done == static_cast<uint32_t>(h.pushWarps)) {
            *reinterpret_cast<volatile uint32_t *>(h.pushDone) = 0u;  // the next pass counts from zero
            __threadfence_system();
            haloRaise(h, haloEpochBase(h) + static_cast<uint32_t>(h.event) + 1u);
        }
    }
}

struct SliceDesc {
    uint4 lo, hi;  // [slot, nbrRow, s0, s1], [s2, s3, s4, flags]
    __device__ __forceinline__ void load(const uint32_t *desc, int slice)
    {
        const uint4 *p = reinterpret_cast<const uint4 *>(desc) + 2 * static_cast<size_t>(slice);
        lo = __ldg(p);
        hi = __ldg(p + 1);
    }
    template <int ST>
    __device__ __forceinline__ uint32_t stream() const
    {
        constexpr int w = streamDescWord(ST);
        return w == 2 ? lo.z : w == 3 ? lo.w : w == 4 ? hi.x : w == 5 ? hi.y : hi.z;
    }
};

constexpr int maxStreamNQ(unsigned smask)
{
    int m = 1;
    if ((smask >> ST_BEND_EDGE) & 1u) m = Stream<ST_BEND_EDGE>::NQ > m ? Stream<ST_BEND_EDGE>::NQ : m;
    if ((smask >> ST_BEND_OPP) & 1u) m = Stream<ST_BEND_OPP>::NQ > m ? Stream<ST_BEND_OPP>::NQ : m;
    if ((smask >> ST_MEMBRANE_STVK) & 1u) m = Stream<ST_MEMBRANE_STVK>::NQ > m ? Stream<ST_MEMBRANE_STVK>::NQ : m;
    if ((smask >> ST_TET_SNH) & 1u) m = Stream<ST_TET_SNH>::NQ > m ? Stream<ST_TET_SNH>::NQ : m;
    if ((smask >> ST_TET_STVK) & 1u) m = Stream<ST_TET_STVK>::NQ > m ? Stream<ST_TET_STVK>::NQ : m;
    if ((smask >> ST_WEDGE) & 1u) m = Stream<ST_WEDGE>::NQ > m ? Stream<ST_WEDGE>::NQ : m;
    return m;
}

// one incidence record of any stream type of the kernel variant (register storage shared between the types)
template <unsigned SMASK>
struct StreamRec {
    uint32_t idx;
    double q[maxStreamNQ(SMASK)];
    template <int ST>
    __device__ __forceinline__ void load(const VbdStreamDev &sd, uint32_t t, int lane)
    {
        using SK = Stream<ST>;
        idx = __ldg(sd.idx + static_cast<size_t>(t) * 32 + lane);
        const double *qp = sd.params + static_cast<size_t>(t) * (SK::NQ * 32) + lane;
#pragma unroll
        for (int k = 0; k < SK::NQ; ++k) q[k] = __ldg(qp + k * 32);
    }
    // P: the warp's neighbour tile, already offset to this lane's vertex column; element (k, c) at P[(3 k + c) * VW]
    template <int ST, int VW>
    __device__ __forceinline__ void eval(const double *P, const V3d &xo, V3d &g, Sym3 &H) const
    {
        using SK = Stream<ST>;
        if (idx == 0xFFFFFFFFu) return;
        V3d xn[SK::NI];
#pragma unroll
        for (int k = 0; k < SK::NI; ++k) {
            const double *p = P + ((idx >> (8 * k)) & 0xFFu) * (3 * VW);
            xn[k] = V3d(p[0], p[VW], p[2 * VW]);
        }
        SK::eval(xo, xn, q, g, H);
    }
};

// The streams of one slice as ONE concatenated step sequence [0, total): step s belongs to the first type whose
// cumulative count exceeds s (a warp-uniform branch).  Running the two-deep register prefetch over the concatenated
// sequence keeps the record loads two steps ahead across stream boundaries as well, so only the first two loads of a
// warp are exposed (and those overlap the neighbour gather).
// position in the concatenated step sequence -> stream type: wedges first -- they carry nearly all of a cloth's steps, so the
// warp-uniform search for a step's type ends at its first comparison
constexpr int streamAt(int i)
{
    return i == 0 ? ST_WEDGE : i - 1;
}

template <unsigned SMASK, int VW>
struct StreamPipe {
    uint32_t base[kNumStreamTypes];  // first step of the type minus the concatenated index where the type begins
    uint32_t cum[kNumStreamTypes];   // concatenated index one past the type's last step
    uint32_t total;

    template <int I>
    __device__ __forceinline__ void init(const SliceDesc &d, const SliceDesc &n, uint32_t &acc)
    {
        constexpr int ST = streamAt(I);
        if constexpr ((SMASK >> ST) & 1u) {
            const uint32_t t0 = d.stream<ST>(), t1 = n.stream<ST>();
            base[ST] = t0 - acc;
            acc += t1 - t0;
            cum[ST] = acc;
        }
    }
    __device__ __forceinline__ StreamPipe(const SliceDesc &d, const SliceDesc &n)
    {
        uint32_t acc = 0;
        init<0>(d, n, acc);
        init<1>(d, n, acc);
        init<2>(d, n, acc);
        init<3>(d, n, acc);
        init<4>(d, n, acc);
        init<5>(d, n, acc);
        total = acc;
    }
    template <int I>
    __device__ __forceinline__ void loadFrom(const VbdArgs &a, StreamRec<SMASK> &r, uint32_t s, int lane) const
    {
        if constexpr (I < kNumStreamTypes) {
            constexpr int ST = streamAt(I);
            if constexpr ((SMASK >> ST) & 1u) {
                if (s < cum[ST]) {
                    r.template load<ST>(a.s[ST], base[ST] + s, lane);
                    return;
                }
            }
            loadFrom<I + 1>(a, r, s, lane);
        }
    }
    template <int I>
    __device__ __forceinline__ void evalFrom(const StreamRec<SMASK> &r, uint32_t s, const double *P, const V3d &xo, V3d &g,
                                             Sym3 &H) const
    {
        if constexpr (I < kNumStreamTypes) {
            constexpr int ST = streamAt(I);
            if constexpr ((SMASK >> ST) & 1u) {
                if (s < cum[ST]) {
                    r.template eval<ST, VW>(P, xo, g, H);
                    return;
                }
            }
            evalFrom<I + 1>(r, s, P, xo, g, H);
        }
    }
};

// returns `lane`, through an instruction that reads `loaded`: whatever is computed from the result is ordered after the
// arrival of the load that produced `loaded` (a record index never equals 0xFFFFFFFE, so the value is always `lane`)
__device__ __forceinline__ int orderAfter(int lane, uint32_t loaded)
{
    asm volatile("{ .reg .pred p; setp.eq.u32 p, %1, 0xFFFFFFFE; selp.s32 %0, -1, %0, p; }" : "+r"(lane) : "r"(loaded));
    return lane;
}

// pull [ptr, ptr + bytes) into L2 with one TMA bulk prefetch (bytes: multiple of 16, ptr 16-byte aligned)
__device__ __forceinline__ void bulkPrefetchL2(const void *ptr, uint32_t bytes)
{
    if (bytes) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(ptr), "r"(bytes));
}
__device__ __forceinline__ void prefetchLineL2(const void *ptr)
{
    asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr));
}


template <unsigned SMASK, int ST>
__device__ __forceinline__ void prefetchStreamIfPresent(const VbdArgs &a, const SliceDesc &d, const SliceDesc &n)
{
    if constexpr ((SMASK >> ST) & 1u) {
        if (a.s[ST].idx) {
            const uint32_t t = d.stream<ST>(), steps = n.stream<ST>() - t;
            bulkPrefetchL2(a.s[ST].idx + static_cast<size_t>(t) * 32, steps * 128u);
            bulkPrefetchL2(a.s[ST].params + static_cast<size_t>(t) * (Stream<ST>::NQ * 32), steps * (Stream<ST>::NQ * 256u));
        }
    }
}

constexpr int kStreamBlockThreads = 128;

// ---- ring path: the record of the next step staged in shared memory ----
// One step of the ring: the 32 index words, then NQmax rows of 32 doubles ([field][lane], as in global memory).
template <unsigned SMASK>
constexpr int ringStepBytes()
{
    return 128 + maxStreamNQ(SMASK) * 256;
}
// bytes of shared memory one warp owns: the neighbour tile (maxNbr rows + own position + inertial position + mass row) and,
// with RING, two ring steps
template <unsigned SMASK, int G, bool RING>
__host__ __device__ inline size_t streamWarpSmemBytes(int maxNbr)
{
    constexpr int VW = 32 / G;
    return static_cast<size_t>(maxNbr + 3) * 3 * VW * sizeof(double) + (RING ? 2 * ringStepBytes<SMASK>() : 0);
}

template <unsigned SMASK, int VW>
struct RingOps {
    // lane copies ITS OWN fields of step t of stream ST into ring step `stage`: nothing another lane wrote is ever read, so
    // cp.async.wait_group alone orders the copies against the reads (no barrier, no __syncwarp); no register is held while
    // the record is in flight, and none of its fields lives longer than the instruction that consumes it
    template <int ST>
    __device__ static __forceinline__ void issueStep(const VbdStreamDev &sd, uint32_t t, char *stage, int lane)
    {
        using SK = Stream<ST>;
        cpAsync4(reinterpret_cast<uint32_t *>(stage) + lane, sd.idx + static_cast<size_t>(t) * 32 + lane);
        const double *qp = sd.params + static_cast<size_t>(t) * (SK::NQ * 32) + lane;
        double *dst = reinterpret_cast<double *>(stage + 128) + lane;
#pragma unroll
        for (int k = 0; k < SK::NQ; ++k) cpAsync8(dst + k * 32, qp + k * 32);
    }
    template <int I>
    __device__ static __forceinline__ void issueFrom(const VbdArgs &a, const StreamPipe<SMASK, VW> &pipe, uint32_t s, char *stage, int lane)
    {
        if constexpr (I < kNumStreamTypes) {
            constexpr int ST = streamAt(I);
            if constexpr ((SMASK >> ST) & 1u) {
                if (s < pipe.cum[ST]) {
                    issueStep<ST>(a.s[ST], pipe.base[ST] + s, stage, lane);
                    return;
                }
            }
            issueFrom<I + 1>(a, pipe, s, stage, lane);
        }
    }
    template <int ST>
    __device__ static __forceinline__ void evalStep(const char *stage, const double *P, const double *Pown, V3d &g, Sym3 &H, int lane)
    {
        using SK = Stream<ST>;
        const uint32_t idx = reinterpret_cast<const uint32_t *>(stage)[lane];
        if (idx == 0xFFFFFFFFu) return;
        const double *qs = reinterpret_cast<const double *>(stage + 128) + lane;
        double q[SK::NQ];
        V3d xn[SK::NI];
#pragma unroll
        for (int k = 0; k < SK::NQ; ++k) q[k] = qs[k * 32];
#pragma unroll
        for (int k = 0; k < SK::NI; ++k) {
            const double *p = P + ((idx >> (8 * k)) & 0xFFu) * (3 * VW);
            xn[k] = V3d(p[0], p[VW], p[2 * VW]);
        }
        const V3d xo(Pown[0], Pown[VW], Pown[2 * VW]);  // re-read per step: six registers fewer across the loop
        SK::eval(xo, xn, q, g, H);
    }
    template <int I>
    __device__ static __forceinline__ void evalFrom(const StreamPipe<SMASK, VW> &pipe, uint32_t s, const char *stage, const double *P,
                                                    const double *Pown, V3d &g, Sym3 &H, int lane)
    {
        if constexpr (I < kNumStreamTypes) {
            constexpr int ST = streamAt(I);
            if constexpr ((SMASK >> ST) & 1u) {
                if (s < pipe.cum[ST]) {
                    evalStep<ST>(stage, P, Pown, g, H, lane);
                    return;
                }
            }
            evalFrom<I + 1>(pipe, s, stage, P, Pown, g, H, lane);
        }
    }
};

// The body of the colour pass; BT = threads per block.  Warps never synchronise with each other (each has its own tile),
// so the block size only sets the granularity of the occupancy / register trade-off (registers are bounded by OCC resident
// blocks of BT threads).
// RING: the record of the next step is prefetched into a two-step shared-memory ring with cp.async (no registers held, no
// scoreboard wait) instead of the two-buffer register pipeline: ~30 registers fewer, which is what a fifth / sixth resident
// block per SM needs.
// NBRFIXED: the unique-neighbour table has a fixed stride of maxNbr rows per slice (padded with slot 0), so its address
// follows from the slice number alone and the ids are requested TOGETHER with the descriptor instead of after it: one L2
// round trip less in the serial prologue descriptor -> neighbour ids -> gather (profiles/r2j_vbd_stream_stalls.txt: that
// chain is most of the 28 % long-scoreboard share).
template <unsigned SMASK, unsigned KMASK, int G, int BT, bool RING, bool NBRFIXED>
__device__ __forceinline__ void vbdStreamBody(const VbdArgs &a, int end, int sliceBase, int nSlicesOfColour, double invdtSquared)
{
    constexpr int VW = 32 / G;  // vertices per warp
    extern __shared__ __align__(128) char smemRaw[];
    const int warpInBlock = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int warp = blockIdx.x * (BT / 32) + warpInBlock;
    asm volatile("griddepcontrol.launch_dependents;");  // the next colour pass may start its prologue
    if (warp >= nSlicesOfColour) return;  // whole warp past the colour
    const int v = lane % VW, sub = lane / VW;
    const int slice = sliceBase + warp;
    // per warp: [ring (RING) | maxNbr neighbour rows | own positions | inertial positions | masses]
    char *warpSmem = smemRaw + static_cast<size_t>(warpInBlock) * streamWarpSmemBytes<SMASK, G, RING>(a.maxNbr);
    char *ring = warpSmem;
    double *P = reinterpret_cast<double *>(warpSmem + (RING ? 2 * ringStepBytes<SMASK>() : 0)) + v;
    double *Pown = P + a.maxNbr * (3 * VW);
    double *Pin = Pown + 3 * VW;
    double *Pmass = Pin + 3 * VW;

    constexpr int PRE = G == 1 ? 12 : (G == 2 ? 8 : 4);  // neighbour ids per lane fetched ahead of the dependency wait
    uint32_t nid[PRE];
    if constexpr (NBRFIXED) {
#pragma unroll
        for (int j = 0; j < PRE; ++j) {
            const int k = sub + j * G;  // rows past the slice's own count hold slot 0
            nid[j] = k < a.maxNbr ? __ldg(a.nbrIds + (static_cast<size_t>(slice) * a.maxNbr + k) * VW + v) : 0u;
        }
    }
    SliceDesc d, n;
    d.load(a.sliceDesc, slice);
    n.load(a.sliceDesc, slice + 1);
    const int slot = static_cast<int>(d.lo.x) + v;
    const bool valid = slot < end;

    // 1. everything that does not depend on the previous colour pass: the first incidence record(s), the ids of the
    // unique neighbours, the vertex's mass and inertial position, and the streams of the slice `lookahead` slices further
    // on (wrapping into the next colour / the next sweep) pulled from HBM into L2 with TMA bulk prefetches
    const StreamPipe<SMASK, VW> pipe(d, n);
    StreamRec<SMASK> r0, r1;
    r0.idx = r1.idx = 0xFFFFFFFFu;
    if (valid && sub == 0) {
        const double *src = a.xInertial + 3 * static_cast<size_t>(slot);
        cpAsync8(Pin, src);
        cpAsync8(Pin + VW, src + 1);
        cpAsync8(Pin + 2 * VW, src + 2);
        cpAsync8(Pmass, a.mass + slot);
    }
    if constexpr (RING) {
        if (pipe.total > 0) RingOps<SMASK, VW>::template issueFrom<0>(a, pipe, 0, ring, lane);
    } else {
        if (pipe.total > 0) pipe.template loadFrom<0>(a, r0, 0, lane);
    }
    cpAsyncCommit();
    // rows [nb, ne) of the neighbour table are the slice's (descriptor word 1 counts rows cumulatively in both layouts)
    const uint32_t nb = NBRFIXED ? static_cast<uint32_t>(slice) * static_cast<uint32_t>(a.maxNbr) : d.lo.y;
    const uint32_t ne = nb + (n.lo.y - d.lo.y);
    if constexpr (!NBRFIXED) {
#pragma unroll
        for (int j = 0; j < PRE; ++j) {
            const uint32_t k = nb + sub + j * G;
            nid[j] = k < ne ? __ldg(a.nbrIds + static_cast<size_t>(k) * VW + v) : 0u;
        }
    }
    if (a.lookahead > 0 && warpInBlock == 0) {
        // ONE lane per block pulls the ranges of the block's BT / 32 consecutive slices, `lookahead` slices further on
        // (issued by every warp this cost 171 issue slots per warp-slice, 9 % of the colour pass: ncu source page)
        int la = slice + a.lookahead;
        if (la >= a.nSlices) la -= a.nSlices;
        const int lb = min(la + BT / 32, a.nSlices);
        if (lane == 0) {
            int la2 = la + a.lookahead;
            if (la2 >= a.nSlices) la2 -= a.nSlices;
            prefetchLineL2(a.sliceDesc + static_cast<size_t>(la2) * kSliceDescWords);
            SliceDesc pd, pn;
            pd.load(a.sliceDesc, la);
            pn.load(a.sliceDesc, lb);
            if constexpr (NBRFIXED)
                bulkPrefetchL2(a.nbrIds + static_cast<size_t>(la) * a.maxNbr * VW, static_cast<uint32_t>(lb - la) * a.maxNbr * (VW * 4u));
            else
                bulkPrefetchL2(a.nbrIds + static_cast<size_t>(pd.lo.y) * VW, (pn.lo.y - pd.lo.y) * (VW * 4u));
            prefetchStreamIfPresent<SMASK, ST_BEND_EDGE>(a, pd, pn);
            prefetchStreamIfPresent<SMASK, ST_BEND_OPP>(a, pd, pn);
            prefetchStreamIfPresent<SMASK, ST_MEMBRANE_STVK>(a, pd, pn);
            prefetchStreamIfPresent<SMASK, ST_TET_SNH>(a, pd, pn);
            prefetchStreamIfPresent<SMASK, ST_TET_STVK>(a, pd, pn);
            prefetchStreamIfPresent<SMASK, ST_WEDGE>(a, pd, pn);
        }
    }
    if (a.halo.nPeers && lane == 0) {
        // multi-GPU: the slice's push entries are read at the very end of the warp's life (haloPushSlice): have them in L2 by
        // then (profiles/r2m_halo_overhead_breakdown_2gpu.txt: the bookkeeping of the boundary slices, not the fences or the
        // remote stores, is the larger half of what the exchange costs)
        const uint32_t pb = d.hi.w >> 3, pe = n.hi.w >> 3;
        if (pe > pb) {
            prefetchLineL2(a.halo.push + pb);
            prefetchLineL2(a.halo.push + pe - 1);
        }
    }
    // 2. positions: wait for the previous colour pass (programmatic dependent launch: this grid may have started
    // while that one was draining), then gather the unique neighbours (columns of padding lanes hold slot 0)
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (a.halo.nPeers) {
        // multi-GPU: a slice that reads halo slots needs the neighbours' stores of the previous event
        if (a.halo.pushWarps == 0 && slice == sliceBase && lane == 0) {
            // no boundary vertex of this colour here: the pass still owes its neighbours the event (the previous
            // pass, stores included, is complete at this point)
            __threadfence_system();
            haloRaise(a.halo, haloEpochBase(a.halo) + static_cast<uint32_t>(a.halo.event) + 1u);
        }
        if (d.hi.w & 2u) {
            if (lane == 0 && !(a.halo.debug & 1)) haloWaitAll(a.halo, haloEpochBase(a.halo) + static_cast<uint32_t>(a.halo.event));
            __syncwarp();
        }
    }
#pragma unroll
    for (int j = 0; j < PRE; ++j) {
        const uint32_t k = nb + sub + j * G;
        if (k < ne) {
            const double *src = a.x + 3 * static_cast<size_t>(nid[j]);
            double *dst = P + (k - nb) * (3 * VW);
            cpAsync8(dst, src);
            cpAsync8(dst + VW, src + 1);
            cpAsync8(dst + 2 * VW, src + 2);
        }
    }
    for (uint32_t k = nb + sub + PRE * G; k < ne; k += G) {
        const uint32_t id = __ldg(a.nbrIds + static_cast<size_t>(k) * VW + v);
        const double *src = a.x + 3 * static_cast<size_t>(id);
        double *dst = P + (k - nb) * (3 * VW);
        cpAsync8(dst, src);
        cpAsync8(dst + VW, src + 1);
        cpAsync8(dst + 2 * VW, src + 2);
    }
    // the vertex's own position rides in the same asynchronous batch (a plain load here gets scheduled BEHIND the wait
    // for the gather and costs a second, fully exposed L2 round trip: profiles/r1f_*); so do its mass and inertial
    // position (first group above), which the epilogue used to fetch with exposed loads after the incidence loop
    if (valid && sub == 0) {
        const double *src = a.x + 3 * static_cast<size_t>(slot);
        cpAsync8(Pown, src);
        cpAsync8(Pown + VW, src + 1);
        cpAsync8(Pown + 2 * VW, src + 2);
    }
    cpAsyncCommit();
    V3d f(0.0, 0.0, 0.0);
    Sym3 H{0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    V3d xo(0.0, 0.0, 0.0);
    if constexpr (RING) {
        // 3a. ring path.  Groups in flight: [own data + step 0] [gather]; every iteration commits exactly one more group
        // (step s + 1, possibly empty) and waits for all but that one.
        for (uint32_t s = 0; s < pipe.total; ++s) {
            char *cur = ring + (s & 1u) * ringStepBytes<SMASK>();
            char *nxt = ring + ((s + 1u) & 1u) * ringStepBytes<SMASK>();
            if (s + 1 < pipe.total) RingOps<SMASK, VW>::template issueFrom<0>(a, pipe, s + 1, nxt, lane);
            cpAsyncCommit();
            cpAsyncWaitAllButOne();
            if constexpr (G > 1) {
                if (s == 0) __syncwarp();  // the tile rows were written by the other lanes of the vertex
            }
            RingOps<SMASK, VW>::template evalFrom<0>(pipe, s, cur, P, Pown, f, H, lane);
        }
        cpAsyncWaitAll();
        if constexpr (G > 1) __syncwarp();
        if (valid) xo = V3d(Pown[0], Pown[VW], Pown[2 * VW]);
    } else {
        cpAsyncWaitAll();
        if constexpr (G > 1) __syncwarp();
        if (valid) xo = V3d(Pown[0], Pown[VW], Pown[2 * VW]);
        // 3b. register pipeline.  All record loads share one scoreboard barrier (ptxas), so a wait for the current record
        // also waits for whatever record load was issued last: the next record is therefore requested right AFTER the
        // current one has arrived (orderAfter) and right BEFORE it is evaluated, which puts one whole evaluation
        // between every request and its wait.
        for (uint32_t s = 0; s < pipe.total; s += 2) {
            // (the tile pointer is routed through orderAfter too, so that the wait dominates the evaluation on every path)
            int l = orderAfter(lane, r0.idx);
            if (s + 1 < pipe.total) pipe.template loadFrom<0>(a, r1, s + 1, l);
            pipe.template evalFrom<0>(r0, s, P + (l - lane), xo, f, H);
            if (s + 1 >= pipe.total) break;
            l = orderAfter(lane, r1.idx);
            if (s + 2 < pipe.total) pipe.template loadFrom<0>(a, r0, s + 2, l);
            pipe.template evalFrom<0>(r1, s + 1, P + (l - lane), xo, f, H);
        }
    }
    // 4. energies without a stream (rare: the descriptor says whether the slice has any such incidence)
    if ((d.hi.w & 1u) && valid) accumulateCsrEnergies<KMASK>(a, slot, sub, G, f, H);
    // 5. reduce, solve, write
    double mass = 0.0;
    V3d xi(0.0, 0.0, 0.0);
    if (valid && sub == 0) {
        mass = Pmass[0];
        xi = V3d(Pin[0], Pin[VW], Pin[2 * VW]);
    }
    const V3d xn = reduceSolveWriteWith<G>(a, slot, valid, sub, xo, f, H, invdtSquared, mass, xi);
    // 6. multi-GPU: boundary vertices of the slice go to the neighbours' halo slots over NVLink
    if (a.halo.nPeers) {
        const uint32_t pb = d.hi.w >> 3, pe = n.hi.w >> 3;
        if (pe > pb && !(a.halo.debug & 16)) haloPushSlice(a.halo, d.lo.x, pb, pe, lane, xn);
    }
}

template <unsigned SMASK, unsigned KMASK, int G, int OCC, bool RING, bool NBRFIXED, int BT = kStreamBlockThreads>
__global__ void __launch_bounds__(BT, OCC)
    vbdStreamKernel(const __grid_constant__ VbdArgs a, int end, int sliceBase, int nSlicesOfColour, double invdtSquared)
{
    vbdStreamBody<SMASK, KMASK, G, BT, RING, NBRFIXED>(a, end, sliceBase, nSlicesOfColour, invdtSquared);
}
template <unsigned SMASK, unsigned KMASK, int G, int OCC, bool RING, bool NBRFIXED, int BT = kStreamBlockThreads>
void launchStreamGO(const VbdArgs &args, int begin, int end, int sliceBase, double invdtSquared, cudaStream_t stream)
{
    constexpr int VW = 32 / G;
    const int n = end - begin;
    if (n <= 0) return;
    const int warps = (n + VW - 1) / VW;
    constexpr int warpsPerBlock = BT / 32;
    const int blocks = (warps + warpsPerBlock - 1) / warpsPerBlock;
    const size_t smem = static_cast<size_t>(warpsPerBlock) * streamWarpSmemBytes<SMASK, G, RING>(args.maxNbr);
    if (smem > 227 * 1024) return;  // rejected at finalize (buildStreams)
    auto kernel = vbdStreamKernel<SMASK, KMASK, G, OCC, RING, NBRFIXED, BT>;
    static size_t configured = 0;  // per instantiation
    if (smem > configured) {
        cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        // shared memory for exactly OCC resident blocks (+ 1 KB per block the runtime reserves), the rest of the 228 KB
        // stays L1: the neighbour gather and the record streams go through it (measured on config 3: 3.08 -> 2.95 ms per
        // frame against the maximum-shared carveout)
        int pct = args.carveout;
        if (pct < 0) pct = static_cast<int>(std::min<size_t>(100, (OCC * (smem + 1024) * 100 + 228 * 1024 - 1) / (228 * 1024) + 2));
        cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
        configured = smem;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(blocks);
    cfg.blockDim = dim3(BT);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (args.pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    if (args.l2PersistBytes) {
        // the positions every gather reads stay in the L2 set-aside; the incidence streams (read once per sweep, 0.65 GB
        // against 126 MB of L2) pass through as streaming and stop evicting them
        attr[na].id = cudaLaunchAttributeAccessPolicyWindow;
        attr[na].val.accessPolicyWindow.base_ptr = args.x;
        attr[na].val.accessPolicyWindow.num_bytes = args.l2PersistBytes;
        attr[na].val.accessPolicyWindow.hitRatio = args.l2HitRatio;
        attr[na].val.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        attr[na].val.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    cudaLaunchKernelEx(&cfg, kernel, args, end, sliceBase, warps, invdtSquared);
}

// variant = occ + 10 * ring: resident 128-thread blocks per SM and whether the records go through the shared-memory ring
// instead of the register pipeline: 4 (register pipeline, <= 128 registers), 14 (ring, 4 blocks, <= 128 registers) and 15
// (ring, 5 blocks, <= 102 registers) are compiled.  A cloth with wedge records (streams.cuh: ST_WEDGE, 11 constants a record)
// runs 14: two such records in flight in registers spill (164 bytes), and so does the wedge evaluator under 102 registers;
// config 3, profiles/r2s_sweep.txt: 14 2.32 ms per frame, 15 2.95, 4 3.37; the ring with 3 / 2 resident blocks (158 / 167
// registers, what ptxas takes without a cap) 2.44 / 3.03 ms: measured, not compiled.  Everything else runs 4.  Measured and rejected, no longer compiled: one lane per vertex, ring with 4 or 6 blocks
// (profiles/r2d ... r2i_sweep.txt); 3 blocks, register caps on the register pipeline, persistent warps, prefetch.global.L1
// (round 1, profiles/r1f_*); two steps evaluated together from a four-step ring (ptxas keeps the two evaluations sequential);
// the register pipeline in 2-warp blocks, 8 resident (2.748 vs 2.702 ms on config 3).
template <unsigned SMASK, unsigned KMASK, int G>
void launchStreamG(const VbdArgs &args, int variant, int begin, int end, int sliceBase, double invdtSquared,
                   cudaStream_t stream)
{
    if (args.nbrFixed) {
        if (variant == 15)
            launchStreamGO<SMASK, KMASK, G, 5, true, true>(args, begin, end, sliceBase, invdtSquared, stream);
        else if (variant == 14)
            launchStreamGO<SMASK, KMASK, G, 4, true, true>(args, begin, end, sliceBase, invdtSquared, stream);
        else
            launchStreamGO<SMASK, KMASK, G, 4, false, true>(args, begin, end, sliceBase, invdtSquared, stream);
    } else {
        if (variant == 15)
            launchStreamGO<SMASK, KMASK, G, 5, true, false>(args, begin, end, sliceBase, invdtSquared, stream);
        else if (variant == 14)
            launchStreamGO<SMASK, KMASK, G, 4, true, false>(args, begin, end, sliceBase, invdtSquared, stream);
        else
            launchStreamGO<SMASK, KMASK, G, 4, false, false>(args, begin, end, sliceBase, invdtSquared, stream);
    }
}

template <unsigned SMASK, unsigned KMASK>
void launchStream(const VbdArgs &args, int lanes, int occ, int begin, int end, int sliceBase, double invdtSquared,
                  cudaStream_t stream)
{
    if (lanes == 4)
        launchStreamG<SMASK, KMASK, 4>(args, occ, begin, end, sliceBase, invdtSquared, stream);
    else
        launchStreamG<SMASK, KMASK, 2>(args, occ, begin, end, sliceBase, invdtSquared, stream);
}

constexpr unsigned sbit(int k)
{
    return 1u << k;
}
constexpr unsigned kStreamsClothSingle = sbit(ST_BEND_EDGE) | sbit(ST_BEND_OPP) | sbit(ST_MEMBRANE_STVK);
constexpr unsigned kStreamsCloth = kStreamsClothSingle | sbit(ST_WEDGE);
constexpr unsigned kStreamsTets = sbit(ST_TET_SNH) | sbit(ST_TET_STVK);
constexpr unsigned kKindsAnchor = sbit(K_ANCHOR);

// one translation unit per combination (compile time)
void launchStreamCloth(const VbdArgs &args, int lanes, int occ, int begin, int end, int sliceBase, double invdtSquared,
                       cudaStream_t stream);
// the same with the wedge evaluator compiled in (its registers would cost the single-incidence kernel its spill-free pipeline)
void launchStreamClothWedge(const VbdArgs &args, int lanes, int occ, int begin, int end, int sliceBase, double invdtSquared,
                            cudaStream_t stream);
void launchStreamTets(const VbdArgs &args, int lanes, int occ, int begin, int end, int sliceBase, double invdtSquared,
                      cudaStream_t stream);
}  // namespace spgx
