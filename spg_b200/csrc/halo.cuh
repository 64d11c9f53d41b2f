// Device side of the peer-memory halo exchange (kernels.h: VbdHaloDev; host side: comm.cu).
//
// Memory-model summary.  Producer (rank A, colour pass j): plain stores of the new positions into B's array, then
// fence.sys by every storing lane, __syncwarp, one relaxed atomic per warp on A's own `pushDone`; the warp that sees
// the full count issues fence.sys and st.release.sys of the event value into B's flag array.  Consumer (rank B,
// colour pass j + 1): lane 0 of every slice that reads halo slots spins on ld.acquire.sys of its own flag array
// (local memory: the peer's store lands in B's L2), then __syncwarp, then gathers.  The acquire also drops what L1
// holds, so a sector that straddles two colours of the halo range cannot be served stale.
#pragma once
#include <cstdint>

#include "kernels.h"

namespace spgx
{
__device__ __forceinline__ uint32_t ldAcquireSys(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void stReleaseSys(uint32_t *p, uint32_t v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long globalTimerNs()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ uint32_t haloEpochBase(const VbdHaloDev &h)
{
    return *reinterpret_cast<const volatile uint32_t *>(h.epochBase);
}

// one thread: block until every neighbour has raised its counter to `need` (wrap-safe); a stuck neighbour costs a
// bounded wait and sets the error word instead of hanging the device
__device__ __forceinline__ void haloWaitAll(const VbdHaloDev &h, uint32_t need)
{
    for (int p = 0; p < h.nPeers; ++p) {
        const uint32_t *f = h.myFlag + p;
        if (static_cast<int32_t>(ldAcquireSys(f) - need) >= 0) continue;
        const unsigned long long t0 = globalTimerNs();
        while (static_cast<int32_t>(ldAcquireSys(f) - need) < 0) {
            if (globalTimerNs() - t0 > h.timeoutNs) {
                *reinterpret_cast<volatile uint32_t *>(h.error) = 1u;
                break;
            }
            __nanosleep(40);
        }
    }
}

// one thread: everything this rank stored into its neighbours before the preceding fences is visible to whoever
// acquires the value
__device__ __forceinline__ void haloRaise(const VbdHaloDev &h, uint32_t value)
{
    for (int p = 0; p < h.nPeers; ++p) stReleaseSys(h.peerFlag[p], value);
}
}  // namespace spgx
