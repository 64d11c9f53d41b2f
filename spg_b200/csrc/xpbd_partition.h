// Domain decomposition plan for XPBD ("next"; reference: XPBD::step, spg/source/sim/solver/xpbd.cpp:16-131,
// which runs every (energy, stencil colour) pass over the whole mesh in one process).  Pure host code.
//
// A pass projects the stencils of one colour of one energy; stencils of a colour share no vertex (xpbd.cpp:133-160), so
// every vertex is written by at most one stencil per pass.  With the vertices split between ranks by `owner`:
//   * a rank HOLDS every stencil that touches a vertex it owns, and keeps halo copies of those stencils' other vertices;
//   * a stencil cut by the partition is executed by EVERY rank that holds it, on identical inputs with identical
//     arithmetic, so each of them gets the same new positions for all of the stencil's vertices: nothing is sent back
//     to the owners (the two-way exchange SURVEY 8e sketches is not needed);
//   * what a rank cannot compute itself is the effect of stencils it does not hold on its halo vertices: after a pass,
//     the OWNER of a vertex the pass moved sends it to every rank that keeps it as halo without holding the stencil that
//     moved it.  One one-way exchange per pass, the same shape as the VBD colour exchange (partition.h).
// Predict and the velocity update run over owned + halo vertices (same arithmetic as on the owner), so a substep needs no
// other exchange.  Passes are numbered in execution order: energy-major, colour-minor.
#pragma once
#include <algorithm>
#include <cstdint>
#include <stdexcept>
#include <vector>

#include "host_build.h"

namespace spgx
{
struct XpbdPlan {
    int nOwned = 0;
    std::vector<int32_t> halo;         // non-owned vertices of the held stencils, ascending
    std::vector<int32_t> peers;        // ranks this one exchanges with, ascending
    int nPasses = 0;
    std::vector<int64_t> heldOffsets;  // nPasses + 1 into held
    std::vector<int32_t> held;         // stencil ids within their energy, ascending per pass
    std::vector<int64_t> sendOffsets;  // nPasses * nPeers + 1 into sendVerts ((pass, peer)-major)
    std::vector<int32_t> sendVerts;    // global vertex ids, ascending per (pass, peer)
    std::vector<int64_t> recvOffsets;  // same shape, into recvVerts: what peer sends in that pass, in its order
    std::vector<int32_t> recvVerts;
};

// colors[e]: the stencil colours of energy e (colorStencils), nColors[e] their count
inline XpbdPlan buildXpbdPlan(int nVerts, const std::vector<FlatStencilsView> &sets, const std::vector<const int32_t *> &colors,
                              const std::vector<int32_t> &nColors, const int32_t *owner, int rank, int nranks)
{
    if (nranks < 1 || rank < 0 || rank >= nranks) throw std::runtime_error("bad rank / rank count");
    if (nranks > 64) throw std::runtime_error("the XPBD plan supports at most 64 ranks");
    auto ownerOf = [&](int v) {
        return owner ? owner[v] : static_cast<int32_t>(static_cast<int64_t>(v) * nranks / std::max(nVerts, 1));
    };
    for (int v = 0; v < nVerts; ++v)
        if (ownerOf(v) < 0 || ownerOf(v) >= nranks) throw std::runtime_error("owner out of range");
    // ranks holding a copy of each vertex (its owner + every rank holding a stencil it belongs to)
    std::vector<uint64_t> holders(static_cast<size_t>(nVerts), 0);
    for (int v = 0; v < nVerts; ++v) holders[v] = 1ull << ownerOf(v);
    auto stencilMask = [&](const FlatStencilsView &fs, int64_t s) {
        uint64_t m = 0;
        for (int l = 0; l < fs.stencilSize; ++l) m |= 1ull << ownerOf(fs.entries[s * fs.stencilSize + l]);
        return m;
    };
    for (const auto &fs : sets) {
        const int64_t n = fs.nEntries / fs.stencilSize;
        for (int64_t s = 0; s < n; ++s) {
            const uint64_t m = stencilMask(fs, s);
            for (int l = 0; l < fs.stencilSize; ++l) holders[fs.entries[s * fs.stencilSize + l]] |= m;
        }
    }
    XpbdPlan plan;
    const uint64_t me = 1ull << rank;
    uint64_t peerMask = 0;
    for (int v = 0; v < nVerts; ++v) {
        if (ownerOf(v) == rank) {
            ++plan.nOwned;
            peerMask |= holders[v] & ~me;
        } else if (holders[v] & me) {
            plan.halo.push_back(v);
            peerMask |= 1ull << ownerOf(v);
        }
    }
    for (int r = 0; r < nranks; ++r)
        if ((peerMask >> r) & 1ull) plan.peers.push_back(r);
    std::vector<int> peerIndex(static_cast<size_t>(nranks), -1);
    for (size_t i = 0; i < plan.peers.size(); ++i) peerIndex[plan.peers[i]] = static_cast<int>(i);
    const size_t nPeers = plan.peers.size();

    for (size_t e = 0; e < sets.size(); ++e) plan.nPasses += nColors[e];
    std::vector<std::vector<int32_t>> held(static_cast<size_t>(plan.nPasses));
    std::vector<std::vector<int32_t>> send(static_cast<size_t>(plan.nPasses) * std::max<size_t>(nPeers, 1)),
        recv(static_cast<size_t>(plan.nPasses) * std::max<size_t>(nPeers, 1));
    int passBase = 0;
    for (size_t e = 0; e < sets.size(); ++e) {
        const auto &fs = sets[e];
        const int64_t n = fs.nEntries / fs.stencilSize;
        for (int64_t s = 0; s < n; ++s) {
            const int c = colors[e][s];
            if (c < 0 || c >= nColors[e]) throw std::runtime_error("stencil colour out of range");
            const size_t pass = static_cast<size_t>(passBase + c);
            const uint64_t m = stencilMask(fs, s);  // the ranks that hold (and execute) this stencil
            if (m & me) held[pass].push_back(static_cast<int32_t>(s));
            for (int l = 0; l < fs.stencilSize; ++l) {
                const int32_t v = fs.entries[s * fs.stencilSize + l];
                const int o = ownerOf(v);
                const uint64_t need = holders[v] & ~m;  // keep v as halo but cannot compute this update themselves
                if (o == rank) {
                    for (size_t i = 0; i < nPeers; ++i)
                        if ((need >> plan.peers[i]) & 1ull) send[pass * nPeers + i].push_back(v);
                } else if (need & me) {
                    recv[pass * nPeers + static_cast<size_t>(peerIndex[o])].push_back(v);
                }
            }
        }
        passBase += nColors[e];
    }
    plan.heldOffsets.assign(static_cast<size_t>(plan.nPasses) + 1, 0);
    for (int p = 0; p < plan.nPasses; ++p) {
        plan.held.insert(plan.held.end(), held[p].begin(), held[p].end());  // ascending: stencils were visited in order
        plan.heldOffsets[p + 1] = static_cast<int64_t>(plan.held.size());
    }
    const size_t cells = static_cast<size_t>(plan.nPasses) * nPeers;
    plan.sendOffsets.assign(cells + 1, 0);
    plan.recvOffsets.assign(cells + 1, 0);
    for (size_t k = 0; k < cells; ++k) {
        std::sort(send[k].begin(), send[k].end());
        std::sort(recv[k].begin(), recv[k].end());
        plan.sendVerts.insert(plan.sendVerts.end(), send[k].begin(), send[k].end());
        plan.recvVerts.insert(plan.recvVerts.end(), recv[k].begin(), recv[k].end());
        plan.sendOffsets[k + 1] = static_cast<int64_t>(plan.sendVerts.size());
        plan.recvOffsets[k + 1] = static_cast<int64_t>(plan.recvVerts.size());
    }
    return plan;
}
}  // namespace spgx
