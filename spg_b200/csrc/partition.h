// Domain decomposition of one particle group across the GPUs of a box (the reference is single
// process: there is nothing to cite but the colour-pass dependency structure of spg/source/sim/solver/vbd.cpp:151-157).
//
// Every rank sees the same global mesh and the same global colouring (bit-exact with the reference), owns the vertices
// `owner[v] == rank`, keeps every stencil that touches an owned vertex, and holds the other vertices of those
// stencils as HALO copies.  Within a colour pass all vertices on all ranks are independent, so the only exchange is:
// after colour pass c every rank sends the positions of its owned colour-c vertices that sit in a neighbour's halo.
// Local slot order:  [ owned, sorted by (colour, boundary first, vertex id) | halo, sorted by (owner, colour, vertex id) ]
// so a colour is a contiguous range of owned slots and what arrives from peer p after colour c lands in one contiguous
// range of halo slots (no unpack); the sender packs the matching list (same order, by construction, on both sides).
#pragma once
#include <algorithm>
#include <cstdint>
#include <stdexcept>
#include <vector>

#include "host_build.h"

namespace spgx
{
struct LocalView {
    int rank = 0, nRanks = 1;
    int nOwned = 0, nLocal = 0, nColors = 0;
    std::vector<int32_t> slot2vert;                   // nLocal global vertex ids
    std::vector<int32_t> vert2slot;                   // global vertex id -> local slot, -1 when not held
    std::vector<int64_t> colorOffsets;                // owned slots per colour, nColors + 1
    std::vector<std::vector<int32_t>> localStencils;  // per energy: ascending global stencil ids that touch an owned vertex
    // halo exchange tables, (colour, peer)-major: entry q = c * nPeers + p
    std::vector<int> peers;            // ranks exchanged with, ascending
    std::vector<int64_t> sendOffsets;  // nColors * nPeers + 1 offsets into sendSlots
    std::vector<int32_t> sendSlots;    // owned slots to pack, ordered by (colour, peer, vertex id)
    std::vector<int64_t> recvSlot;     // nColors * nPeers: first halo slot of what peer p sends after colour c
    std::vector<int64_t> recvCount;    // nColors * nPeers
};

// owner == nullptr: a single domain (everything owned, no halo, no peers)
inline LocalView buildLocalView(int nVerts, const std::vector<int32_t> &colors, const std::vector<FlatStencilsView> &sets,
                                const int32_t *owner, int rank, int nRanks)
{
    LocalView lv;
    lv.rank = rank;
    lv.nRanks = nRanks;
    int nColors = 0;
    for (int32_t c : colors) nColors = std::max(nColors, c + 1);
    lv.nColors = nColors;
    lv.vert2slot.assign(static_cast<size_t>(nVerts), -1);
    lv.localStencils.resize(sets.size());
    auto mine = [&](int v) { return owner == nullptr || owner[v] == rank; };
    if (owner)
        for (int v = 0; v < nVerts; ++v)
            if (owner[v] < 0 || owner[v] >= nRanks) throw std::runtime_error("vertex owner out of range");

    // stencils kept, halo vertices, and which of my vertices each peer needs
    std::vector<uint8_t> isHalo(static_cast<size_t>(nVerts), 0);
    std::vector<int> peerIndex(static_cast<size_t>(nRanks), -1);
    std::vector<std::vector<uint8_t>> needs;  // per peer: my vertices it holds as halo
    for (size_t e = 0; e < sets.size(); ++e) {
        const auto &fs = sets[e];
        const int S = fs.stencilSize;
        const int64_t n = fs.nEntries / S;
        for (int64_t i = 0; i < n; ++i) {
            const int32_t *st = fs.entries + i * S;
            bool any = false, foreign = false;
            for (int l = 0; l < S; ++l) {
                if (mine(st[l]))
                    any = true;
                else
                    foreign = true;
            }
            if (!any) continue;
            lv.localStencils[e].push_back(static_cast<int32_t>(i));
            if (!foreign) continue;
            for (int l = 0; l < S; ++l) {
                const int u = st[l];
                if (mine(u)) continue;
                isHalo[u] = 1;
                const int r = owner[u];
                if (peerIndex[r] < 0) {
                    peerIndex[r] = static_cast<int>(needs.size());
                    needs.emplace_back(static_cast<size_t>(nVerts), 0);
                }
                for (int m = 0; m < S; ++m)
                    if (mine(st[m])) needs[peerIndex[r]][st[m]] = 1;
            }
        }
    }
    for (int r = 0; r < nRanks; ++r)
        if (peerIndex[r] >= 0) lv.peers.push_back(r);
    const int nPeers = static_cast<int>(lv.peers.size());

    // owned slots: (colour, boundary first, id).  A boundary vertex is one some neighbour holds as halo; putting them at
    // the head of their colour makes them the first warps of the colour pass, so their peer-memory stores (and the
    // system-scope fence behind them) overlap the interior work instead of trailing the pass (halo.cuh).
    std::vector<uint8_t> isBoundary(static_cast<size_t>(nVerts), 0);
    for (const auto &need : needs)
        for (int v = 0; v < nVerts; ++v) isBoundary[v] |= need[v];
    lv.colorOffsets.assign(static_cast<size_t>(nColors) + 1, 0);
    std::vector<int64_t> boundaryCount(static_cast<size_t>(nColors), 0);
    for (int v = 0; v < nVerts; ++v)
        if (mine(v)) {
            lv.colorOffsets[colors[v] + 1]++;
            boundaryCount[colors[v]] += isBoundary[v];
        }
    for (int c = 0; c < nColors; ++c) lv.colorOffsets[c + 1] += lv.colorOffsets[c];
    lv.nOwned = static_cast<int>(lv.colorOffsets[nColors]);
    // halo slots: (owner, colour, id)
    std::vector<int64_t> haloOffsets(static_cast<size_t>(nPeers) * nColors + 1, 0);
    auto haloBucket = [&](int v) { return static_cast<size_t>(std::lower_bound(lv.peers.begin(), lv.peers.end(), owner[v]) - lv.peers.begin()) * nColors + colors[v]; };
    for (int v = 0; v < nVerts; ++v)
        if (isHalo[v]) haloOffsets[haloBucket(v) + 1]++;
    for (size_t b = 0; b + 1 < haloOffsets.size(); ++b) haloOffsets[b + 1] += haloOffsets[b];
    lv.nLocal = lv.nOwned + static_cast<int>(haloOffsets.back());
    lv.slot2vert.assign(static_cast<size_t>(lv.nLocal), -1);
    {
        std::vector<int64_t> cursor(lv.colorOffsets.begin(), lv.colorOffsets.end() - 1);  // boundary vertices
        std::vector<int64_t> icursor(static_cast<size_t>(nColors), 0);                      // interior vertices
        for (int c = 0; c < nColors; ++c) icursor[c] = lv.colorOffsets[c] + boundaryCount[c];
        std::vector<int64_t> hcursor(haloOffsets.begin(), haloOffsets.end() - 1);
        for (int v = 0; v < nVerts; ++v) {
            if (mine(v)) {
                const int s = static_cast<int>(isBoundary[v] ? cursor[colors[v]]++ : icursor[colors[v]]++);
                lv.slot2vert[s] = v;
                lv.vert2slot[v] = s;
            } else if (isHalo[v]) {
                const int s = lv.nOwned + static_cast<int>(hcursor[haloBucket(v)]++);
                lv.slot2vert[s] = v;
                lv.vert2slot[v] = s;
            }
        }
    }
    // exchange tables
    lv.sendOffsets.assign(static_cast<size_t>(nColors) * nPeers + 1, 0);
    lv.recvSlot.assign(static_cast<size_t>(nColors) * nPeers, 0);
    lv.recvCount.assign(static_cast<size_t>(nColors) * nPeers, 0);
    if (nPeers > 0) {
        for (int p = 0; p < nPeers; ++p) {
            const auto &need = needs[peerIndex[lv.peers[p]]];
            for (int v = 0; v < nVerts; ++v)
                if (need[v]) lv.sendOffsets[static_cast<size_t>(colors[v]) * nPeers + p + 1]++;
        }
        for (size_t q = 0; q + 1 < lv.sendOffsets.size(); ++q) lv.sendOffsets[q + 1] += lv.sendOffsets[q];
        lv.sendSlots.resize(static_cast<size_t>(lv.sendOffsets.back()));
        std::vector<int64_t> cursor(lv.sendOffsets.begin(), lv.sendOffsets.end() - 1);
        for (int p = 0; p < nPeers; ++p) {
            const auto &need = needs[peerIndex[lv.peers[p]]];
            for (int v = 0; v < nVerts; ++v)
                if (need[v]) lv.sendSlots[cursor[static_cast<size_t>(colors[v]) * nPeers + p]++] = lv.vert2slot[v];
        }
        for (int c = 0; c < nColors; ++c)
            for (int p = 0; p < nPeers; ++p) {
                const size_t hb = static_cast<size_t>(p) * nColors + c;
                lv.recvSlot[static_cast<size_t>(c) * nPeers + p] = lv.nOwned + haloOffsets[hb];
                lv.recvCount[static_cast<size_t>(c) * nPeers + p] = haloOffsets[hb + 1] - haloOffsets[hb];
            }
    }
    return lv;
}
}  // namespace spgx
