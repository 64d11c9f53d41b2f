// Host-side structure building: greedy colouring (bit-exact with the reference, but linear time),
// vertex -> (stencil, local) incidence, colour-sorted orderings.
//
// Reference: spg/source/utils/graphColoring.cpp:21-49 (Graph::greedyColoring), :51-70 (colorStencils),
// :72-84 (colorVertices); spg/source/sim/solver/vbd.cpp:203-268; spg/source/sim/solver/xpbd.cpp:133-160.
//
// The reference colours node n with the smallest colour not used by an *already coloured* neighbour.
// Nodes are visited in index order, so "already coloured" == "index < n" and the explicit std::list
// adjacency (with duplicate edges) is unnecessary: the neighbours of n are enumerated on the fly
// from the incidence tables.  The O(N) std::fill per node (graphColoring.cpp:29) becomes a stamp array.
#pragma once
#include <algorithm>
#include <cstdint>
#include <stdexcept>
#include <vector>

namespace spgx
{
struct FlatStencilsView {
    const int32_t *entries;
    int64_t nEntries;
    int stencilSize;
};

// CSR of flat-stencil positions per vertex: for vertex v, the ascending list of flat indices s with entries[s] == v
struct Incidence {
    std::vector<int64_t> rowptr;  // nVerts + 1
    std::vector<int64_t> flat;    // flat index s (stencil = s / S, local = s % S)
};

inline Incidence buildIncidence(int nVerts, const FlatStencilsView &fs)
{
    Incidence inc;
    inc.rowptr.assign(static_cast<size_t>(nVerts) + 1, 0);
    for (int64_t s = 0; s < fs.nEntries; ++s) {
        const int v = fs.entries[s];
        if (v < 0 || v >= nVerts) throw std::runtime_error("Trying to add edge to graph with invalid node indices");
        inc.rowptr[v + 1]++;
    }
    for (int v = 0; v < nVerts; ++v) inc.rowptr[v + 1] += inc.rowptr[v];
    inc.flat.resize(static_cast<size_t>(fs.nEntries));
    std::vector<int64_t> cursor(inc.rowptr.begin(), inc.rowptr.end() - 1);
    for (int64_t s = 0; s < fs.nEntries; ++s) inc.flat[cursor[fs.entries[s]]++] = s;
    return inc;
}

// the same table restricted to the stencils `kept` (ascending ids): complete for every vertex all of whose stencils are kept
// (a partitioned rank keeps every stencil touching a vertex it owns), O(nVerts + kept) instead of O(nVerts + all)
inline Incidence buildIncidenceOf(int nVerts, const FlatStencilsView &fs, const std::vector<int32_t> &kept)
{
    const int S = fs.stencilSize;
    Incidence inc;
    inc.rowptr.assign(static_cast<size_t>(nVerts) + 1, 0);
    for (int32_t st : kept)
        for (int l = 0; l < S; ++l) inc.rowptr[fs.entries[static_cast<int64_t>(st) * S + l] + 1]++;
    for (int v = 0; v < nVerts; ++v) inc.rowptr[v + 1] += inc.rowptr[v];
    inc.flat.resize(kept.size() * static_cast<size_t>(S));
    std::vector<int64_t> cursor(inc.rowptr.begin(), inc.rowptr.end() - 1);
    for (int32_t st : kept)
        for (int l = 0; l < S; ++l) {
            const int64_t s = static_cast<int64_t>(st) * S + l;
            inc.flat[cursor[fs.entries[s]]++] = s;  // kept is ascending, so flat indices ascend within a row
        }
    return inc;
}

struct StampSet {
    std::vector<int64_t> stamp;
    void forbid(int colour, int64_t tag)
    {
        if (colour >= static_cast<int>(stamp.size())) stamp.resize(colour + 1, -1);
        stamp[colour] = tag;
    }
    int firstFree(int64_t tag) const
    {
        int c = 0;
        while (c < static_cast<int>(stamp.size()) && stamp[c] == tag) ++c;
        return c;
    }
};

// graphColoring.cpp:72-84 + :21-49.  Two vertices are adjacent iff they appear in a common stencil of any set.
inline std::vector<int32_t> colorVertices(int nVerts, const std::vector<FlatStencilsView> &sets,
                                          const std::vector<Incidence> *incidences = nullptr)
{
    std::vector<int32_t> colour(static_cast<size_t>(nVerts), -1);
    if (nVerts == 0) return colour;
    std::vector<Incidence> local;
    if (!incidences) {
        for (const auto &fs : sets) local.push_back(buildIncidence(nVerts, fs));
        incidences = &local;
    }
    StampSet forbidden;
    for (int n = 0; n < nVerts; ++n) {
        for (size_t e = 0; e < sets.size(); ++e) {
            const auto &fs = sets[e];
            const auto &inc = (*incidences)[e];
            const int S = fs.stencilSize;
            for (int64_t q = inc.rowptr[n]; q < inc.rowptr[n + 1]; ++q) {
                const int64_t base = (inc.flat[q] / S) * S;
                for (int l = 0; l < S; ++l) {
                    const int u = fs.entries[base + l];
                    if (u < n) forbidden.forbid(colour[u], n);
                }
            }
        }
        colour[n] = forbidden.firstFree(n);
    }
    return colour;
}

// graphColoring.cpp:51-70 + :21-49.  Two stencils are adjacent iff they share a vertex.
inline std::vector<int32_t> colorStencils(const FlatStencilsView &fs)
{
    const int S = fs.stencilSize;
    const int64_t nStencils = fs.nEntries / S;
    std::vector<int32_t> colour(static_cast<size_t>(nStencils), -1);
    if (nStencils == 0) return colour;
    int maxV = -1;
    for (int64_t s = 0; s < fs.nEntries; ++s) {
        if (fs.entries[s] < 0) throw std::runtime_error("Trying to add edge to graph with invalid node indices");
        maxV = std::max(maxV, fs.entries[s]);
    }
    const Incidence inc = buildIncidence(maxV + 1, fs);
    StampSet forbidden;
    for (int64_t i = 0; i < nStencils; ++i) {
        for (int l = 0; l < S; ++l) {
            const int v = fs.entries[i * S + l];
            for (int64_t q = inc.rowptr[v]; q < inc.rowptr[v + 1]; ++q) {
                const int64_t j = inc.flat[q] / S;
                if (j >= i) break;  // flat indices ascend within a vertex row
                forbidden.forbid(colour[j], i);
            }
        }
        colour[i] = forbidden.firstFree(i);
    }
    return colour;
}

// stable counting sort of ids by key (colour): returns order (ids grouped by key, ascending id inside) and
// the nKeys + 1 group offsets
inline void groupByKey(const std::vector<int32_t> &key, std::vector<int32_t> &order, std::vector<int64_t> &offsets)
{
    int nKeys = 0;
    for (int32_t k : key) nKeys = std::max(nKeys, k + 1);
    offsets.assign(static_cast<size_t>(nKeys) + 1, 0);
    for (int32_t k : key) offsets[k + 1]++;
    for (int k = 0; k < nKeys; ++k) offsets[k + 1] += offsets[k];
    order.resize(key.size());
    std::vector<int64_t> cursor(offsets.begin(), offsets.end() - 1);
    for (size_t i = 0; i < key.size(); ++i) order[cursor[key[i]]++] = static_cast<int32_t>(i);
}
}  // namespace spgx
