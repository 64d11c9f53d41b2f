// Per-vertex incidence records of the streaming VBD colour pass (streams.cuh), host side, no CUDA: which incidence of a
// vertex goes to which stream type, in which order, and the canonical record of each.  buildStreams (api_build.cu) lays the
// result out as sliced ELL; patchStreamRecords re-runs it for one vertex after a live record edit; tests/native/
// host_stream_records.cpp replays the records of a whole mesh on the host and compares the per-vertex sums with the generic
// evaluation (reference: accumulateVertexForce / accumulateVertexHessian, spg/include/spg/sim/energy/energy.h:240-258).
#pragma once
#include <cstdint>
#include <vector>

#include "streams.cuh"

namespace spgx
{
// one energy as the record builder sees it (vertex ids in any ONE id space; `inc*`: for vertex v the ascending flat indices
// stencil * S + local at which v appears, host_build.h: Incidence)
struct StreamEnergyView {
    int kind = 0, S = 0, NP = 0;
    const int32_t *stencils = nullptr;
    const double *records = nullptr;
    const int64_t *incRowptr = nullptr;
    const int64_t *incFlat = nullptr;
};

struct SlotScratch {
    std::vector<int32_t> seen;  // unique neighbour vertex ids of the current vertex, first-seen order (a handful: a linear
                                // scan beats a V-sized mark array)
    std::vector<uint8_t> usedB, usedM;
};

// Appends the records of vertex v to idx[type] / q[type] (type: StreamType; streamNQ(type) doubles per record) and leaves its
// unique neighbours in sc.seen -- the 8-bit fields of an idx word index that list.  wedgeB / wedgeM: indices into `en` of the
// discrete-bending and the StVK-membrane energy when wedges are to be formed (-1: every incidence is a record of its own).
// Wedges (streams.cuh: Stream<ST_WEDGE>): every incident triangle (v, A, B) looks for the hinge on v-A whose opposite vertices
// include B and for the hinge on A-B opposite to v, among the incidences no earlier triangle has taken; A is the vertex after
// v in the triangle's cyclic order on the first pass (on a consistently oriented manifold mesh every interior edge at v is
// "the edge to the next vertex" of exactly one of its two triangles, so all 3 n incidences of an interior vertex pair up) and
// the vertex before v on a second pass over the triangles that are left.  What completes no wedge -- mesh boundary,
// non-manifold input, energies that do not cover the same triangles -- stays a single-incidence record, in the order
// (energy, flat index) of the CSR walk.  Returns true when v has incidences of a kind without a stream (CSR path).
inline bool appendSlotRecords(const StreamEnergyView *en, int nEnergies, int v, int wedgeB, int wedgeM, SlotScratch &sc,
                              std::vector<uint32_t> *idx, std::vector<double> *q)
{
    sc.seen.clear();
    auto emit = [&](int type, const int32_t *nbr, const double *rq) {
        uint32_t word = 0;
        for (int k = 0; k < streamNI(type); ++k) {
            size_t li = 0;
            while (li < sc.seen.size() && sc.seen[li] != nbr[k]) ++li;
            if (li == sc.seen.size()) sc.seen.push_back(nbr[k]);
            word |= static_cast<uint32_t>(li & 0xFFu) << (8 * k);
        }
        idx[type].push_back(word);
        q[type].insert(q[type].end(), rq, rq + streamNQ(type));
    };
    const bool wedges = wedgeB >= 0 && wedgeM >= 0;
    if (wedges) {
        const StreamEnergyView &eb = en[wedgeB], &em = en[wedgeM];
        const int64_t b0 = eb.incRowptr[v], nb = eb.incRowptr[v + 1] - b0;
        const int64_t m0 = em.incRowptr[v], nm = em.incRowptr[v + 1] - m0;
        sc.usedB.assign(static_cast<size_t>(nb), 0);
        sc.usedM.assign(static_cast<size_t>(nm), 0);
        for (int pass = 0; pass < 2; ++pass)
            for (int64_t im = 0; im < nm; ++im) {
                if (sc.usedM[static_cast<size_t>(im)]) continue;
                const int64_t ft = em.incFlat[m0 + im];
                const int32_t *tri = em.stencils + (ft / 3) * 3;
                const int localT = static_cast<int>(ft % 3);
                const bool aNext = pass == 0;
                const int32_t A = tri[(localT + (aNext ? 1 : 2)) % 3], B = tri[(localT + (aNext ? 2 : 1)) % 3];
                int64_t je = -1, jo = -1;
                for (int64_t j = 0; j < nb && (je < 0 || jo < 0); ++j) {
                    if (sc.usedB[static_cast<size_t>(j)]) continue;
                    const int64_t fb = eb.incFlat[b0 + j];
                    const int32_t *h = eb.stencils + (fb / 4) * 4;
                    const int l = static_cast<int>(fb % 4);
                    if (l < 2) {
                        if (je < 0 && h[1 - l] == A && (h[2] == B || h[3] == B)) je = j;
                    } else {
                        if (jo < 0 && ((h[0] == A && h[1] == B) || (h[0] == B && h[1] == A))) jo = j;
                    }
                }
                if (je < 0 || jo < 0) continue;
                const int64_t fe = eb.incFlat[b0 + je], fo = eb.incFlat[b0 + jo];
                int32_t nbr[4];
                double rq[16];
                if (!makeWedgeRecord(tri, em.records + (ft / 3) * em.NP, localT, aNext, eb.stencils + (fe / 4) * 4,
                                     eb.records + (fe / 4) * eb.NP, static_cast<int>(fe % 4), eb.stencils + (fo / 4) * 4,
                                     eb.records + (fo / 4) * eb.NP, static_cast<int>(fo % 4), nbr, rq))
                    continue;
                emit(ST_WEDGE, nbr, rq);
                sc.usedM[static_cast<size_t>(im)] = sc.usedB[static_cast<size_t>(je)] = sc.usedB[static_cast<size_t>(jo)] = 1;
            }
    }
    bool hasCsr = false;
    for (int ei = 0; ei < nEnergies; ++ei) {
        const StreamEnergyView &e = en[ei];
        if (streamTypeOf(e.kind, 0) < 0) {
            if (e.incRowptr[v + 1] > e.incRowptr[v]) hasCsr = true;
            continue;
        }
        const std::vector<uint8_t> *used = !wedges ? nullptr : (ei == wedgeB ? &sc.usedB : (ei == wedgeM ? &sc.usedM : nullptr));
        for (int64_t r = e.incRowptr[v]; r < e.incRowptr[v + 1]; ++r) {
            if (used && (*used)[static_cast<size_t>(r - e.incRowptr[v])]) continue;
            const int64_t stencil = e.incFlat[r] / e.S;
            const int local = static_cast<int>(e.incFlat[r] % e.S);
            int32_t nbr[3];
            double rq[16];
            makeStreamRecord(e.kind, e.stencils + stencil * e.S, e.records + stencil * e.NP, local, nbr, rq);
            emit(streamTypeOf(e.kind, local), nbr, rq);
        }
    }
    return hasCsr;
}
}  // namespace spgx
