// C-ABI, part 1: library queries, pure-host structure functions, context, group construction and
// finalize (colouring + incidence + upload).  See include/spgx.h for the reference interface each
// entry point replaces.
#include <algorithm>
#include <array>
#include <chrono>
#include <cstdio>
#include <exception>
#include <thread>
#include <sched.h>
#include <cstdlib>
#include <cstring>

#include "energy_prepare.h"
#include "group.h"
#include "host_build.h"
#include "partition.h"
#include "xpbd_partition.h"
#include "stream_records.h"
#include "streams.cuh"

namespace spgx
{
static thread_local std::string g_lastError;
void setLastError(const std::string &msg)
{
    g_lastError = msg;
}

// SPGX_FINALIZE_TIMING=1: where spgx_group_finalize spends its host time, one line per phase on stderr
struct PhaseTimer {
    bool on = getenv("SPGX_FINALIZE_TIMING") && atoi(getenv("SPGX_FINALIZE_TIMING")) != 0;
    std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now();
    void lap(const char *what)
    {
        if (!on) return;
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[spgx finalize] %-28s %8.3f s\n", what, std::chrono::duration<double>(now - t).count());
        t = now;
    }
};

// fn(begin, end, chunk) over [0, n) in contiguous chunks on host threads (SPGX_HOST_THREADS, default: the cores this process
// may use, at most 32).  Chunk boundaries are a function of n and the thread count only.
template <class F>
static void parallelChunks(int64_t n, int nChunks, F &&fn)
{
    nChunks = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(nChunks, n)));
    std::vector<std::thread> pool;
    std::vector<std::exception_ptr> err(static_cast<size_t>(nChunks));
    for (int c = 0; c < nChunks; ++c) {
        const int64_t b = n * c / nChunks, e = n * (c + 1) / nChunks;
        auto body = [&fn, &err, b, e, c] {
            try {
                fn(b, e, c);
            } catch (...) {
                err[static_cast<size_t>(c)] = std::current_exception();
            }
        };
        if (c + 1 < nChunks)
            pool.emplace_back(body);
        else
            body();
    }
    for (auto &t : pool) t.join();
    for (auto &e : err)
        if (e) std::rethrow_exception(e);
}
static int hostThreads()
{
    if (const char *env = getenv("SPGX_HOST_THREADS")) return std::max(1, atoi(env));
    cpu_set_t set;
    CPU_ZERO(&set);
    int n = 0;
    if (sched_getaffinity(0, sizeof(set), &set) == 0) n = CPU_COUNT(&set);
    if (n <= 0) n = static_cast<int>(std::thread::hardware_concurrency());
    return std::max(1, std::min(n, 32));
}

static void checkKind(int kind)
{
    if (kind < 0 || kind >= K_COUNT) throw std::runtime_error("unknown energy kind");
}

// the group's energies as stream_records.h sees them (vertex ids, reference order)
static std::vector<StreamEnergyView> streamEnergyViews(const spgx_group &grp, const std::vector<Incidence> &inc)
{
    std::vector<StreamEnergyView> views(grp.energies.size());
    for (size_t ei = 0; ei < grp.energies.size(); ++ei) {
        const auto &e = grp.energies[ei];
        views[ei] = {e.kind, e.S, e.NP, e.stencils.data(), e.records.data(), inc[ei].rowptr.data(), inc[ei].flat.data()};
    }
    return views;
}
// which energies form wedges (streams.cuh: Stream<ST_WEDGE>): exactly one discrete-bending and one StVK-membrane energy and no
// tetrahedra (the wedge stream takes the descriptor word of ST_TET_STVK); -1, -1 otherwise or with SPGX_VBD_WEDGE=0
static void streamWedgeEnergies(const spgx_group &grp, int &wedgeB, int &wedgeM)
{
    wedgeB = wedgeM = -1;
    if (!grp.vbdWedge) return;
    int nb = 0, nm = 0, nt = 0, b = -1, m = -1;
    for (size_t ei = 0; ei < grp.energies.size(); ++ei) {
        const int k = grp.energies[ei].kind;
        if (k == K_DISCRETE_BENDING) ++nb, b = static_cast<int>(ei);
        if (k == K_MEMBRANE_STVK) ++nm, m = static_cast<int>(ei);
        if (k == K_TET_SNH || k == K_TET_STVK) ++nt;
    }
    if (nb == 1 && nm == 1 && nt == 0) wedgeB = b, wedgeM = m;
}

// Streaming layout of the VBD colour pass (kernels.h, streams.cuh).  For every vertex (in slot order): the list of its
// unique neighbours over all streamed incidences (first-seen order), and per stream type the canonical incidence
// records referencing that list by 8-bit index.  Both are packed as sliced ELL over warp-slices of VW = 32 / G
// consecutive slots of one colour: incidence j of the slice's vertex v sits in step j / G at lane (j % G) * VW + v;
// neighbour k of vertex v sits in row k at column v.  One 32-byte descriptor per slice holds where each of its
// ranges starts.
// Returns false (and leaves the group on the CSR path) when a vertex has more than 255 unique neighbours, two
// energies map to the same stream type, or the neighbour tile would not fit in shared memory.
static bool buildStreams(spgx_group &grp, const std::vector<Incidence> &inc, cudaStream_t st)
{
    PhaseTimer pt;
    const int V = grp.nVerts, NO = grp.nOwned, G = grp.vbdLanes, VW = 32 / G;  // V: global vertices, NO: owned slots
    const int nColors = static_cast<int>(grp.colorOffsets.size()) - 1;
    struct HostStream {
        int energy = -1, NQ = 0;
        std::vector<uint32_t> rowptr;  // per slot
        std::vector<uint32_t> idx;     // per incidence
        std::vector<double> q;         // NQ per incidence
    };
    HostStream hs[ST_COUNT];
    bool any = false;
    for (size_t ei = 0; ei < grp.energies.size(); ++ei) {
        const auto &e = grp.energies[ei];
        for (int local = 0; local < e.S; ++local) {
            const int t = streamTypeOf(e.kind, local);
            if (t < 0) continue;
            if (hs[t].energy >= 0 && hs[t].energy != static_cast<int>(ei)) return false;
            if (hs[t].energy < 0) {
                hs[t].energy = static_cast<int>(ei);
                hs[t].NQ = streamNQ(t);
                hs[t].rowptr.assign(static_cast<size_t>(NO) + 1, 0);
                any = true;
            }
        }
    }
    if (!any) return false;
    // wedges (streams.cuh: Stream<ST_WEDGE>): a cloth whose discrete bending and StVK membrane cover the same triangles
    const std::vector<StreamEnergyView> views = streamEnergyViews(grp, inc);
    int wedgeB = -1, wedgeM = -1;
    streamWedgeEnergies(grp, wedgeB, wedgeM);
    if (wedgeB >= 0) {
        hs[ST_WEDGE].energy = wedgeM;
        hs[ST_WEDGE].NQ = streamNQ(ST_WEDGE);
        hs[ST_WEDGE].rowptr.assign(static_cast<size_t>(NO) + 1, 0);
    }

    // per slot: unique neighbours (first-seen order) and the canonical records of its incidences (stream_records.h) -- slots
    // are independent, so contiguous chunks of slots are built on host threads into chunk-local arrays and concatenated in
    // slot order
    std::vector<uint32_t> nbrRowptr(static_cast<size_t>(NO) + 1, 0);
    std::vector<uint32_t> nbrList;
    std::vector<uint8_t> hasCsr(static_cast<size_t>(NO), 0);
    int maxNbr = 0;
    (void)V;
    {
        const int nChunks = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(hostThreads(), NO / 4096 + 1)));
        struct Chunk {
            std::vector<uint32_t> nbr;
            std::vector<uint32_t> idx[ST_COUNT];
            std::vector<double> q[ST_COUNT];
            int maxNbr = 0;
            bool tooMany = false;
        };
        std::vector<Chunk> chunks(static_cast<size_t>(nChunks));
        parallelChunks(NO, nChunks, [&](int64_t s0, int64_t s1, int c) {
            Chunk &ck = chunks[static_cast<size_t>(c)];
            SlotScratch sc;
            for (int64_t s = s0; s < s1; ++s) {
                const int v = grp.slot2vert[static_cast<size_t>(s)];
                for (int t = 0; t < ST_COUNT; ++t)
                    if (hs[t].energy >= 0) hs[t].rowptr[static_cast<size_t>(s)] = static_cast<uint32_t>(ck.idx[t].size());  // chunk-local for now
                if (appendSlotRecords(views.data(), static_cast<int>(views.size()), v, wedgeB, wedgeM, sc, ck.idx, ck.q))
                    hasCsr[static_cast<size_t>(s)] = 1;
                for (const int32_t u : sc.seen) ck.nbr.push_back(static_cast<uint32_t>(grp.vert2slot[u]));
                const int cnt = static_cast<int>(sc.seen.size());
                if (cnt > 254) ck.tooMany = true;
                ck.maxNbr = std::max(ck.maxNbr, cnt);
                nbrRowptr[static_cast<size_t>(s) + 1] = static_cast<uint32_t>(cnt);  // a count for now
            }
        });
        for (const auto &ck : chunks) {
            if (ck.tooMany) return false;
            maxNbr = std::max(maxNbr, ck.maxNbr);
        }
        // chunk-local offsets -> global ones, then the chunks' arrays side by side
        for (int s = 0; s < NO; ++s) nbrRowptr[static_cast<size_t>(s) + 1] += nbrRowptr[static_cast<size_t>(s)];
        nbrList.resize(nbrRowptr[static_cast<size_t>(NO)]);
        std::vector<size_t> idxBase[ST_COUNT];
        for (int t = 0; t < ST_COUNT; ++t) {
            if (hs[t].energy < 0) continue;
            idxBase[t].assign(static_cast<size_t>(nChunks) + 1, 0);
            for (int c = 0; c < nChunks; ++c) idxBase[t][static_cast<size_t>(c) + 1] = idxBase[t][static_cast<size_t>(c)] + chunks[static_cast<size_t>(c)].idx[t].size();
            hs[t].idx.resize(idxBase[t][static_cast<size_t>(nChunks)]);
            hs[t].q.resize(idxBase[t][static_cast<size_t>(nChunks)] * static_cast<size_t>(hs[t].NQ));
        }
        parallelChunks(NO, nChunks, [&](int64_t s0, int64_t s1, int c) {
            const Chunk &ck = chunks[static_cast<size_t>(c)];
            std::copy(ck.nbr.begin(), ck.nbr.end(), nbrList.begin() + nbrRowptr[static_cast<size_t>(s0)]);
            for (int t = 0; t < ST_COUNT; ++t) {
                if (hs[t].energy < 0) continue;
                const size_t base = idxBase[t][static_cast<size_t>(c)];
                std::copy(ck.idx[t].begin(), ck.idx[t].end(), hs[t].idx.begin() + base);
                std::copy(ck.q[t].begin(), ck.q[t].end(), hs[t].q.begin() + base * static_cast<size_t>(hs[t].NQ));
                for (int64_t s = s0; s < s1; ++s) hs[t].rowptr[static_cast<size_t>(s)] += static_cast<uint32_t>(base);
            }
        });
    }
    pt.lap("  streams: records per slot");
    maxNbr = std::max(maxNbr, 1);
    // shared memory of a 4-warp block (vbd_stream.cuh: tile rows + the pair path's record ring, sized for the widest record)
    if ((static_cast<size_t>(maxNbr + 3) * 3 * VW * 8 + 2 * (128 + 12 * 256)) * 4 > 200 * 1024) return false;
    for (auto &h : hs)
        if (h.energy >= 0) h.rowptr[NO] = static_cast<uint32_t>(h.idx.size());

    // ---- slices and their descriptors ----
    std::vector<int> sliceSlot;  // first slot of every slice
    std::vector<int> sliceEnd;   // one past its last slot
    std::vector<int> sliceColour;
    for (int c = 0; c < nColors; ++c) {
        const int b = static_cast<int>(grp.colorOffsets[c]), en = static_cast<int>(grp.colorOffsets[c + 1]);
        for (int s0 = b; s0 < en; s0 += VW) {
            sliceSlot.push_back(s0);
            sliceEnd.push_back(std::min(s0 + VW, en));
            sliceColour.push_back(c);
        }
    }
    // push list of the peer-memory halo exchange (comm.cu), ascending slot = slice order
    std::vector<spgx_group::PushEntry> pushAll;
    std::vector<int> pushWarps(static_cast<size_t>(std::max(nColors, 1)), 0);
    {
        const int nPeers = static_cast<int>(grp.peers.size());
        for (int c = 0; c < nColors && nPeers > 0; ++c)
            for (int p = 0; p < nPeers; ++p) {
                const size_t q = static_cast<size_t>(c) * nPeers + p;
                for (int64_t i = grp.sendOffsets[q]; i < grp.sendOffsets[q + 1]; ++i)
                    pushAll.push_back({grp.p2p.sendSlots[static_cast<size_t>(i)], p, c, static_cast<int32_t>(i - grp.sendOffsets[q])});
            }
        std::stable_sort(pushAll.begin(), pushAll.end(),
                         [](const spgx_group::PushEntry &a, const spgx_group::PushEntry &b) { return a.slot < b.slot; });
    }
    size_t pushCursor = 0, pushBefore = 0;
    int64_t nUniform = 0;
    const size_t nSlices = sliceSlot.size();
    std::vector<uint32_t> desc((nSlices + 1) * kSliceDescWords, 0u);
    double bytes = 4.0 * desc.size();
    {
        uint32_t row = 0;
        for (size_t sl = 0; sl < nSlices; ++sl) {
            desc[sl * kSliceDescWords + 0] = static_cast<uint32_t>(sliceSlot[sl]);
            desc[sl * kSliceDescWords + 1] = row;
            uint32_t rows = 0, csr = 0, readsHalo = 0;
            // uniform: a full slice whose vertices all have the same number of incidences in every stream, a multiple
            // of 2 G -- every lane then has a record in every step and the steps pair up (the pair path of the kernel)
            uint32_t uniform = (sliceEnd[sl] - sliceSlot[sl] == VW) ? 1u : 0u;
            for (auto &h : hs) {
                if (h.energy < 0 || !uniform) continue;
                const uint32_t c0 = h.rowptr[sliceSlot[sl] + 1] - h.rowptr[sliceSlot[sl]];
                if (c0 % (2u * G) != 0u) uniform = 0;
                for (int s = sliceSlot[sl]; s < sliceEnd[sl] && uniform; ++s)
                    if (h.rowptr[s + 1] - h.rowptr[s] != c0) uniform = 0;
            }
            nUniform += uniform;
            for (int s = sliceSlot[sl]; s < sliceEnd[sl]; ++s) {
                rows = std::max(rows, nbrRowptr[s + 1] - nbrRowptr[s]);
                csr |= hasCsr[s];
                for (uint32_t k = nbrRowptr[s]; k < nbrRowptr[s + 1]; ++k)
                    if (nbrList[k] >= static_cast<uint32_t>(NO)) readsHalo = 1;
            }
            // boundary vertices of the slice: one push entry per (vertex, neighbouring rank that holds it as halo)
            while (pushCursor < pushAll.size() && pushAll[pushCursor].slot < sliceEnd[sl]) ++pushCursor;
            if (pushCursor >= (1u << 29)) throw std::runtime_error("too many boundary vertices for the slice descriptor");
            if (pushCursor > pushBefore) pushWarps[sliceColour[sl]]++;
            desc[sl * kSliceDescWords + 7] = csr | (readsHalo << 1) | (uniform << 2) | (static_cast<uint32_t>(pushBefore) << 3);
            pushBefore = pushCursor;
            row += rows;
        }
        desc[nSlices * kSliceDescWords + 0] = static_cast<uint32_t>(NO);
        desc[nSlices * kSliceDescWords + 1] = row;
        desc[nSlices * kSliceDescWords + 7] = static_cast<uint32_t>(pushBefore) << 3;
        grp.p2p.pushLocal = pushAll;
        grp.p2p.pushWarps = pushWarps;
        // packed: slice sl starts at its cumulative row; fixed stride (vbdNbrFixed): at row sl * maxNbr, so that the kernel can
        // request the ids before the descriptor has arrived (unused rows hold slot 0, a harmless gather target)
        const bool fixed = grp.vbdNbrFixed != 0;
        std::vector<uint32_t> ids((fixed ? nSlices * static_cast<size_t>(maxNbr) : static_cast<size_t>(row)) * VW, 0u);
        parallelChunks(static_cast<int64_t>(nSlices), hostThreads(), [&](int64_t a, int64_t b, int) {
            for (size_t sl = static_cast<size_t>(a); sl < static_cast<size_t>(b); ++sl) {
                const size_t row0 = fixed ? sl * static_cast<size_t>(maxNbr) : static_cast<size_t>(desc[sl * kSliceDescWords + 1]);
                for (int s = sliceSlot[sl]; s < sliceEnd[sl]; ++s)
                    for (uint32_t k = nbrRowptr[s]; k < nbrRowptr[s + 1]; ++k)
                        ids[(row0 + (k - nbrRowptr[s])) * VW + (s - sliceSlot[sl])] = nbrList[k];
            }
        });
        grp.d_nbrIds.upload(ids, st);
        bytes += 4.0 * ids.size();
        SPGX_CUDA(cudaStreamSynchronize(st));
    }
    pt.lap("  streams: descriptors + ids");
    grp.streams.clear();
    grp.streams.resize(ST_COUNT);
    grp.streamMask = 0;
    for (int type = 0; type < ST_COUNT; ++type) {
        const HostStream &h = hs[type];
        if (h.energy < 0) continue;
        const int NQ = h.NQ;
        uint32_t step = 0;
        for (size_t sl = 0; sl < nSlices; ++sl) {
            desc[sl * kSliceDescWords + streamDescWord(type)] = step;
            uint32_t steps = 0;
            for (int s = sliceSlot[sl]; s < sliceEnd[sl]; ++s) steps = std::max(steps, (h.rowptr[s + 1] - h.rowptr[s] + G - 1) / G);
            step += steps;
        }
        desc[nSlices * kSliceDescWords + streamDescWord(type)] = step;
        const size_t T = step;
        std::vector<uint32_t> idx(T * 32, 0xFFFFFFFFu);
        std::vector<double> params(T * NQ * 32, 0.0);
        parallelChunks(static_cast<int64_t>(nSlices), hostThreads(), [&](int64_t a, int64_t b, int) {
            for (size_t sl = static_cast<size_t>(a); sl < static_cast<size_t>(b); ++sl)
                for (int s = sliceSlot[sl]; s < sliceEnd[sl]; ++s) {
                    uint32_t j = 0;
                    for (uint32_t r = h.rowptr[s]; r < h.rowptr[s + 1]; ++r, ++j) {
                        const size_t t = desc[sl * kSliceDescWords + streamDescWord(type)] + j / G;
                        const int lane = static_cast<int>(j % G) * VW + (s - sliceSlot[sl]);
                        idx[t * 32 + lane] = h.idx[r];
                        for (int k = 0; k < NQ; ++k) params[(t * NQ + k) * 32 + lane] = h.q[static_cast<size_t>(r) * NQ + k];
                    }
                }
        });
        auto &out = grp.streams[type];
        out.type = type;
        out.d_idx.upload(idx, st);
        out.d_params.upload(params, st);
        if (T > 0) grp.streamMask |= 1u << type;  // (a type without a single record -- e.g. every membrane incidence joined a wedge -- has no stream)
        bytes += 4.0 * idx.size() + 8.0 * params.size();
        SPGX_CUDA(cudaStreamSynchronize(st));
    }
    pt.lap("  streams: SoA fill + uploads");
    grp.d_sliceDesc.upload(desc, st);
    SPGX_CUDA(cudaStreamSynchronize(st));
    grp.sliceDescHost = std::move(desc);
    grp.nSlices = static_cast<int>(nSlices);
    grp.maxNbr = maxNbr;
    grp.streamBytes = bytes;
    grp.wedgeRecords = static_cast<int64_t>(hs[ST_WEDGE].idx.size());
    for (auto &e : grp.energies) e.streamed = streamTypeOf(e.kind, 0) >= 0;
    return true;
}

// spgx_energy_update_record on an energy whose incidences live in streams: the colour pass never reads d_records for it,
// it reads the canonical per-incidence records baked into the streams at finalize (buildStreams).  For every vertex of the
// edited stencil that this rank owns, the vertex's records are built again (stream_records.h: the same function, so the same
// assignment of incidences to stream types, wedges included, and the same order) and their constants overwritten in place.
// Where record j of a type sits: the vertex's slot fixes the warp-slice and the column, j the step and the lane group.
// The incidence lists come from a scan of the stencil arrays (record edits are rare; a table would cost 8 bytes per incidence
// at every finalize).
static void patchStreamRecords(spgx_group &grp, const EnergyState &edited, int stencil)
{
    const int G = grp.vbdLanes, VW = 32 / G;
    cudaStream_t st = grp.ctx->stream;
    int wedgeB = -1, wedgeM = -1;
    streamWedgeEnergies(grp, wedgeB, wedgeM);
    for (int local = 0; local < edited.S; ++local) {
        const int v = edited.stencils[static_cast<size_t>(stencil) * edited.S + local];
        bool earlier = false;  // the same vertex twice in the stencil: patched once
        for (int l = 0; l < local; ++l) earlier |= edited.stencils[static_cast<size_t>(stencil) * edited.S + l] == v;
        const int slot = grp.vert2slot[v];
        if (earlier || slot < 0 || slot >= grp.nOwned) continue;  // a halo copy: its owner patches it
        // incidence lists of v alone: a one-row CSR per energy (appendSlotRecords uses its vertex argument as the row only)
        std::vector<std::vector<int64_t>> flat(grp.energies.size());
        std::vector<std::array<int64_t, 2>> rowptr(grp.energies.size());
        std::vector<StreamEnergyView> views(grp.energies.size());
        for (size_t ei = 0; ei < grp.energies.size(); ++ei) {
            const auto &e = grp.energies[ei];
            if (streamTypeOf(e.kind, 0) >= 0)
                for (size_t q = 0; q < e.stencils.size(); ++q)
                    if (e.stencils[q] == v) flat[ei].push_back(static_cast<int64_t>(q));
            rowptr[ei] = {0, static_cast<int64_t>(flat[ei].size())};
            views[ei] = {e.kind, e.S, e.NP, e.stencils.data(), e.records.data(), rowptr[ei].data(), flat[ei].data()};
        }
        SlotScratch sc;
        std::vector<uint32_t> idx[ST_COUNT];
        std::vector<double> q[ST_COUNT];
        appendSlotRecords(views.data(), static_cast<int>(views.size()), /*row*/ 0, wedgeB, wedgeM, sc, idx, q);
        int colour = 0;
        while (grp.colorOffsets[colour + 1] <= slot) ++colour;
        const int inColour = slot - static_cast<int>(grp.colorOffsets[colour]);
        const size_t slice = static_cast<size_t>(grp.sliceBase[colour]) + inColour / VW;
        for (int type = 0; type < ST_COUNT; ++type) {
            const int NQ = streamNQ(type);
            for (size_t j = 0; j < idx[type].size(); ++j) {
                const uint32_t t = grp.sliceDescHost[slice * kSliceDescWords + streamDescWord(type)] + static_cast<uint32_t>(j) / G;
                const int lane = static_cast<int>(j % G) * VW + inColour % VW;
                double *dst = grp.streams[type].d_params.ptr + static_cast<size_t>(t) * NQ * 32 + lane;
                // q[k] goes to params[t][k][lane]: NQ doubles, 32 doubles apart
                SPGX_CUDA(cudaMemcpy2DAsync(dst, 32 * sizeof(double), q[type].data() + j * NQ, sizeof(double), sizeof(double), NQ,
                                            cudaMemcpyHostToDevice, st));
            }
        }
        SPGX_CUDA(cudaStreamSynchronize(st));
    }
}
}  // namespace spgx

using namespace spgx;

extern "C" {

const char *spgx_last_error(void)
{
    return g_lastError.c_str();
}
const char *spgx_version(void)
{
    return "spgx 0.1 (sm_100a)";
}
int spgx_kind_stencil_size(int kind)
{
    return (kind >= 0 && kind < K_COUNT) ? kindStencilSize(kind) : -1;
}
int spgx_kind_num_constraints(int kind)
{
    return kindNumConstraints(kind);
}
int spgx_kind_num_model_params(int kind)
{
    return kindNumModelParams(kind);
}
int spgx_kind_num_record_params(int kind)
{
    return kindNumParams(kind);
}

int spgx_color_vertices(int nVerts, int nSets, const int32_t *const *entries, const int64_t *nEntries,
                        const int32_t *stencilSizes, int32_t *colors)
{
    return guarded([&] {
        std::vector<FlatStencilsView> sets;
        for (int s = 0; s < nSets; ++s) sets.push_back({entries[s], nEntries[s], stencilSizes[s]});
        const auto c = colorVertices(nVerts, sets);
        std::copy(c.begin(), c.end(), colors);
    });
}

int spgx_color_stencils(const int32_t *entries, int64_t nEntries, int stencilSize, int32_t *colors)
{
    return guarded([&] {
        const auto c = colorStencils({entries, nEntries, stencilSize});
        std::copy(c.begin(), c.end(), colors);
    });
}

int spgx_prepare_energy(int kind, int nStencils, const int32_t *stencils, const double *modelParams, int nVerts,
                        const double *x0, double *records)
{
    return guarded([&] {
        checkKind(kind);
        const int S = kindStencilSize(kind), NP = kindNumParams(kind), NM = kindNumModelParams(kind);
        for (int i = 0; i < nStencils; ++i) {
            for (int s = 0; s < S; ++s)
                if (stencils[i * S + s] < 0 || stencils[i * S + s] >= nVerts)
                    throw std::runtime_error("stencil vertex index out of range");
            prepareStencil(kind, stencils + static_cast<size_t>(i) * S, modelParams + static_cast<size_t>(i) * NM, x0,
                           records + static_cast<size_t>(i) * NP);
        }
    });
}

int spgx_ctx_create(int device, spgx_ctx **out)
{
    return guarded([&] {
        int count = 0;
        cudaError_t err = cudaGetDeviceCount(&count);
        if (err != cudaSuccess || count == 0)
            throw CudaError(std::string("no CUDA device available (there is no CPU fallback): ") +
                            cudaGetErrorString(err));
        if (device < 0 || device >= count) throw CudaError("CUDA device index out of range");
        SPGX_CUDA(cudaSetDevice(device));
        auto *ctx = new spgx_ctx;
        ctx->device = device;
        SPGX_CUDA(cudaGetDeviceProperties(&ctx->prop, device));
        SPGX_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
        *out = ctx;
    });
}

int spgx_comm_unique_id(void *id128)
{
    return guarded([&] { commUniqueId(id128); });
}

int spgx_ctx_init_comm(spgx_ctx *ctx, int rank, int nranks, const void *id128)
{
    return guarded([&] {
        if (!ctx) throw StateError("null context");
        if (nranks < 1 || rank < 0 || rank >= nranks) throw std::runtime_error("invalid rank / nranks");
        SPGX_CUDA(cudaSetDevice(ctx->device));
        commInit(ctx, rank, nranks, id128);
    });
}

int spgx_group_set_partition(spgx_group *grp, int rank, int nranks, const int32_t *owner)
{
    return guarded([&] {
        if (!grp) throw StateError("null group");
        if (nranks < 1 || rank < 0 || rank >= nranks) throw std::runtime_error("invalid rank / nranks");
        grp->rank = rank;
        grp->nRanks = nranks;
        grp->partitioned = nranks > 1;
        grp->owner.resize(static_cast<size_t>(grp->nVerts));
        for (int v = 0; v < grp->nVerts; ++v) {
            // equal contiguous index ranges: row strips of a grid, slabs of a beam
            const int o = owner ? owner[v] : static_cast<int>(static_cast<int64_t>(v) * nranks / std::max(grp->nVerts, 1));
            if (o < 0 || o >= nranks) throw std::runtime_error("vertex owner out of range");
            grp->owner[v] = o;
        }
        grp->finalized = false;
    });
}

int spgx_group_partition_info(spgx_group *grp, int64_t *out8)
{
    return guarded([&] {
        if (!grp || !grp->finalized) throw StateError("group is not finalized");
        out8[0] = grp->nOwned, out8[1] = grp->nLocal, out8[2] = static_cast<int64_t>(grp->peers.size());
        out8[3] = grp->localLo, out8[4] = grp->localHi, out8[5] = grp->ownedLo, out8[6] = grp->ownedHi;
        out8[7] = static_cast<int64_t>(grp->colorOffsets.size()) - 1;
    });
}

int spgx_group_p2p_export(spgx_group *grp, void *blob, int64_t capacity, int64_t *size)
{
    return guarded([&] {
        if (!grp) throw StateError("null group");
        const int64_t n = p2pExport(grp, blob, capacity);
        if (size) *size = n;
    });
}

int spgx_group_p2p_connect(spgx_group *grp, int nBlobs, const void *const *blobs, const int64_t *sizes)
{
    return guarded([&] {
        if (!grp || !blobs || !sizes) throw StateError("null argument");
        p2pConnect(grp, nBlobs, blobs, sizes);
    });
}

int spgx_group_p2p_disconnect(spgx_group *grp)
{
    return guarded([&] {
        if (!grp) throw StateError("null group");
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        SPGX_CUDA(cudaStreamSynchronize(grp->ctx->stream));
        p2pDisconnect(grp);
    });
}

int spgx_group_set_halo_mode(spgx_group *grp, int mode)
{
    return guarded([&] {
        if (!grp || !grp->finalized) throw StateError("group is not finalized");
        if (mode == SPGX_HALO_P2P && !grp->p2p.connected) throw StateError("the group is not connected to its neighbours (spgx_group_p2p_connect)");
        if (mode == SPGX_HALO_NONE && grp->partitioned && !grp->peers.empty()) throw StateError("a partitioned group needs a halo exchange");
        if (mode != SPGX_HALO_NONE && mode != SPGX_HALO_NCCL && mode != SPGX_HALO_P2P) throw std::runtime_error("unknown halo mode");
        if (!grp->partitioned) mode = SPGX_HALO_NONE;
        if (mode != grp->p2p.mode && grp->graphExec) {
            cudaGraphExecDestroy(grp->graphExec);
            grp->graphExec = nullptr;
        }
        grp->p2p.mode = mode;
    });
}

int spgx_group_get_halo_mode(spgx_group *grp, int *mode)
{
    return guarded([&] {
        if (!grp || !mode) throw StateError("null argument");
        *mode = grp->p2p.mode;
    });
}

int spgx_group_kernel_info(spgx_group *grp, int64_t *out8)
{
    return guarded([&] {
        if (!grp || !grp->finalized || !out8) throw StateError("group is not finalized");
        out8[0] = static_cast<int64_t>(grp->streamMask) | (grp->wedgeRecords << 8), out8[1] = grp->vbdLanes, out8[2] = grp->nSlices, out8[3] = grp->maxNbr;
        out8[4] = grp->p2p.mode, out8[5] = static_cast<int64_t>(grp->p2p.sendSlots.size()), out8[6] = grp->colouringRounds, out8[7] = grp->vbdOcc;
    });
}

int spgx_partition_plan(int nVerts, int nSets, const int32_t *const *entries, const int64_t *nEntries,
                        const int32_t *stencilSizes, const int32_t *owner, int rank, int nranks, int64_t *sizes,
                        int32_t *slot2vert, int32_t *peers, int64_t *sendOffsets, int32_t *sendVerts, int64_t *recvSlot,
                        int64_t *recvCount)
{
    return guarded([&] {
        std::vector<FlatStencilsView> sets;
        for (int s = 0; s < nSets; ++s) sets.push_back({entries[s], nEntries[s], stencilSizes[s]});
        const auto colors = colorVertices(nVerts, sets);
        std::vector<int32_t> own(static_cast<size_t>(nVerts));
        for (int v = 0; v < nVerts; ++v)
            own[v] = owner ? owner[v] : static_cast<int>(static_cast<int64_t>(v) * nranks / std::max(nVerts, 1));
        const LocalView lv = buildLocalView(nVerts, colors, sets, nranks > 1 ? own.data() : nullptr, rank, nranks);
        if (sizes) {
            sizes[0] = lv.nOwned, sizes[1] = lv.nLocal, sizes[2] = static_cast<int64_t>(lv.peers.size());
            sizes[3] = lv.nColors, sizes[4] = static_cast<int64_t>(lv.sendSlots.size());
        }
        if (slot2vert) std::copy(lv.slot2vert.begin(), lv.slot2vert.end(), slot2vert);
        if (peers) std::copy(lv.peers.begin(), lv.peers.end(), peers);
        if (sendOffsets) std::copy(lv.sendOffsets.begin(), lv.sendOffsets.end(), sendOffsets);
        if (sendVerts)
            for (size_t i = 0; i < lv.sendSlots.size(); ++i) sendVerts[i] = lv.slot2vert[lv.sendSlots[i]];
        if (recvSlot) std::copy(lv.recvSlot.begin(), lv.recvSlot.end(), recvSlot);
        if (recvCount) std::copy(lv.recvCount.begin(), lv.recvCount.end(), recvCount);
    });
}

int spgx_xpbd_partition_plan(int nVerts, int nSets, const int32_t *const *entries, const int64_t *nEntries,
                             const int32_t *stencilSizes, const int32_t *owner, int rank, int nranks, int64_t *sizes,
                             int32_t *halo, int32_t *peers, int64_t *heldOffsets, int32_t *held, int64_t *sendOffsets,
                             int32_t *sendVerts, int64_t *recvOffsets, int32_t *recvVerts)
{
    return guarded([&] {
        std::vector<FlatStencilsView> sets;
        std::vector<std::vector<int32_t>> colors;
        std::vector<const int32_t *> colorPtrs;
        std::vector<int32_t> nColors;
        for (int s = 0; s < nSets; ++s) {
            sets.push_back({entries[s], nEntries[s], stencilSizes[s]});
            colors.push_back(colorStencils(sets.back()));  // xpbd.cpp:133-160
            int32_t nc = 0;
            for (int32_t c : colors.back()) nc = std::max(nc, c + 1);
            nColors.push_back(nc);
        }
        for (auto &c : colors) colorPtrs.push_back(c.data());
        const XpbdPlan plan = buildXpbdPlan(nVerts, sets, colorPtrs, nColors, owner, rank, nranks);
        if (sizes) {
            sizes[0] = plan.nOwned, sizes[1] = static_cast<int64_t>(plan.halo.size()), sizes[2] = static_cast<int64_t>(plan.peers.size());
            sizes[3] = plan.nPasses, sizes[4] = static_cast<int64_t>(plan.held.size());
            sizes[5] = static_cast<int64_t>(plan.sendVerts.size()), sizes[6] = static_cast<int64_t>(plan.recvVerts.size());
        }
        if (halo) std::copy(plan.halo.begin(), plan.halo.end(), halo);
        if (peers) std::copy(plan.peers.begin(), plan.peers.end(), peers);
        if (heldOffsets) std::copy(plan.heldOffsets.begin(), plan.heldOffsets.end(), heldOffsets);
        if (held) std::copy(plan.held.begin(), plan.held.end(), held);
        if (sendOffsets) std::copy(plan.sendOffsets.begin(), plan.sendOffsets.end(), sendOffsets);
        if (sendVerts) std::copy(plan.sendVerts.begin(), plan.sendVerts.end(), sendVerts);
        if (recvOffsets) std::copy(plan.recvOffsets.begin(), plan.recvOffsets.end(), recvOffsets);
        if (recvVerts) std::copy(plan.recvVerts.begin(), plan.recvVerts.end(), recvVerts);
    });
}

void spgx_ctx_destroy(spgx_ctx *ctx)
{
    if (!ctx) return;
    commDestroy(ctx);
    if (ctx->h2dStream) cudaStreamDestroy(ctx->h2dStream);
    if (ctx->d2hStream) cudaStreamDestroy(ctx->d2hStream);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int spgx_ctx_sync(spgx_ctx *ctx)
{
    return guarded([&] {
        if (!ctx) throw StateError("null context");
        SPGX_CUDA(cudaSetDevice(ctx->device));
        SPGX_CUDA(cudaStreamSynchronize(ctx->stream));
        if (ctx->h2dStream) SPGX_CUDA(cudaStreamSynchronize(ctx->h2dStream));
        if (ctx->d2hStream) SPGX_CUDA(cudaStreamSynchronize(ctx->d2hStream));
        for (spgx_group *g : ctx->groups) throwIfHaloTimedOut(g);
    });
}

int spgx_host_register(void *ptr, uint64_t bytes)
{
    return guarded([&] { SPGX_CUDA(cudaHostRegister(ptr, bytes, cudaHostRegisterDefault)); });
}
int spgx_host_unregister(void *ptr)
{
    return guarded([&] { SPGX_CUDA(cudaHostUnregister(ptr)); });
}
int spgx_host_alloc(uint64_t bytes, void **ptr)
{
    return guarded([&] {
        if (!ptr) throw StateError("null argument");
        *ptr = nullptr;
        if (bytes) SPGX_CUDA(cudaHostAlloc(ptr, bytes, cudaHostAllocPortable));
    });
}
int spgx_host_free(void *ptr)
{
    return guarded([&] {
        if (ptr) SPGX_CUDA(cudaFreeHost(ptr));
    });
}

int spgx_group_create(spgx_ctx *ctx, int nVerts, const double *x, const double *x0, const double *mass,
                      spgx_group **out)
{
    return guarded([&] {
        if (!ctx) throw StateError("null context");
        if (nVerts < 0) throw std::runtime_error("negative vertex count");
        auto *g = new spgx_group;
        g->ctx = ctx;
        g->nVerts = nVerts;
        g->xInitial.assign(x, x + 3 * static_cast<size_t>(nVerts));
        g->xPending = g->xInitial;
        g->vPending.assign(3 * static_cast<size_t>(nVerts), 0.0);
        g->x0.assign(x0, x0 + 3 * static_cast<size_t>(nVerts));
        g->mass.assign(mass, mass + nVerts);
        ctx->groups.push_back(g);
        *out = g;
    });
}

void spgx_group_destroy(spgx_group *grp)
{
    if (!grp) return;
    cudaSetDevice(grp->ctx->device);
    auto &gs = grp->ctx->groups;
    gs.erase(std::remove(gs.begin(), gs.end(), grp), gs.end());
    delete grp;
}

int spgx_group_add_energy(spgx_group *grp, int kind, int nStencils, const int32_t *stencils,
                          const double *modelParams, int *energyId)
{
    return guarded([&] {
        checkKind(kind);
        if (static_cast<int>(grp->energies.size()) >= kMaxEnergies)
            throw std::runtime_error("too many energies in one particle group");
        EnergyState e;
        e.kind = kind;
        e.S = kindStencilSize(kind);
        e.NP = kindNumParams(kind);
        e.NC = kindNumConstraints(kind);
        e.n = nStencils;
        e.stencils.assign(stencils, stencils + static_cast<size_t>(nStencils) * e.S);
        e.records.resize(static_cast<size_t>(nStencils) * e.NP);
        const int NM = kindNumModelParams(kind);
        for (int i = 0; i < nStencils; ++i) {
            for (int s = 0; s < e.S; ++s) {
                const int v = stencils[static_cast<size_t>(i) * e.S + s];
                if (v < 0 || v >= grp->nVerts) throw std::runtime_error("stencil vertex index out of range");
            }
            prepareStencil(kind, stencils + static_cast<size_t>(i) * e.S, modelParams + static_cast<size_t>(i) * NM,
                           grp->x0.data(), e.records.data() + static_cast<size_t>(i) * e.NP);
        }
        grp->energies.push_back(std::move(e));
        grp->finalized = false;
        if (energyId) *energyId = static_cast<int>(grp->energies.size()) - 1;
    });
}

int spgx_group_remove_energy(spgx_group *grp, int energyId)
{
    return guarded([&] {
        if (energyId < 0 || energyId >= static_cast<int>(grp->energies.size())) throw std::runtime_error("energy id out of range");
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        SPGX_CUDA(cudaStreamSynchronize(grp->ctx->stream));  // a step still in flight reads the energy's device buffers
        grp->energies.erase(grp->energies.begin() + energyId);
        grp->finalized = false;
    });
}

int spgx_energy_get_records(spgx_group *grp, int energyId, double *records)
{
    return guarded([&] {
        const auto &e = grp->energies.at(energyId);
        std::copy(e.records.begin(), e.records.end(), records);
    });
}

int spgx_energy_update_record(spgx_group *grp, int energyId, int stencil, int offset, const double *values, int count)
{
    return guarded([&] {
        auto &e = grp->energies.at(energyId);
        if (stencil < 0 || stencil >= e.n || offset < 0 || offset + count > e.NP)
            throw std::runtime_error("record update out of range");
        std::copy(values, values + count, e.records.begin() + static_cast<size_t>(stencil) * e.NP + offset);
        if (grp->finalized) {
            SPGX_CUDA(cudaSetDevice(grp->ctx->device));
            const int pos = e.positionOf.empty() ? stencil : e.positionOf[stencil];
            if (pos >= 0) {  // a partitioned rank keeps only the stencils touching its vertices
                SPGX_CUDA(cudaMemcpyAsync(e.d_records.ptr + static_cast<size_t>(pos) * e.NP + offset, values,
                                          count * sizeof(double), cudaMemcpyHostToDevice, grp->ctx->stream));
                SPGX_CUDA(cudaStreamSynchronize(grp->ctx->stream));
            }
            if (e.streamed) patchStreamRecords(*grp, e, stencil);
        }
    });
}

}  // extern "C"

namespace
{
// Domain-decomposed XPBD (xpbd_partition.h; reference: the single-process pass loop of xpbd.cpp:82-103): global stencil
// colours (bit-exact), the stencils this rank holds in (colour, id) order, local slots [owned ascending | halo ascending],
// and per (energy, colour) pass the owned slots to send to / the halo slots to receive from every neighbour.
void finalizeXpbdPartitioned(spgx_group *grp, const std::vector<FlatStencilsView> &sets, cudaStream_t st)
{
    const int V = grp->nVerts;
    const size_t nE = grp->energies.size();
    std::vector<const int32_t *> colorPtrs;
    std::vector<int32_t> nColors;
    for (size_t ei = 0; ei < nE; ++ei) {
        auto &e = grp->energies[ei];
        e.stencilColors = colorStencils(sets[ei]);  // xpbd.cpp:133-160, global: identical on every rank
        int nc = 0;
        for (int32_t c : e.stencilColors) nc = std::max(nc, c + 1);
        colorPtrs.push_back(e.stencilColors.data());
        nColors.push_back(nc);
    }
    const XpbdPlan plan = buildXpbdPlan(V, sets, colorPtrs, nColors, grp->owner.data(), grp->rank, grp->nRanks);
    if (static_cast<int>(plan.peers.size()) > kMaxPeers) throw std::runtime_error("partitioned XPBD: too many neighbouring ranks");
    grp->nOwned = plan.nOwned;
    grp->nLocal = plan.nOwned + static_cast<int>(plan.halo.size());
    grp->slot2vert.clear();
    grp->vert2slot.assign(static_cast<size_t>(V), -1);
    for (int v = 0; v < V; ++v)
        if (grp->owner[v] == grp->rank) {
            grp->vert2slot[v] = static_cast<int32_t>(grp->slot2vert.size());
            grp->slot2vert.push_back(v);
        }
    for (int32_t v : plan.halo) {
        grp->vert2slot[v] = static_cast<int32_t>(grp->slot2vert.size());
        grp->slot2vert.push_back(v);
    }
    grp->peers.assign(plan.peers.begin(), plan.peers.end());
    grp->vertexColors.clear();
    grp->colorOffsets.clear();
    grp->sendOffsets.clear(), grp->recvSlot.clear(), grp->recvCount.clear();
    auto &xh = grp->xh;
    xh = spgx_group::XpbdHaloState{};
    xh.nPasses = plan.nPasses;
    xh.halo = plan.halo;
    xh.sendOffsets = plan.sendOffsets;
    xh.recvOffsets = plan.recvOffsets;
    xh.sendVerts = plan.sendVerts;
    for (int32_t v : plan.sendVerts) xh.sendSlots.push_back(grp->vert2slot[v]);
    for (int32_t v : plan.recvVerts) xh.recvSlots.push_back(grp->vert2slot[v]);
    grp->p2p.sendSlots = xh.sendSlots;
    grp->d_sendSlots.upload(xh.sendSlots, st);
    grp->d_sendBuf.alloc(3 * xh.sendSlots.size());
    xh.d_recvSlots.upload(xh.recvSlots, st);
    xh.d_recvBuf.alloc(3 * xh.recvSlots.size());
    grp->haloBytesPerSweep = static_cast<int64_t>(24 * xh.sendSlots.size());
    int passBase = 0;
    for (size_t ei = 0; ei < nE; ++ei) {
        auto &e = grp->energies[ei];
        const int nc = nColors[ei];
        e.colorOffsets.assign(static_cast<size_t>(nc) + 1, 0);
        e.order.clear();
        for (int c = 0; c < nc; ++c) {
            const int64_t a = plan.heldOffsets[passBase + c], b = plan.heldOffsets[passBase + c + 1];
            e.order.insert(e.order.end(), plan.held.begin() + a, plan.held.begin() + b);
            e.colorOffsets[c + 1] = static_cast<int64_t>(e.order.size());
        }
        passBase += nc;
        const size_t nHeld = e.order.size();
        e.positionOf.assign(static_cast<size_t>(e.n), -1);
        std::vector<int32_t> st2(nHeld * e.S);
        std::vector<double> rec2(nHeld * e.NP);
        for (size_t pos = 0; pos < nHeld; ++pos) {
            const int i = e.order[pos];
            e.positionOf[i] = static_cast<int32_t>(pos);
            for (int l = 0; l < e.S; ++l) st2[pos * e.S + l] = grp->vert2slot[e.stencils[static_cast<size_t>(i) * e.S + l]];
            std::copy(e.records.begin() + static_cast<size_t>(i) * e.NP, e.records.begin() + static_cast<size_t>(i + 1) * e.NP,
                      rec2.begin() + pos * e.NP);
        }
        e.d_stencils.upload(st2, st);
        e.d_records.upload(rec2, st);
        e.d_lambda.alloc(nHeld * e.NC);
        e.d_stencilW.alloc(st2.size());
        e.d_rowptr.release();
        e.d_entries.release();
        e.streamed = false;
        SPGX_CUDA(cudaStreamSynchronize(st));
    }
}
}  // namespace

extern "C" {

int spgx_group_finalize(spgx_group *grp, int solverKind)
{
    return guarded([&] {
        if (solverKind != SPGX_VBD && solverKind != SPGX_XPBD && solverKind != SPGX_XPBD_SERIAL && solverKind != SPGX_NEWTON)
            throw std::runtime_error("unknown solver kind");
        SPGX_CUDA(cudaSetDevice(grp->ctx->device));
        cudaStream_t st = grp->ctx->stream;
        const int V = grp->nVerts;
        // preserve current state across a re-finalize: every slot the OLD layout holds, owned and halo (a partitioned rank's
        // spgx_download returns owned entries only; its halo copies would restart from the initial state and the first
        // colour passes of the next step would read them before the owners' pushes arrive)
        std::vector<double> xh, vh;
        if ((grp->finalized || grp->d_x.ptr) && grp->d_x.ptr && grp->d_v.ptr) {
            xh = grp->xInitial;
            vh.assign(3 * static_cast<size_t>(V), 0.0);
            const size_t nOld = grp->d_x.count / 3;
            std::vector<double> xs(3 * nOld), vs(3 * nOld);
            SPGX_CUDA(cudaMemcpyAsync(xs.data(), grp->d_x.ptr, xs.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
            SPGX_CUDA(cudaMemcpyAsync(vs.data(), grp->d_v.ptr, vs.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
            SPGX_CUDA(cudaStreamSynchronize(st));
            for (size_t s = 0; s < nOld && s < grp->slot2vert.size(); ++s) {
                const size_t o = 3 * static_cast<size_t>(grp->slot2vert[s]);
                for (int c = 0; c < 3; ++c) xh[o + c] = xs[3 * s + c], vh[o + c] = vs[3 * s + c];
            }
        } else {
            xh = grp->xPending;
            vh = grp->vPending;
        }
        grp->solverKind = solverKind;
        grp->identityOrder = false;
        if (grp->graphExec) {
            cudaGraphExecDestroy(grp->graphExec);
            grp->graphExec = nullptr;
        }
        const size_t nE = grp->energies.size();
        std::vector<FlatStencilsView> sets;
        for (auto &e : grp->energies) sets.push_back({e.stencils.data(), static_cast<int64_t>(e.stencils.size()), e.S});

        grp->slot2vert.assign(V, 0);
        grp->vert2slot.assign(V, 0);
        grp->uniqueNeighbourSum = 0;
        if (solverKind == SPGX_VBD) {
            PhaseTimer pt;
            std::vector<Incidence> inc;
            // vbd.cpp:203-229 (global: identical on every rank).  Large groups are coloured on the device (colour.cu: the same
            // colours bit for bit, tests/test_gpu_colouring.py); SPGX_DEVICE_COLOURING=0 / 1 forces the host / device form
            bool onDevice = V >= 500000;  // 3 x faster than the host pass at 1 M vertices, 6 x at 16 M (profiles/r2o_device_colouring.txt)
            if (const char *env = getenv("SPGX_DEVICE_COLOURING")) onDevice = atoi(env) != 0;
            grp->colouringRounds = 0;
            if (!onDevice)
                for (auto &fs : sets) inc.push_back(buildIncidence(V, fs));
            grp->vertexColors = onDevice ? colorVerticesDevice(grp->ctx, V, sets, &grp->colouringRounds) : colorVertices(V, sets, &inc);
            pt.lap("colouring (+ host incidence)");
            LocalView lv = buildLocalView(V, grp->vertexColors, sets, grp->partitioned ? grp->owner.data() : nullptr,
                                          grp->rank, grp->nRanks);
            pt.lap("local view");
            // vertex -> (stencil, local) incidence (vbd.cpp:231-268): the host colouring needed the global table; otherwise only
            // the rows of owned vertices are read below, and those are complete over the stencils this rank keeps
            if (onDevice) {
                inc.resize(nE);  // one host thread per energy
                parallelChunks(static_cast<int64_t>(nE), static_cast<int>(nE), [&](int64_t a, int64_t b, int) {
                    for (int64_t ei = a; ei < b; ++ei)
                        inc[ei] = grp->partitioned ? buildIncidenceOf(V, sets[ei], lv.localStencils[ei]) : buildIncidence(V, sets[ei]);
                });
            }
            grp->nOwned = lv.nOwned;
            grp->nLocal = lv.nLocal;
            grp->slot2vert = lv.slot2vert;
            grp->vert2slot = lv.vert2slot;
            grp->colorOffsets = lv.colorOffsets;
            const int nOwned = lv.nOwned, nLocal = lv.nLocal;
            // halo exchange tables
            grp->peers = lv.peers;
            grp->sendOffsets = lv.sendOffsets;
            grp->recvSlot = lv.recvSlot;
            grp->recvCount = lv.recvCount;
            grp->p2p.sendSlots = lv.sendSlots;
            grp->p2p.pushLocal.clear();
            grp->p2p.pushWarps.clear();
            grp->d_sendSlots.upload(lv.sendSlots, st);
            grp->d_sendBuf.alloc(3 * lv.sendSlots.size());
            SPGX_CUDA(cudaStreamSynchronize(st));
            {
                // threads per vertex (measured on B200, profiles/): two lanes per vertex for big colours (shorter
                // per-warp chains, half the neighbour tile, more resident slices), four below 100 k vertices per colour
                int64_t largest = 0;
                for (size_t c = 0; c + 1 < grp->colorOffsets.size(); ++c)
                    largest = std::max(largest, grp->colorOffsets[c + 1] - grp->colorOffsets[c]);
                grp->vbdLanes = largest >= 100000 ? 2 : 4;
            }
            if (const char *env = getenv("SPGX_VBD_LANES")) {
                const int l = atoi(env);
                if (l == 2 || l == 4) grp->vbdLanes = l;
            }
            // kernel variant (vbd_stream.cuh launchStreamG): resident blocks per SM + 10 x "records through the shared-memory ring"
            grp->vbdOcc = 0;  // 0: chosen once the streams are built (4, or 14 for a group with wedge records)
            if (const char *env = getenv("SPGX_VBD_VARIANT")) {
                const int o = atoi(env);
                if (o == 4 || o == 14 || o == 15) grp->vbdOcc = o;
            }
            grp->vbdNbrFixed = 1;
            if (const char *env = getenv("SPGX_VBD_NBRFIXED")) grp->vbdNbrFixed = atoi(env) != 0;
            grp->vbdWedge = 1;
            if (const char *env = getenv("SPGX_VBD_WEDGE")) grp->vbdWedge = atoi(env) != 0;
            grp->sliceBase.assign(grp->colorOffsets.size(), 0);
            for (size_t c = 0; c + 1 < grp->colorOffsets.size(); ++c) {
                const int64_t n = grp->colorOffsets[c + 1] - grp->colorOffsets[c];
                const int VW = 32 / grp->vbdLanes;
                grp->sliceBase[c + 1] = grp->sliceBase[c] + static_cast<int32_t>((n + VW - 1) / VW);
            }
            // incidence in slot order (vbd.cpp:231-268 keeps ascending flat index per vertex and energy): the host tables of the
            // energies are built on one thread each, then uploaded
            struct EnergyTables {
                std::vector<uint32_t> rowptr, entries;
                std::vector<int32_t> slotStencils;
                std::vector<double> keptRecords;
            };
            std::vector<EnergyTables> tables(nE);
            parallelChunks(static_cast<int64_t>(nE), static_cast<int>(nE), [&](int64_t ea, int64_t eb, int) {
                for (int64_t ei = ea; ei < eb; ++ei) {
                    auto &e = grp->energies[ei];
                    auto &tb = tables[ei];
                    if (static_cast<uint64_t>(e.n) >= (1ull << 30)) throw std::runtime_error("too many stencils in one energy");
                    // global stencil id -> position among the stencils this rank keeps (-1: not kept)
                    const auto &kept = lv.localStencils[ei];
                    e.positionOf.assign(static_cast<size_t>(e.n), -1);
                    for (size_t k = 0; k < kept.size(); ++k) e.positionOf[kept[k]] = static_cast<int32_t>(k);
                    tb.rowptr.assign(static_cast<size_t>(nLocal) + 1, 0);
                    uint32_t cursor = 0;
                    for (int s = 0; s < nLocal; ++s) {
                        const int v = grp->slot2vert[s];
                        tb.rowptr[s] = cursor;
                        if (s >= nOwned) continue;  // halo vertices are never updated here
                        for (int64_t q = inc[ei].rowptr[v]; q < inc[ei].rowptr[v + 1]; ++q) {
                            const int64_t flat = inc[ei].flat[q];
                            tb.entries.push_back((static_cast<uint32_t>(e.positionOf[flat / e.S]) << 2) | static_cast<uint32_t>(flat % e.S));
                            ++cursor;
                        }
                    }
                    tb.rowptr[nLocal] = cursor;
                    e.ownedIncidences = cursor;
                    tb.slotStencils.resize(kept.size() * e.S);
                    tb.keptRecords.resize(kept.size() * e.NP);
                    for (size_t k = 0; k < kept.size(); ++k) {
                        for (int l = 0; l < e.S; ++l)
                            tb.slotStencils[k * e.S + l] = grp->vert2slot[e.stencils[static_cast<size_t>(kept[k]) * e.S + l]];
                        std::copy(e.records.begin() + static_cast<size_t>(kept[k]) * e.NP,
                                  e.records.begin() + static_cast<size_t>(kept[k] + 1) * e.NP, tb.keptRecords.begin() + k * e.NP);
                    }
                }
            });
            for (size_t ei = 0; ei < nE; ++ei) {
                auto &e = grp->energies[ei];
                auto &tb = tables[ei];
                e.d_stencils.upload(tb.slotStencils, st);
                e.d_records.upload(tb.keptRecords, st);
                e.d_rowptr.upload(tb.rowptr, st);
                e.d_entries.upload(tb.entries, st);
                e.streamed = false;
                e.order.clear();
                SPGX_CUDA(cudaStreamSynchronize(st));
            }
            tables.clear();
            pt.lap("incidence + CSR + uploads");
            // streaming layout, when the group's mix of energies has a compiled streaming kernel
            grp->streams.clear();
            grp->streamMask = 0;
            grp->maxNbr = 0;
            grp->wedgeRecords = 0;
            grp->streamBytes = 0;
            grp->d_sliceDesc.release();
            grp->d_nbrIds.release();
            grp->nSlices = 0;
            grp->vbdLookahead = 512;
            if (const char *env = getenv("SPGX_VBD_LOOKAHEAD")) grp->vbdLookahead = std::max(0, atoi(env));
            grp->vbdPdl = 1;
            if (const char *env = getenv("SPGX_VBD_PDL")) grp->vbdPdl = atoi(env) != 0;
            grp->vbdCarveout = -1;
            if (const char *env = getenv("SPGX_VBD_CARVEOUT")) grp->vbdCarveout = std::min(100, std::max(-1, atoi(env)));
            {
                // L2 set-aside for the persisting window over the positions (api_step.cu): SPGX_VBD_L2PERSIST_MB, 0 = off
                int mb = 32;  // measured on config 3: 16 / 32 / 48 / 64 / 96 MB -> 2.756 / 2.709 / 2.711 / 2.728 / 2.736 ms (off: 2.798)
                if (const char *env = getenv("SPGX_VBD_L2PERSIST_MB")) mb = std::max(0, atoi(env));
                size_t want = static_cast<size_t>(mb) << 20;
                want = std::min(want, static_cast<size_t>(grp->ctx->prop.persistingL2CacheMaxSize));
                if (want) SPGX_CUDA(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want));
                grp->vbdL2Persist = static_cast<int64_t>(want);
            }
            {
                unsigned smask = 0, kmask = 0;
                for (auto &e : grp->energies) {
                    bool any = false;
                    for (int l = 0; l < e.S; ++l)
                        if (streamTypeOf(e.kind, l) >= 0) smask |= 1u << streamTypeOf(e.kind, l), any = true;
                    if (!any) kmask |= 1u << e.kind;
                }
                if (!getenv("SPGX_NO_STREAMS") && vbdStreamComboSupported(smask, kmask)) {
                    if (!buildStreams(*grp, inc, st)) {
                        grp->streams.clear();
                        grp->streamMask = 0;
                        for (auto &e : grp->energies) e.streamed = false;
                    }
                }
            }
            if (grp->vbdOcc == 0) grp->vbdOcc = grp->wedgeRecords > 0 ? 14 : 4;  // vbd_stream.cuh: launchStreamG
            pt.lap("stream layout (buildStreams)");
            // distinct stencil neighbours per vertex (for the byte model)
            {
                const int nChunks = hostThreads();
                std::vector<double> partial(static_cast<size_t>(nChunks), 0.0);
                parallelChunks(V, nChunks, [&](int64_t a, int64_t b, int c) {
                    std::vector<int32_t> nb;
                    double sum = 0;
                    for (int64_t v = a; v < b; ++v) {
                        if (grp->vert2slot[v] < 0 || grp->vert2slot[v] >= nOwned) continue;
                        nb.clear();
                        for (size_t ei = 0; ei < nE; ++ei) {
                            const auto &e = grp->energies[ei];
                            for (int64_t q = inc[ei].rowptr[v]; q < inc[ei].rowptr[v + 1]; ++q) {
                                const int64_t base = (inc[ei].flat[q] / e.S) * e.S;
                                for (int l = 0; l < e.S; ++l)
                                    if (e.stencils[base + l] != v) nb.push_back(e.stencils[base + l]);
                            }
                        }
                        std::sort(nb.begin(), nb.end());
                        sum += static_cast<double>(std::unique(nb.begin(), nb.end()) - nb.begin());
                    }
                    partial[static_cast<size_t>(c)] = sum;
                });
                for (double p : partial) grp->uniqueNeighbourSum += p;  // integer-valued: the order of the sum is immaterial
            }
            pt.lap("unique-neighbour count");
        } else if (grp->partitioned && solverKind == SPGX_XPBD) {
            finalizeXpbdPartitioned(grp, sets, st);
        } else {
            if (grp->partitioned)
                throw std::runtime_error("domain decomposition is implemented for the VBD and the (parallel Gauss-Seidel) XPBD solver only");
            grp->identityOrder = true;
            grp->xh = spgx_group::XpbdHaloState{};
            {
                CgState *pinned = grp->newton.hostState;
                double *pinnedScalar = grp->newton.hostScalar;
                cudaEvent_t ev[3] = {grp->newton.ev[0], grp->newton.ev[1], grp->newton.ev[2]};
                grp->newton = spgx_group::NewtonState{};
                grp->newton.hostState = pinned;
                grp->newton.hostScalar = pinnedScalar;
                for (int k = 0; k < 3; ++k) grp->newton.ev[k] = ev[k];
            }
            grp->nOwned = grp->nLocal = V;
            grp->peers.clear();
            for (int s = 0; s < V; ++s) grp->slot2vert[s] = grp->vert2slot[s] = s;
            grp->vertexColors.clear();
            grp->colorOffsets.clear();
            for (size_t ei = 0; ei < nE; ++ei) {
                auto &e = grp->energies[ei];
                if (solverKind == SPGX_XPBD) {
                    e.stencilColors = colorStencils(sets[ei]);  // xpbd.cpp:133-160
                    groupByKey(e.stencilColors, e.order, e.colorOffsets);
                } else {
                    e.stencilColors.assign(e.n, 0);
                    e.order.resize(e.n);
                    for (int i = 0; i < e.n; ++i) e.order[i] = i;
                    e.colorOffsets = {0, e.n};
                }
                e.positionOf.resize(e.n);
                std::vector<int32_t> st2(e.stencils.size());
                std::vector<double> rec2(e.records.size());
                for (int pos = 0; pos < e.n; ++pos) {
                    const int i = e.order[pos];
                    e.positionOf[i] = pos;
                    std::copy(e.stencils.begin() + static_cast<size_t>(i) * e.S,
                              e.stencils.begin() + static_cast<size_t>(i + 1) * e.S, st2.begin() + static_cast<size_t>(pos) * e.S);
                    std::copy(e.records.begin() + static_cast<size_t>(i) * e.NP,
                              e.records.begin() + static_cast<size_t>(i + 1) * e.NP, rec2.begin() + static_cast<size_t>(pos) * e.NP);
                }
                e.d_stencils.upload(st2, st);
                e.d_records.upload(rec2, st);
                e.d_lambda.alloc(solverKind == SPGX_NEWTON ? 0 : static_cast<size_t>(e.n) * e.NC);
                e.d_stencilW.alloc(solverKind == SPGX_NEWTON ? 0 : st2.size());
                e.d_rowptr.release();
                e.d_entries.release();
                SPGX_CUDA(cudaStreamSynchronize(st));
            }
            if (solverKind == SPGX_NEWTON) newtonBuild(grp, st);
        }
        // state: local vertices (owned + halo) on the device; the staging buffer spans the global index range
        const int nLocalAll = grp->nLocal;
        const size_t n3 = 3 * static_cast<size_t>(nLocalAll);
        // [x | xInertial | mass] contiguous (x first: its address is what the peer-memory exchange exports)
        grp->d_x.release(), grp->d_xInertial.release(), grp->d_mass.release();
        grp->d_stateArena.alloc(2 * n3 + static_cast<size_t>(nLocalAll) + 2);
        grp->d_x.view(grp->d_stateArena.ptr, n3);
        grp->d_xInertial.view(grp->d_stateArena.ptr + n3, n3);
        grp->d_mass.view(grp->d_stateArena.ptr + 2 * n3, static_cast<size_t>(nLocalAll));
        grp->d_v.alloc(n3);
        grp->d_xOld.alloc(n3);
        grp->d_stage.alloc(3 * static_cast<size_t>(V));
        grp->d_g.alloc(3);
        grp->d_eval.alloc(64);
        std::vector<double> ms(nLocalAll), ws(nLocalAll);
        for (int s = 0; s < nLocalAll; ++s) {
            const double m = grp->mass[grp->slot2vert[s]];
            ms[s] = m;
            ws[s] = (m == 0) ? 0 : 1 / m;  // particleGroup.cpp:81
        }
        grp->d_mass.upload(ms, st);
        grp->d_invMass.upload(ws, st);
        if (solverKind == SPGX_XPBD || solverKind == SPGX_XPBD_SERIAL)
            for (auto &e : grp->energies)  // inverse masses per stencil vertex: a pass requests them with the ids, not after them
                launchGatherScalar(grp->d_invMass.ptr, e.d_stencils.ptr, static_cast<long long>(e.d_stencilW.count), e.d_stencilW.ptr, st);
        // global index range the local vertices span (upload / download move only that range)
        grp->localLo = V;
        grp->localHi = 0;
        grp->ownedLo = V;
        grp->ownedHi = 0;
        for (int s = 0; s < nLocalAll; ++s) {
            const int v = grp->slot2vert[s];
            grp->localLo = std::min(grp->localLo, v);
            grp->localHi = std::max(grp->localHi, v + 1);
            if (s < grp->nOwned) {
                grp->ownedLo = std::min(grp->ownedLo, v);
                grp->ownedHi = std::max(grp->ownedHi, v + 1);
            }
        }
        if (nLocalAll == 0) grp->localLo = grp->localHi = grp->ownedLo = grp->ownedHi = 0;
        grp->ownedContiguous = (grp->ownedHi - grp->ownedLo) == grp->nOwned;
        grp->d_slot2vert.upload(grp->slot2vert, st);
        SPGX_CUDA(cudaStreamSynchronize(st));
        p2pPrepare(grp, st);  // the position array is new: earlier peer mappings of it are void
        if (!grp->evStart) {
            SPGX_CUDA(cudaEventCreate(&grp->evStart));
            SPGX_CUDA(cudaEventCreate(&grp->evStop));
        }
        grp->finalized = true;
        if (spgx_upload(grp, xh.data(), vh.data()) != SPGX_OK) throw CudaError(spgx_last_error());
        grp->xPending.clear();
        grp->vPending.clear();
    });
}

int spgx_group_num_colors(spgx_group *grp, int energyId)
{
    if (!grp || !grp->finalized) return -1;
    if (grp->solverKind == SPGX_VBD) return static_cast<int>(grp->colorOffsets.size()) - 1;
    if (energyId < 0 || energyId >= static_cast<int>(grp->energies.size())) return -1;
    return static_cast<int>(grp->energies[energyId].colorOffsets.size()) - 1;
}

int spgx_get_vertex_colors(spgx_group *grp, int32_t *colors)
{
    return guarded([&] {
        if (!grp->finalized || grp->solverKind != SPGX_VBD) throw StateError("group is not finalized for VBD");
        std::copy(grp->vertexColors.begin(), grp->vertexColors.end(), colors);
    });
}

int spgx_get_vertex_incidence(spgx_group *grp, int64_t *rowptr, int32_t *entries, int64_t *nEntries)
{
    return guarded([&] {
        const int V = grp->nVerts;
        const size_t nE = grp->energies.size();
        std::vector<Incidence> inc;
        for (auto &e : grp->energies)
            inc.push_back(buildIncidence(V, {e.stencils.data(), static_cast<int64_t>(e.stencils.size()), e.S}));
        int64_t n = 0;
        for (int v = 0; v < V; ++v)
            for (size_t ei = 0; ei < nE; ++ei) {
                if (rowptr) rowptr[static_cast<size_t>(v) * nE + ei] = n;
                const int S = grp->energies[ei].S;
                for (int64_t q = inc[ei].rowptr[v]; q < inc[ei].rowptr[v + 1]; ++q) {
                    if (entries) {
                        entries[2 * n] = static_cast<int32_t>(inc[ei].flat[q] / S);
                        entries[2 * n + 1] = static_cast<int32_t>(inc[ei].flat[q] % S);
                    }
                    ++n;
                }
            }
        if (rowptr) rowptr[static_cast<size_t>(V) * nE] = n;
        if (nEntries) *nEntries = n;
    });
}

int spgx_get_stencil_colors(spgx_group *grp, int energyId, int32_t *colors)
{
    return guarded([&] {
        if (!grp->finalized || grp->solverKind == SPGX_VBD) throw StateError("group is not finalized for XPBD");
        const auto &e = grp->energies.at(energyId);
        std::copy(e.stencilColors.begin(), e.stencilColors.end(), colors);
    });
}

int spgx_model_bytes(spgx_group *grp, double *bytesPerIteration, double *bytesLowerBound)
{
    return guarded([&] {
        if (!grp->finalized) throw StateError("group is not finalized");
        const double V = grp->nOwned;
        double gather = 0, lower = 0;
        if (grp->solverKind == SPGX_VBD) {
            gather = 92.0 * V + 24.0 * grp->uniqueNeighbourSum;
            lower = (24.0 + 24.0 + 24.0 + 8.0) * V;
            for (const auto &e : grp->energies) {
                const double incidences = static_cast<double>(e.ownedIncidences);
                gather += incidences * (4.0 + modelRecordBytes(e.kind));
                lower += modelRecordBytes(e.kind) * (grp->partitioned ? incidences / e.S : static_cast<double>(e.n)) + 4.0 * incidences;
            }
        } else {
            gather = 6.0 * 24.0 * V;
            for (const auto &e : grp->energies)  // a partitioned rank projects the stencils it holds
                gather += static_cast<double>(grp->partitioned ? static_cast<int64_t>(e.order.size()) : e.n) * (modelRecordBytes(e.kind) + e.S * 56.0);
            lower = gather;
        }
        if (bytesPerIteration) *bytesPerIteration = gather;
        if (bytesLowerBound) *bytesLowerBound = lower;
    });
}

}  // extern "C"
