// Host-side replay of the streaming colour pass's record layout (spg_b200/csrc/stream_records.h, the function buildStreams
// and patchStreamRecords call): on triangulated grids -- regular diagonals, alternating diagonals (valence 4 / 8, the
// reference's createCloth pattern), with randomly rotated triangle stencils and randomly reordered hinge stencils -- every
// vertex's records are built with and without wedges, replayed with the canonical evaluators (streams.cuh), and the
// per-vertex gradient and 3 x 3 block compared with the generic second-order-dual evaluation of every incidence (the
// restatement of accumulateVertexForce / accumulateVertexHessian, spg/include/spg/sim/energy/energy.h:240-258).
// Also checked: no incidence is lost or duplicated (3 x wedges + single records == incidences), and every interior vertex
// of a consistently oriented mesh pairs up completely (no single-incidence record left).
// Output: one line "<case> max_rel <value> wedges <n> singles <n> interior_singles <n>"; exit code = number of failed cases.
#include <cmath>
#include <cstdio>
#include <map>
#include <random>
#include <utility>
#include <vector>

#include "../../spg_b200/csrc/energy_prepare.h"
#include "../../spg_b200/csrc/host_build.h"
#include "../../spg_b200/csrc/stream_records.h"

using namespace spgx;
using D2 = Dual<3, true>;

struct Mesh {
    int n = 0;  // n x n vertices
    std::vector<double> x0, x;
    std::vector<int32_t> tris, hinges;  // 3 / 4 ids per stencil
    std::vector<double> recT, recH;
};

static Mesh makeMesh(int n, bool alternate, bool shuffle, std::mt19937_64 &rng)
{
    Mesh m;
    m.n = n;
    std::uniform_real_distribution<double> U(-1.0, 1.0);
    const double h = 1.0 / (n - 1);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
            const double px = j * h + 0.15 * h * U(rng), py = i * h + 0.15 * h * U(rng);
            m.x0.insert(m.x0.end(), {px, py, 0.0});
            m.x.insert(m.x.end(), {px + 0.2 * h * U(rng), py + 0.2 * h * U(rng), 0.3 * h * U(rng)});
        }
    auto id = [n](int i, int j) { return i * n + j; };
    for (int i = 0; i + 1 < n; ++i)
        for (int j = 0; j + 1 < n; ++j) {
            const int a = id(i, j), b = id(i, j + 1), c = id(i + 1, j + 1), d = id(i + 1, j);  // counter-clockwise
            int t[2][3];
            if (alternate && ((i + j) & 1)) {
                const int u[2][3] = {{a, b, d}, {b, c, d}};
                std::copy(&u[0][0], &u[0][0] + 6, &t[0][0]);
            } else {
                const int u[2][3] = {{a, b, c}, {a, c, d}};
                std::copy(&u[0][0], &u[0][0] + 6, &t[0][0]);
            }
            for (auto &tt : t) {
                const int r = shuffle ? static_cast<int>(rng() % 3) : 0;
                for (int k = 0; k < 3; ++k) m.tris.push_back(tt[(k + r) % 3]);
            }
        }
    // hinges: interior edges with the two opposite vertices
    std::map<std::pair<int, int>, std::vector<int>> edgeOpp;
    for (size_t t = 0; t < m.tris.size() / 3; ++t)
        for (int k = 0; k < 3; ++k) {
            const int u = m.tris[3 * t + k], v = m.tris[3 * t + (k + 1) % 3], w = m.tris[3 * t + (k + 2) % 3];
            edgeOpp[{std::min(u, v), std::max(u, v)}].push_back(w);
        }
    for (auto &kv : edgeOpp) {
        if (kv.second.size() != 2) continue;
        int s[4] = {kv.first.first, kv.first.second, kv.second[0], kv.second[1]};
        if (shuffle) {
            if (rng() & 1) std::swap(s[0], s[1]);
            if (rng() & 1) std::swap(s[2], s[3]);
        }
        m.hinges.insert(m.hinges.end(), s, s + 4);
    }
    const double modelT[4] = {100.0, 0.3, 0, 0};
    m.recT.resize(m.tris.size() / 3 * EnergyKind<K_MEMBRANE_STVK>::NP);
    for (size_t t = 0; t < m.tris.size() / 3; ++t)
        prepareStencil(K_MEMBRANE_STVK, m.tris.data() + 3 * t, modelT, m.x0.data(), m.recT.data() + t * EnergyKind<K_MEMBRANE_STVK>::NP);
    m.recH.resize(m.hinges.size() / 4 * EnergyKind<K_DISCRETE_BENDING>::NP);
    for (size_t t = 0; t < m.hinges.size() / 4; ++t) {
        const double modelH[4] = {0.0, 1e-2 * (1.0 + 0.5 * U(rng)), 0, 0};
        double *p = m.recH.data() + t * EnergyKind<K_DISCRETE_BENDING>::NP;
        prepareStencil(K_DISCRETE_BENDING, m.hinges.data() + 4 * t, modelH, m.x0.data(), p);
        p[0] = 0.3 * U(rng);  // a rest angle of either sign
    }
    return m;
}

template <int KIND>
static void addGeneric(const Mesh &m, const int32_t *st, const double *rec, int local, double *g, double *H)
{
    constexpr int S = EnergyKind<KIND>::S;
    V3<D2> xs[S];
    for (int s = 0; s < S; ++s) {
        const bool own = s == local;
        const double *p = m.x.data() + 3 * st[s];
        xs[s] = V3<D2>(D2(p[0], own ? 0 : -1), D2(p[1], own ? 1 : -1), D2(p[2], own ? 2 : -1));
    }
    const D2 E = stencilEnergy<KIND, D2>(xs, rec);
    for (int i = 0; i < 3; ++i) g[i] += E.g[i];
    for (int i = 0; i < 6; ++i) H[i] += E.h[i];
}

static int runCase(const char *name, int n, bool alternate, bool shuffle, std::mt19937_64 &rng)
{
    const Mesh m = makeMesh(n, alternate, shuffle, rng);
    const int V = n * n;
    const Incidence incH = buildIncidence(V, {m.hinges.data(), static_cast<int64_t>(m.hinges.size()), 4});
    const Incidence incT = buildIncidence(V, {m.tris.data(), static_cast<int64_t>(m.tris.size()), 3});
    // energies in the reference scene's insertion order: bending, membrane
    const StreamEnergyView views[2] = {
        {K_DISCRETE_BENDING, 4, EnergyKind<K_DISCRETE_BENDING>::NP, m.hinges.data(), m.recH.data(), incH.rowptr.data(), incH.flat.data()},
        {K_MEMBRANE_STVK, 3, EnergyKind<K_MEMBRANE_STVK>::NP, m.tris.data(), m.recT.data(), incT.rowptr.data(), incT.flat.data()}};
    double worst = 0;
    long wedges = 0, singles = 0, interiorSingles = 0, lost = 0;
    SlotScratch sc;
    for (int pass = 0; pass < 2; ++pass) {  // 0: wedges, 1: single-incidence records only
        for (int v = 0; v < V; ++v) {
            std::vector<uint32_t> idx[ST_COUNT];
            std::vector<double> q[ST_COUNT];
            appendSlotRecords(views, 2, v, pass == 0 ? 0 : -1, pass == 0 ? 1 : -1, sc, idx, q);
            double g[3] = {0, 0, 0}, H[6] = {0, 0, 0, 0, 0, 0};
            V3d gs(0, 0, 0);
            Sym3 Hs{0, 0, 0, 0, 0, 0};
            long records = 0, vw = 0;
            const V3d xo(m.x[3 * v], m.x[3 * v + 1], m.x[3 * v + 2]);
            for (int t = 0; t < ST_COUNT; ++t)
                for (size_t r = 0; r < idx[t].size(); ++r) {
                    V3d xn[4];
                    for (int k = 0; k < streamNI(t); ++k) {
                        const int u = sc.seen[(idx[t][r] >> (8 * k)) & 0xFFu];
                        xn[k] = V3d(m.x[3 * u], m.x[3 * u + 1], m.x[3 * u + 2]);
                    }
                    evalStream(t, xo, xn, q[t].data() + r * streamNQ(t), gs, Hs);
                    records += t == ST_WEDGE ? 3 : 1;
                    vw += t == ST_WEDGE;
                }
            long incidences = 0;
            for (int64_t r = incH.rowptr[v]; r < incH.rowptr[v + 1]; ++r, ++incidences)
                addGeneric<K_DISCRETE_BENDING>(m, m.hinges.data() + (incH.flat[r] / 4) * 4, m.recH.data() + (incH.flat[r] / 4) * EnergyKind<K_DISCRETE_BENDING>::NP,
                                               static_cast<int>(incH.flat[r] % 4), g, H);
            for (int64_t r = incT.rowptr[v]; r < incT.rowptr[v + 1]; ++r, ++incidences)
                addGeneric<K_MEMBRANE_STVK>(m, m.tris.data() + (incT.flat[r] / 3) * 3, m.recT.data() + (incT.flat[r] / 3) * EnergyKind<K_MEMBRANE_STVK>::NP,
                                            static_cast<int>(incT.flat[r] % 3), g, H);
            if (records != incidences) ++lost;
            const int i = v / n, j = v % n;
            const bool interior = i >= 2 && j >= 2 && i < n - 2 && j < n - 2;  // two rings in: every hinge of the fan exists
            if (pass == 0) {
                wedges += vw;
                singles += records - 3 * vw;
                if (interior) interiorSingles += records - 3 * vw;
            } else if (vw != 0)
                ++lost;
            const double a[3] = {gs.x, gs.y, gs.z}, b[6] = {Hs.xx, Hs.xy, Hs.xz, Hs.yy, Hs.yz, Hs.zz};
            double scaleG = 1e-300, scaleH = 1e-300, eg = 0, eh = 0;
            for (int k = 0; k < 3; ++k) scaleG = std::fmax(scaleG, std::fabs(g[k])), eg = std::fmax(eg, std::fabs(g[k] - a[k]));
            for (int k = 0; k < 6; ++k) scaleH = std::fmax(scaleH, std::fabs(H[k])), eh = std::fmax(eh, std::fabs(H[k] - b[k]));
            worst = std::fmax(worst, std::fmax(eg / scaleG, eh / scaleH));
        }
    }
    std::printf("%s max_rel %.3e wedges %ld singles %ld interior_singles %ld lost %ld\n", name, worst, wedges, singles, interiorSingles, lost);
    return (worst < 1e-10 && lost == 0 && interiorSingles == 0 && wedges > 0) ? 0 : 1;
}

int main()
{
    std::mt19937_64 rng(2024);
    int bad = 0;
    bad += runCase("regular", 12, false, false, rng);
    bad += runCase("alternating", 13, true, false, rng);
    bad += runCase("regular_shuffled", 11, false, true, rng);
    bad += runCase("alternating_shuffled", 14, true, true, rng);
    return bad;
}
