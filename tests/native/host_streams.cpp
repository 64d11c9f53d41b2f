// Host-side check of the canonical stream evaluators (spg_b200/csrc/streams.cuh) against the generic second-order
// dual-number evaluation of the same energy (energies.cuh, the restatement of energy.h:240-258) on random stencils,
// for every streamed kind and every owner position.  Compiled with g++ (the evaluators are __host__ __device__).
// Prints "max_rel <value>" per (kind, local) and exits non-zero above the tolerance.
#include <cmath>
#include <cstdio>
#include <random>
#include <type_traits>

#include "../../spg_b200/csrc/energy_prepare.h"
#include "../../spg_b200/csrc/streams.cuh"

using namespace spgx;
using D2 = Dual<3, true>;

template <int KIND>
static double check(std::mt19937_64 &rng, int trials)
{
    using EK = EnergyKind<KIND>;
    constexpr int S = EK::S;
    std::uniform_real_distribution<double> U(-1.0, 1.0);
    double worst = 0;
    for (int t = 0; t < trials; ++t) {
        // rest pose: a well-shaped random stencil (z = 0 for membranes), current pose: perturbed
        double x0[3 * S], x[3 * S];
        const double base[4][3] = {{0, 0, 0}, {1, 0, 0}, {0.3, 0.9, 0}, {0.4, -0.8, 0}};
        const double baseT[4][3] = {{0, 0, 0}, {1, 0, 0}, {0.2, 0.9, 0.1}, {0.3, 0.2, 0.8}};
        for (int s = 0; s < S; ++s)
            for (int c = 0; c < 3; ++c) {
                const bool tet = (KIND == K_TET_SNH || KIND == K_TET_STVK);
                x0[3 * s + c] = (tet ? baseT : base)[s][c] + 0.15 * U(rng);
                if (KIND == K_MEMBRANE_STVK && c == 2) x0[3 * s + c] = 0;
                x[3 * s + c] = x0[3 * s + c] + 0.2 * U(rng);
            }
        int st[S];
        for (int s = 0; s < S; ++s) st[s] = s;
        double model[4] = {0.3 * U(rng), 1.0 + U(rng) * 0.5, 0, 0};
        if (KIND == K_MEMBRANE_STVK) model[0] = 100.0, model[1] = 0.3;
        if (KIND == K_TET_SNH || KIND == K_TET_STVK) model[0] = 1000.0, model[1] = 0.2;
        double rec[EK::NP];
        prepareStencil(KIND, st, model, x0, rec);
        for (int local = 0; local < S; ++local) {
            V3<D2> xs[S];
            for (int s = 0; s < S; ++s) {
                const bool own = s == local;
                xs[s] = V3<D2>(D2(x[3 * s], own ? 0 : -1), D2(x[3 * s + 1], own ? 1 : -1), D2(x[3 * s + 2], own ? 2 : -1));
            }
            const D2 E = stencilEnergy<KIND, D2>(xs, rec);
            const int stype = streamTypeOf(KIND, local);
            int32_t nbr[3];
            double q[16];
            makeStreamRecord(KIND, st, rec, local, nbr, q);
            V3d xn[3];
            for (int k = 0; k < streamNI(stype); ++k) xn[k] = V3d(x[3 * nbr[k]], x[3 * nbr[k] + 1], x[3 * nbr[k] + 2]);
            V3d g(0, 0, 0);
            Sym3 H{0, 0, 0, 0, 0, 0};
            evalStream(stype, V3d(x[3 * local], x[3 * local + 1], x[3 * local + 2]), xn, q, g, H);
            const double gs[3] = {g.x, g.y, g.z};
            const double hs[6] = {H.xx, H.xy, H.xz, H.yy, H.yz, H.zz};
            double scaleG = 1e-300, scaleH = 1e-300, eg = 0, eh = 0;
            for (int i = 0; i < 3; ++i) scaleG = std::fmax(scaleG, std::fabs(E.g[i])), eg = std::fmax(eg, std::fabs(E.g[i] - gs[i]));
            for (int i = 0; i < 6; ++i) scaleH = std::fmax(scaleH, std::fabs(E.h[i])), eh = std::fmax(eh, std::fabs(E.h[i] - hs[i]));
            worst = std::fmax(worst, std::fmax(eg / scaleG, eh / scaleH));
        }
    }
    return worst;
}

// the wedge evaluator (membrane + the two hinges that share its normal) against the sum of the three generic evaluations,
// with the three stencils in every vertex order the canonicalisation has to undo
static double checkWedge(std::mt19937_64 &rng, int trials)
{
    std::uniform_real_distribution<double> U(-1.0, 1.0);
    double worst = 0;
    int made = 0;
    for (int t = 0; t < trials; ++t) {
        // vertices: 0 = o, 1 = A, 2 = B, 3 = C (beyond o-A), 4 = R (beyond A-B)
        const double base[5][3] = {{0, 0, 0}, {1, 0, 0}, {0.3, 0.9, 0}, {0.4, -0.8, 0}, {1.2, 1.0, 0}};
        double x0[15], x[15];
        const int rot = t % 3;
        const bool rev = (t / 3) % 2;  // the triangle listed as (o, B, A): mirrored rest pose, so that its rest area stays positive
        for (int s = 0; s < 5; ++s)
            for (int c = 0; c < 3; ++c) {
                x0[3 * s + c] = (c == 1 && rev ? -1.0 : 1.0) * base[s][c] + (c < 2 ? 0.15 * U(rng) : 0.0);
                x[3 * s + c] = x0[3 * s + c] + 0.2 * U(rng);
            }
        const bool flipE = (t / 6) % 2, swapE = (t / 12) % 2, flipO = (t / 24) % 2, swapO = (t / 48) % 2;
        int tri0[3] = {0, rev ? 2 : 1, rev ? 1 : 2}, tri[3];
        for (int k = 0; k < 3; ++k) tri[(k + rot) % 3] = tri0[k];
        const int localT = rot;
        int he[4] = {flipE ? 1 : 0, flipE ? 0 : 1, swapE ? 3 : 2, swapE ? 2 : 3};
        const int localE = flipE ? 1 : 0;
        int ho[4] = {flipO ? 2 : 1, flipO ? 1 : 2, swapO ? 4 : 0, swapO ? 0 : 4};
        const int localO = swapO ? 3 : 2;
        const double modelM[4] = {100.0, 0.3, 0, 0};
        const double modelE[4] = {0.3 * U(rng), 1.0 + 0.5 * U(rng), 0, 0}, modelO[4] = {0.3 * U(rng), 1.0 + 0.5 * U(rng), 0, 0};
        double pm[EnergyKind<K_MEMBRANE_STVK>::NP], pe[EnergyKind<K_DISCRETE_BENDING>::NP], po[EnergyKind<K_DISCRETE_BENDING>::NP];
        prepareStencil(K_MEMBRANE_STVK, tri, modelM, x0, pm);
        prepareStencil(K_DISCRETE_BENDING, he, modelE, x0, pe);
        prepareStencil(K_DISCRETE_BENDING, ho, modelO, x0, po);
        pe[0] = modelE[0], po[0] = modelO[0];  // a rest angle of either sign (the rest pose is flat)
        double rg[3] = {0, 0, 0}, rh[6] = {0, 0, 0, 0, 0, 0};
        auto addGeneric = [&](auto kindTag, const int *st, const double *rec, int local) {
            constexpr int KIND = decltype(kindTag)::value;
            constexpr int S = EnergyKind<KIND>::S;
            V3<D2> xs[S];
            for (int s = 0; s < S; ++s) {
                const bool own = s == local;
                xs[s] = V3<D2>(D2(x[3 * st[s]], own ? 0 : -1), D2(x[3 * st[s] + 1], own ? 1 : -1), D2(x[3 * st[s] + 2], own ? 2 : -1));
            }
            const D2 E = stencilEnergy<KIND, D2>(xs, rec);
            for (int i = 0; i < 3; ++i) rg[i] += E.g[i];
            for (int i = 0; i < 6; ++i) rh[i] += E.h[i];
        };
        addGeneric(std::integral_constant<int, K_MEMBRANE_STVK>{}, tri, pm, localT);
        addGeneric(std::integral_constant<int, K_DISCRETE_BENDING>{}, he, pe, localE);
        addGeneric(std::integral_constant<int, K_DISCRETE_BENDING>{}, ho, po, localO);
        int32_t nbr[4];
        double q[16];
        // A is the vertex after o in the triangle's cyclic order exactly when the triangle was not reversed
        if (makeWedgeRecord(tri, pm, localT, rev, he, pe, localE, ho, po, localO, nbr, q)) return 1.0;  // wrong side: must refuse
        if (!makeWedgeRecord(tri, pm, localT, !rev, he, pe, localE, ho, po, localO, nbr, q)) return 2.0;
        ++made;
        if (nbr[0] != 1 || nbr[1] != 2 || nbr[2] != 3 || nbr[3] != 4) return 3.0;
        V3d xn[4];
        for (int k = 0; k < 4; ++k) xn[k] = V3d(x[3 * nbr[k]], x[3 * nbr[k] + 1], x[3 * nbr[k] + 2]);
        V3d g(0, 0, 0);
        Sym3 H{0, 0, 0, 0, 0, 0};
        evalStream(ST_WEDGE, V3d(x[0], x[1], x[2]), xn, q, g, H);
        const double gs[3] = {g.x, g.y, g.z};
        const double hs[6] = {H.xx, H.xy, H.xz, H.yy, H.yz, H.zz};
        double scaleG = 1e-300, scaleH = 1e-300, eg = 0, eh = 0;
        for (int i = 0; i < 3; ++i) scaleG = std::fmax(scaleG, std::fabs(rg[i])), eg = std::fmax(eg, std::fabs(rg[i] - gs[i]));
        for (int i = 0; i < 6; ++i) scaleH = std::fmax(scaleH, std::fabs(rh[i])), eh = std::fmax(eh, std::fabs(rh[i] - hs[i]));
        worst = std::fmax(worst, std::fmax(eg / scaleG, eh / scaleH));
    }
    return made == trials ? worst : 4.0;
}

int main()
{
    std::mt19937_64 rng(12345);
    const double tol = 1e-10;
    int bad = 0;
    auto report = [&](const char *name, double w) {
        std::printf("%s max_rel %.3e\n", name, w);
        if (!(w < tol)) ++bad;
    };
    report("discrete_bending", check<K_DISCRETE_BENDING>(rng, 2000));
    report("membrane_stvk", check<K_MEMBRANE_STVK>(rng, 2000));
    report("tet_snh", check<K_TET_SNH>(rng, 2000));
    report("tet_stvk", check<K_TET_STVK>(rng, 2000));
    report("wedge", checkWedge(rng, 1920));
    return bad;
}
