// Host-side check of the canonical stream evaluators (spg_b200/csrc/streams.cuh) against the generic second-order
// dual-number evaluation of the same energy (energies.cuh, the restatement of energy.h:240-258) on random stencils,
// for every streamed kind and every owner position.  Compiled with g++ (the evaluators are __host__ __device__).
// Prints "max_rel <value>" per (kind, local) and exits non-zero above the tolerance.
#include <cmath>
#include <cstdio>
#include <random>

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
    return bad;
}
