// Host-side check of the hand-derived 9 x 9 StVK membrane Hessian (spg_b200/csrc/membrane_hessian.cuh) against the
// second-order dual-number evaluation of the same energy over all nine coordinates (energies.cuh: the restatement of what
// TinyAD computes, energy.h:196-237) on random triangles.  Compiled with g++.  Prints "max_rel <value>".
#include <cmath>
#include <cstdio>
#include <random>

#include "../../spg_b200/csrc/energy_prepare.h"
#include "../../spg_b200/csrc/membrane_hessian.cuh"

using namespace spgx;
using D = Dual<9, true>;

int main()
{
    std::mt19937_64 rng(99);
    std::uniform_real_distribution<double> U(-1.0, 1.0);
    double worst = 0;
    for (int t = 0; t < 2000; ++t) {
        const double base[3][3] = {{0, 0, 0}, {1, 0, 0}, {0.3, 0.9, 0}};
        double x0[9], xs[9];
        for (int s = 0; s < 3; ++s)
            for (int c = 0; c < 3; ++c) {
                x0[3 * s + c] = c == 2 ? 0.0 : base[s][c] + 0.15 * U(rng);
                xs[3 * s + c] = x0[3 * s + c] + 0.3 * U(rng);
            }
        const int st[3] = {0, 1, 2};
        const double model[2] = {100.0 * (1.0 + 0.5 * U(rng)), 0.3 * (1.0 + U(rng))};
        double rec[EnergyKind<K_MEMBRANE_STVK>::NP];
        prepareStencil(K_MEMBRANE_STVK, st, model, x0, rec);
        V3<D> xd[3];
        V3d xv[3];
        for (int s = 0; s < 3; ++s) {
            xd[s] = V3<D>(D(xs[3 * s], 3 * s), D(xs[3 * s + 1], 3 * s + 1), D(xs[3 * s + 2], 3 * s + 2));
            xv[s] = V3d(xs[3 * s], xs[3 * s + 1], xs[3 * s + 2]);
        }
        const D E = stencilEnergy<K_MEMBRANE_STVK, D>(xd, rec);
        double out[81];
        membraneStvkElementHessian(xv, rec, out);
        double scale = 0, err = 0;
        for (int i = 0; i < 9; ++i)
            for (int j = 0; j < 9; ++j) {
                const double ref = E.hess(i, j);
                const double got = out[((i / 3) * 3 + (j / 3)) * 9 + 3 * (i % 3) + (j % 3)];
                scale = std::fmax(scale, std::fabs(ref));
                err = std::fmax(err, std::fabs(ref - got));
            }
        worst = std::fmax(worst, err / scale);
    }
    std::printf("max_rel %.3e\n", worst);
    return worst <= 1e-11 ? 0 : 1;
}
