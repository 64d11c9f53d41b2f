// Host-side check of the hand-derived 12 x 12 discrete-bending Hessian (spg_b200/csrc/hinge_hessian.cuh) against the
// second-order dual-number evaluation of the same energy over all twelve coordinates (energies.cuh: the restatement of
// what TinyAD computes, energy.h:196-237) on random hinges, flat and folded.  Compiled with g++ (the evaluators are
// __host__ __device__).  Prints "max_rel <value>" and exits non-zero above the tolerance.
#include <cmath>
#include <cstdio>
#include <random>

#include "../../spg_b200/csrc/hinge_hessian.cuh"

using namespace spgx;
using D = Dual<12, true>;

int main()
{
    std::mt19937_64 rng(1234);
    std::uniform_real_distribution<double> U(-1.0, 1.0);
    double worst = 0;
    for (int t = 0; t < 2000; ++t) {
        const double base[4][3] = {{0, 0, 0}, {1, 0, 0}, {0.3, 0.9, 0}, {0.4, -0.8, 0}};
        const double fold = t % 4 == 0 ? 0.0 : (t % 4 == 1 ? 1e-3 : 0.6);  // flat, nearly flat, folded
        double xs[12];
        for (int s = 0; s < 4; ++s)
            for (int c = 0; c < 3; ++c) xs[3 * s + c] = base[s][c] + (c == 2 ? fold : 0.2) * U(rng);
        const double p[3] = {0.4 * U(rng), 1.0 + 0.5 * U(rng), 0.0};
        V3<D> xd[4];
        V3d xv[4];
        for (int s = 0; s < 4; ++s) {
            xd[s] = V3<D>(D(xs[3 * s], 3 * s), D(xs[3 * s + 1], 3 * s + 1), D(xs[3 * s + 2], 3 * s + 2));
            xv[s] = V3d(xs[3 * s], xs[3 * s + 1], xs[3 * s + 2]);
        }
        const D E = stencilEnergy<K_DISCRETE_BENDING, D>(xd, p);
        double out[16 * 9];
        hingeElementHessian(xv, p, out);
        double scale = 0, err = 0;
        for (int i = 0; i < 12; ++i)
            for (int j = 0; j < 12; ++j) {
                const double ref = E.hess(i, j);
                const double got = out[((i / 3) * 4 + (j / 3)) * 9 + 3 * (i % 3) + (j % 3)];
                scale = std::fmax(scale, std::fabs(ref));
                err = std::fmax(err, std::fabs(ref - got));
            }
        worst = std::fmax(worst, err / scale);
    }
    std::printf("max_rel %.3e\n", worst);
    return worst <= 1e-11 ? 0 : 1;
}
