// Minimal caller of the C++ facade, written the way a caller of the reference is (apps/elasticPendulum.cpp:1-34,
// apps/solverCompare.cpp:56-102 createRope): a rope of `n` particles joined by springs, pinned at particle 0 by an
// anchor spring, stepped with VBD or XPBD.  Prints the final positions, one particle per line, with 17 digits.
//   rope_facade <vbd|xpbd> <nParticles> <frames> <substeps> <iterations>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "spgx/spgx.hpp"

int main(int argc, char **argv)
{
    if (argc < 6) {
        std::fprintf(stderr, "usage: %s <vbd|xpbd> n frames substeps iterations\n", argv[0]);
        return 2;
    }
    const bool vbd = std::strcmp(argv[1], "vbd") == 0;
    const int n = std::atoi(argv[2]), frames = std::atoi(argv[3]), substeps = std::atoi(argv[4]), iters = std::atoi(argv[5]);
    try {
        using namespace spgx;
        ParticleGroup rope;
        const Real length = 1.0, mass = 1.0, dl = length / (n - 1);
        for (int i = 0; i < n; ++i) rope.addParticle({i * dl, 0, 0}, {i * dl, 0, 0}, mass / n);
        auto springs = std::make_shared<SpringEnergy>();
        for (int i = 0; i + 1 < n; ++i) springs->addStencil({i, i + 1}, dl, 1e3);
        auto anchor = std::make_shared<SpringAnchorEnergy>();
        anchor->addStencil({0}, rope.positions()[0], 1e5);
        springs->preparePrecomputations(rope);
        rope.addEnergy(springs);
        rope.addEnergy(anchor);

        std::unique_ptr<solver::BaseSolver> s;
        if (vbd) {
            auto v = std::make_unique<solver::VBD>();
            v->setIterations(iters);
            s = std::move(v);
        } else {
            s = std::make_unique<solver::XPBD>();
        }
        s->addObject(rope);
        s->setDt(0.01);
        s->setNumSubsteps(substeps);
        for (int f = 0; f < frames; ++f) {
            if (f == frames / 2) anchor->updateAnchor(0, {0, 0.05, 0});  // live edit, solverCompare.cpp:1524
            s->step();
        }
        const auto &x = static_cast<const solver::BaseSolver &>(*s).particleGroups()[0].positions();
        for (const auto &p : x) std::printf("%.17g %.17g %.17g\n", p.x, p.y, p.z);
    } catch (const std::runtime_error &e) {
        std::fprintf(stderr, "runtime_error: %s\n", e.what());
        return 1;
    }
    return 0;
}
