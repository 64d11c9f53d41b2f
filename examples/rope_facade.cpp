// Minimal caller of the C++ facade, written the way a caller of the reference is (apps/elasticPendulum.cpp:1-34,
// apps/solverCompare.cpp:56-102 createRope): a rope of `n` particles joined by springs, pinned at particle 0 by an
// anchor spring, stepped with VBD or XPBD.  Prints the final positions, one particle per line, with 17 digits.
//   rope_facade <vbd|xpbd> <nParticles> <frames> <substeps> <iterations> [pick]
// With `pick`, the last particle is picked for the middle third of the run the way the reference application does it
// (apps/solverCompare.cpp:717-726, :1446-1455, :1471-1473, :1499-1502): a stiff anchor spring joins the SOLVER's copy of
// the group, the solver is flagged with requirePrecomputationUpdate, and the energy is removed again on release.
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
    const bool pick = argc > 6 && std::strcmp(argv[6], "pick") == 0;
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
        solver::VBD *asVbd = nullptr;
        solver::XPBD *asXpbd = nullptr;
        if (vbd) {
            auto v = std::make_unique<solver::VBD>();
            v->setIterations(iters);
            asVbd = v.get();
            s = std::move(v);
        } else {
            auto x = std::make_unique<solver::XPBD>();
            asXpbd = x.get();
            s = std::move(x);
        }
        auto flagTopologyChange = [&] {
            if (asVbd) asVbd->requirePrecomputationUpdate();
            if (asXpbd) asXpbd->requirePrecomputationUpdate();
        };
        s->addObject(rope);
        s->setDt(0.01);
        s->setNumSubsteps(substeps);
        std::shared_ptr<SpringAnchorEnergy> picking;
        for (int f = 0; f < frames; ++f) {
            if (!pick && f == frames / 2) anchor->updateAnchor(0, {0, 0.05, 0});  // live edit, solverCompare.cpp:1524
            if (pick && f == frames / 3) {
                ParticleGroup &live = s->particleGroups()[0];
                picking = std::make_shared<SpringAnchorEnergy>();
                picking->addStencil({n - 1}, live.positions()[n - 1], 1e4);
                live.addEnergy(picking);
                flagTopologyChange();
            }
            if (pick && picking && f > frames / 3) picking->updateAnchor(0, {1.0, 0.01 * f, 0.02});  // the drag
            if (pick && f == 2 * frames / 3) {
                s->particleGroups()[0].removeEnergy(picking);
                picking = nullptr;
                flagTopologyChange();
            }
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
