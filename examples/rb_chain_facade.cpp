// Rigid bodies through the C++ facade, written the way the reference application builds its chain scene
// (apps/solverCompare.cpp:605-660 createRigidBodyChain): boxes hung corner to corner by RBSpringEnergy, the first one from
// an RBSpringAnchorEnergy, stepped with VBD or XPBD; half way the anchor is dragged (RBSpringAnchorEnergy::updateAnchor).
// Prints, per body, the centre of mass and the axis-angle vector with 17 digits.
//   rb_chain_facade <vbd|xpbd> <nBodies> <frames> <substeps> <iterations>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "spgx/spgx.hpp"

int main(int argc, char **argv)
{
    if (argc < 6) {
        std::fprintf(stderr, "usage: %s <vbd|xpbd> nBodies frames substeps iterations\n", argv[0]);
        return 2;
    }
    const bool vbd = std::strcmp(argv[1], "vbd") == 0;
    const int n = std::atoi(argv[2]), frames = std::atoi(argv[3]), substeps = std::atoi(argv[4]), iters = std::atoi(argv[5]);
    try {
        using namespace spgx;
        const Real mass = 1;
        const float width = 1, height = 1, depth = 1;
        Matrix3 localInertia{};
        localInertia[0] = 1.0 / 12.0 * mass * (height * height + depth * depth);
        localInertia[4] = 1.0 / 12.0 * mass * (width * width + depth * depth);
        localInertia[8] = 1.0 / 12.0 * mass * (width * width + height * height);
        RigidBodyGroup rbGroup;
        rbGroup.addBody({0, 0, 0}, {0, 0, 0}, mass, {0, 0, 0}, {0, 0, 0}, localInertia);
        auto springAnchorEnergy = std::make_shared<RBSpringAnchorEnergy>();
        springAnchorEnergy->addStencil({0}, {width * 0.5, height * 0.5, 0}, {width * 0.5, height * 0.5, 0}, 100);
        auto springEnergy = std::make_shared<RBSpringEnergy>();
        for (int i = 1; i < n; ++i) {
            rbGroup.addBody({-(height + 1) * i, 0, -(height + 1) * i}, {-(height + 1) * i, 0, -(height + 1) * i}, mass, {0, 0, 0}, {0, 0, 0},
                            localInertia);
            springEnergy->addStencil({i - 1, i}, {Vector3{-width * 0.5, -height * 0.5, 0}, Vector3{width * 0.5, height * 0.5, 0}},
                                     std::sqrt(width * 2), 100);
        }
        rbGroup.addEnergy(springAnchorEnergy);
        if (n > 1) rbGroup.addEnergy(springEnergy);

        std::unique_ptr<solver::BaseSolver> s;
        if (vbd) {
            auto v = std::make_unique<solver::VBD>();
            v->setIterations(iters);
            s = std::move(v);
        } else {
            s = std::make_unique<solver::XPBD>();
        }
        s->addObject(rbGroup);
        s->setDt(0.01);
        s->setNumSubsteps(substeps);
        for (int f = 0; f < frames; ++f) {
            if (f == frames / 2) springAnchorEnergy->updateAnchor(0, {0.7, 0.6, 0.1});
            s->step();
        }
        const RigidBodyGroup &out = static_cast<const solver::BaseSolver &>(*s).rbGroups()[0];
        for (int i = 0; i < out.size(); ++i)
            std::printf("%.17g %.17g %.17g %.17g %.17g %.17g\n", out.positions()[i].x, out.positions()[i].y, out.positions()[i].z,
                        out.thetas()[i].x, out.thetas()[i].y, out.thetas()[i].z);
    } catch (const std::exception &e) {
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    return 0;
}
