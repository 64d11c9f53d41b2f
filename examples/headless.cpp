// Headless counterpart of the reference's solverCompare application: loads a triangle mesh
// (.obj) or a tetrahedral mesh (.msh) through spgx::io, dresses it the way createCloth / createTetrahedralBeam do
// (apps/solverCompare.cpp:367-498, :260-330), steps it with the chosen solver on the GPU and dumps frames instead of drawing
// them (there is no window: polyscope / OpenGL are unavailable on this box).
//
//   headless <mesh.obj|mesh.msh> <solver> <frames> <substeps> <iterations> <anchor ids, comma separated> [framePrefix]
//
//   .obj : the file is the REST pose (planar, z = 0: the reference's membrane energies insist on it); the initial pose is the
//          rest pose laid flat, (x, y, 0) -> (x, 0, -y), as buildTriangleSquareMesh does (solverCompare.cpp:332-365);
//          DiscreteBendingEnergy (rest angle 0, k 1e-2) on every interior edge in TriangleMesh edge order (triangleMesh.h:36-50:
//          unique edges in (min, max) order, opposing vertices in face order), MembraneStvkEnergy (k 1e2, nu 0) on every face,
//          area-weighted masses summing to 1; energies in createCloth's insertion order: bending, membrane, anchors (k 1e5)
//   .msh : anchors (k 1e4), StableNeoHookeanEnergy (E 1e3, nu 0.2) on every tetrahedron; volume-weighted masses summing to 1
//   solver: vbd xpbd baraff_witkin newton_dx newton_dv bdf2 newton_robust symplectic
// Prints the final positions (17 digits, one particle per line); with framePrefix also writes <prefix>NNNN.obj per frame.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "spgx/geom.hpp"

using namespace spgx;

static Vector3 sub(const Vector3 &a, const Vector3 &b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static Vector3 cross(const Vector3 &a, const Vector3 &b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
static Real dot(const Vector3 &a, const Vector3 &b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

static std::unique_ptr<solver::BaseSolver> makeSolver(const std::string &name, int iterations)
{
    if (name == "vbd") {
        auto s = std::make_unique<solver::VBD>();
        s->setIterations(iterations);
        return s;
    }
    if (name == "xpbd") return std::make_unique<solver::XPBD>();
    if (name == "baraff_witkin") return std::make_unique<solver::ImplicitEulerBaraffWitkin>();
    if (name == "newton_dx") return std::make_unique<solver::ImplicitEulerNewtonDx>();
    if (name == "newton_dv") return std::make_unique<solver::ImplicitEulerNewtonDv>();
    if (name == "bdf2") return std::make_unique<solver::BDF2>();
    if (name == "newton_robust") return std::make_unique<solver::ImplicitEulerNewtonRobust>();
    if (name == "symplectic") return std::make_unique<solver::SimplecticEuler>();
    throw std::runtime_error("unknown solver " + name);
}

int main(int argc, char **argv)
{
    if (argc < 7) {
        std::fprintf(stderr, "usage: %s <mesh.obj|mesh.msh> <solver> frames substeps iterations anchorIds [framePrefix]\n", argv[0]);
        return 2;
    }
    const std::string path = argv[1], solverName = argv[2];
    const int frames = std::atoi(argv[3]), substeps = std::atoi(argv[4]), iterations = std::atoi(argv[5]);
    std::vector<int> anchors;
    for (char *tok = std::strtok(argv[6], ","); tok; tok = std::strtok(nullptr, ",")) anchors.push_back(std::atoi(tok));
    const std::string prefix = argc > 7 ? argv[7] : "";
    try {
        ParticleGroup object;
        std::vector<Int3> faces;
        const bool isObj = path.size() > 4 && path.substr(path.size() - 4) == ".obj";
        if (isObj) {
            std::vector<Vector3> verts;
            std::tie(verts, faces) = io::loadObj(path);
            std::vector<Real> area(faces.size()), particleArea(verts.size(), 0.0);
            Real total = 0;
            for (size_t f = 0; f < faces.size(); ++f) {
                const Vector3 n = cross(sub(verts[faces[f][1]], verts[faces[f][0]]), sub(verts[faces[f][2]], verts[faces[f][1]]));
                area[f] = 0.5 * std::sqrt(dot(n, n));
                total += area[f];
            }
            for (size_t f = 0; f < faces.size(); ++f)
                for (int c : faces[f]) particleArea[c] += area[f] / 3.0;
            std::vector<Vector3> flat(verts.size());
            for (size_t i = 0; i < verts.size(); ++i) flat[i] = {verts[i].x, 0.0, -verts[i].y};
            for (size_t i = 0; i < verts.size(); ++i) object.addParticle(flat[i], verts[i], particleArea[i] * (1.0 / total));
            const TriangleMesh mesh(flat, verts, faces);  // edges in (min, max) order, opposing vertices in face order
            auto bending = std::make_shared<DiscreteBendingEnergy>();
            for (size_t e = 0; e < mesh.edges().size(); ++e) {
                const auto &opp = mesh.opposingVerticesPerEdge()[e];
                if (opp.size() == 2) bending->addStencil({mesh.edges()[e][0], mesh.edges()[e][1], opp[0].vertexId, opp[1].vertexId}, 0.0, 1e-2);
            }
            auto membrane = std::make_shared<MembraneStvkEnergy>();
            for (const auto &f : faces) membrane->addStencil({f[0], f[1], f[2]}, 1e2, 0.0);
            auto pins = std::make_shared<SpringAnchorEnergy>();
            for (int a : anchors) pins->addStencil({a}, flat.at(a), 1e5);
            object.addEnergy(bending);
            bending->preparePrecomputations(object);
            object.addEnergy(membrane);
            membrane->preparePrecomputations(object);
            object.addEnergy(pins);
        } else {
            std::vector<Vector3> verts;
            std::vector<Int4> tets;
            std::tie(verts, tets) = io::loadMsh(path);
            std::vector<Real> vol(tets.size()), particleVol(verts.size(), 0.0);
            Real total = 0;
            for (size_t t = 0; t < tets.size(); ++t) {
                const Vector3 &p0 = verts[tets[t][0]];
                vol[t] = std::fabs(dot(sub(verts[tets[t][1]], p0), cross(sub(verts[tets[t][2]], p0), sub(verts[tets[t][3]], p0)))) / 6.0;
                total += vol[t];
            }
            for (size_t t = 0; t < tets.size(); ++t)
                for (int c : tets[t]) particleVol[c] += vol[t] / 4.0;
            for (size_t i = 0; i < verts.size(); ++i) object.addParticle(verts[i], verts[i], particleVol[i] * (1.0 / total));
            auto pins = std::make_shared<SpringAnchorEnergy>();
            for (int a : anchors) pins->addStencil({a}, verts.at(a), 1e4);
            auto elastic = std::make_shared<StableNeoHookeanEnergy>();
            const Real young = static_cast<Real>(1e3f), poisson = static_cast<Real>(0.2f);  // createTetrahedralBeam takes floats
            for (const auto &t : tets) elastic->addStencil({t[0], t[1], t[2], t[3]}, young, poisson);
            object.addEnergy(pins);
            object.addEnergy(elastic);
            elastic->preparePrecomputations(object);
        }

        auto s = makeSolver(solverName, iterations);
        s->addObject(object);
        s->setDt(0.01);
        s->setNumSubsteps(substeps);
        for (int f = 0; f < frames; ++f) {
            s->step();
            if (!prefix.empty()) {
                char name[32];
                std::snprintf(name, sizeof name, "%04d.obj", f);
                io::saveObj(prefix + name, static_cast<const solver::BaseSolver &>(*s).particleGroups()[0].positions(), faces);
            }
        }
        const auto &x = static_cast<const solver::BaseSolver &>(*s).particleGroups()[0].positions();
        for (const auto &p : x) std::printf("%.17g %.17g %.17g\n", p.x, p.y, p.z);
    } catch (const std::exception &e) {
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    return 0;
}
