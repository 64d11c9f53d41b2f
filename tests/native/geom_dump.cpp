// Prints what spgx::TriangleMesh / spgx::TetrahedralMesh (include/spgx/geom.hpp) derive from a mesh file
// (tests/test_meshio.py): for an .obj "E n" + n lines "lo hi k (triangle opposite)*k"; for a .msh "E n" + n lines "lo hi".
// `selftest` checks the two validation errors of each class and prints their messages.
#include <cstdio>
#include <cstring>
#include <string>

#include "spgx/geom.hpp"

using namespace spgx;

int main(int argc, char **argv)
{
    if (argc < 2) return 2;
    try {
        if (std::strcmp(argv[1], "selftest") == 0) {
            const std::vector<Vector3> v{{0, 0, 0}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
            const char *expected[4] = {"Triangle mesh face contains index to invalid vertex", "Triangle mesh face contains repeated vertex index",
                                       "Tetrahedral mesh cell contains index to invalid vertex",
                                       "Tetrahedral mesh cell contains repeated vertex index"};
            for (int k = 0; k < 4; ++k) {
                try {
                    if (k == 0) TriangleMesh(v, v, {{0, 1, 4}});
                    if (k == 1) TriangleMesh(v, v, {{0, 1, 2}, {2, 2, 3}});
                    if (k == 2) TetrahedralMesh(v, v, {{0, 1, 2, -1}});
                    if (k == 3) TetrahedralMesh(v, v, {{0, 1, 2, 1}});
                    std::printf("case %d did not throw\n", k);
                    return 1;
                } catch (const std::runtime_error &e) {
                    if (std::strcmp(e.what(), expected[k]) != 0) {
                        std::printf("case %d: %s\n", k, e.what());
                        return 1;
                    }
                }
            }
            std::printf("ok\n");
            return 0;
        }
        const std::string path = argv[1];
        if (path.size() > 4 && path.substr(path.size() - 4) == ".obj") {
            auto [v, f] = io::loadObj(path);
            const TriangleMesh mesh(v, v, f);
            std::printf("E %zu\n", mesh.edges().size());
            for (size_t e = 0; e < mesh.edges().size(); ++e) {
                const auto &opp = mesh.opposingVerticesPerEdge()[e];
                std::printf("%d %d %zu", mesh.edges()[e][0], mesh.edges()[e][1], opp.size());
                for (const auto &o : opp) std::printf(" %d %d", o.triangleId, o.vertexId);
                std::printf("\n");
            }
        } else {
            auto [v, t] = io::loadMsh(path);
            const TetrahedralMesh mesh(v, v, t);
            std::printf("E %zu\n", mesh.edges().size());
            for (const auto &e : mesh.edges()) std::printf("%d %d\n", e[0], e[1]);
        }
    } catch (const std::exception &e) {
        std::fprintf(stderr, "%s\n", e.what());
        return 1;
    }
    return 0;
}
