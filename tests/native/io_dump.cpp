// Prints what spgx::io::loadObj / loadMsh keep of a file (tests/test_meshio.py): "V n" + n vertex lines, "E m k" + m element
// lines of k indices.  Exit code 1 with the message on stderr when the loader throws.
#include <cstdio>
#include <string>

#include "spgx/io.hpp"

int main(int argc, char **argv)
{
    if (argc < 2) return 2;
    const std::string path = argv[1];
    try {
        if (path.size() > 4 && path.substr(path.size() - 4) == ".obj") {
            auto [v, f] = spgx::io::loadObj(path);
            std::printf("V %zu\n", v.size());
            for (auto &p : v) std::printf("%.17g %.17g %.17g\n", p.x, p.y, p.z);
            std::printf("E %zu 3\n", f.size());
            for (auto &t : f) std::printf("%d %d %d\n", t[0], t[1], t[2]);
        } else {
            auto [v, t] = spgx::io::loadMsh(path);
            std::printf("V %zu\n", v.size());
            for (auto &p : v) std::printf("%.17g %.17g %.17g\n", p.x, p.y, p.z);
            std::printf("E %zu 4\n", t.size());
            for (auto &e : t) std::printf("%d %d %d %d\n", e[0], e[1], e[2], e[3]);
        }
    } catch (const std::exception &e) {
        std::fprintf(stderr, "%s\n", e.what());
        return 1;
    }
    return 0;
}
