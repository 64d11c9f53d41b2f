// Mesh ingestion for the facade: spg::io::loadObj / spg::io::loadMsh
// (spg/include/spg/geom/io/obj.h:7, msh.h; spg/source/geom/io/obj.cpp:10-35, msh.cpp:10-41).
//
// The reference delegates the parsing to tinyobjloader and MshIO behind compile-time switches (TINYOBJLOADER_ACTIVE /
// MSHIO_ACTIVE, off by default, in which case both functions throw); neither library is in this image, so the two formats
// are read here directly.  What the reference keeps of a file is kept here:
//   loadObj: every `v x y z` of the file, and the faces of the FIRST shape only (obj.cpp:26-31 reads shapes.front()),
//            fan-triangulated as tinyobjloader does by default; only the vertex index of a `v/vt/vn` corner is used;
//            negative (relative) indices are resolved; a new `o` / `g` statement after the first face starts the
//            second shape, which is ignored.
//   loadMsh: all node blocks in file order, and the first four nodes of every element turned 0-based (msh.cpp:30-35 reads
//            block.data[j * 5 + 1 .. 4] - 1, i.e. it assumes 4-node tetrahedra): elements of any other type are skipped
//            here instead of being misread.  ASCII MSH 4.1 and 2.2.
#pragma once
#include <array>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#include "spgx.hpp"

namespace spgx
{
using Int3 = std::array<int, 3>;
using Int4 = std::array<int, 4>;
namespace io
{
inline std::tuple<std::vector<Vector3>, std::vector<Int3>> loadObj(const std::string &filePath)
{
    std::ifstream in(filePath);
    if (!in) throw std::runtime_error("loadObj: cannot open " + filePath);
    std::vector<Vector3> vertices;
    std::vector<Int3> faces;
    bool firstShapeHasFaces = false, firstShapeClosed = false;
    std::string line;
    while (std::getline(in, line)) {
        std::istringstream ls(line);
        std::string tag;
        if (!(ls >> tag)) continue;
        if (tag == "v") {
            Real x = 0, y = 0, z = 0;
            ls >> x >> y >> z;
            vertices.push_back({x, y, z});
        } else if (tag == "o" || tag == "g") {
            if (firstShapeHasFaces) firstShapeClosed = true;
        } else if (tag == "f" && !firstShapeClosed) {
            std::vector<int> corner;
            std::string tok;
            while (ls >> tok) {
                const int idx = std::stoi(tok.substr(0, tok.find('/')));
                corner.push_back(idx > 0 ? idx - 1 : static_cast<int>(vertices.size()) + idx);
            }
            if (corner.size() < 3) throw std::runtime_error("loadObj: face with fewer than three corners in " + filePath);
            for (size_t k = 1; k + 1 < corner.size(); ++k) faces.push_back({corner[0], corner[k], corner[k + 1]});
            firstShapeHasFaces = true;
        }
    }
    for (const auto &f : faces)
        for (int c : f)
            if (c < 0 || c >= static_cast<int>(vertices.size())) throw std::runtime_error("loadObj: vertex index out of range in " + filePath);
    return {std::move(vertices), std::move(faces)};
}

inline std::tuple<std::vector<Vector3>, std::vector<Int4>> loadMsh(const std::string &filePath)
{
    std::ifstream in(filePath);
    if (!in) throw std::runtime_error("loadMsh: cannot open " + filePath);
    auto expect = [&](const std::string &what) {
        std::string tok;
        while (in >> tok)
            if (tok == what) return;
        throw std::runtime_error("loadMsh: section " + what + " not found in " + filePath);
    };
    expect("$MeshFormat");
    double version = 0;
    int fileType = 0, dataSize = 0;
    in >> version >> fileType >> dataSize;
    if (fileType != 0) throw std::runtime_error("loadMsh: binary MSH files are not supported: " + filePath);
    std::vector<Vector3> vertices;
    std::vector<Int4> tets;
    std::vector<long long> tags;  // node tag of every vertex, in file order
    expect("$Nodes");
    if (version >= 4.0) {
        long long nBlocks = 0, nNodes = 0, minTag = 0, maxTag = 0;
        in >> nBlocks >> nNodes >> minTag >> maxTag;
        for (long long b = 0; b < nBlocks; ++b) {
            long long dim = 0, entity = 0, parametric = 0, n = 0;
            in >> dim >> entity >> parametric >> n;
            for (long long j = 0; j < n; ++j) {
                long long t = 0;
                in >> t;
                tags.push_back(t);
            }
            for (long long j = 0; j < n; ++j) {
                Real x = 0, y = 0, z = 0;
                in >> x >> y >> z;
                for (long long p = 0; p < (parametric ? dim : 0); ++p) {
                    Real skip;
                    in >> skip;
                }
                vertices.push_back({x, y, z});
            }
        }
    } else {
        long long nNodes = 0;
        in >> nNodes;
        for (long long j = 0; j < nNodes; ++j) {
            long long t = 0;
            Real x = 0, y = 0, z = 0;
            in >> t >> x >> y >> z;
            tags.push_back(t);
            vertices.push_back({x, y, z});
        }
    }
    if (!in) throw std::runtime_error("loadMsh: malformed $Nodes section in " + filePath);
    expect("$Elements");
    auto nodesOf = [](long long type) {  // gmsh element types
        static const int n[] = {0, 2, 3, 4, 4, 8, 6, 5, 3, 6, 9, 10, 27, 18, 14, 1};
        return type >= 1 && type <= 15 ? n[type] : -1;
    };
    if (version >= 4.0) {
        long long nBlocks = 0, nElems = 0, minTag = 0, maxTag = 0;
        in >> nBlocks >> nElems >> minTag >> maxTag;
        for (long long b = 0; b < nBlocks; ++b) {
            long long dim = 0, entity = 0, type = 0, n = 0;
            in >> dim >> entity >> type >> n;
            const int nn = nodesOf(type);
            if (nn < 0) throw std::runtime_error("loadMsh: unsupported element type in " + filePath);
            for (long long j = 0; j < n; ++j) {
                long long tag = 0, node[27] = {0};
                in >> tag;
                for (int k = 0; k < nn; ++k) in >> node[k];
                if (type == 4)  // the reference keeps nodes 1..4 of the element, 0-based (msh.cpp:31-34)
                    tets.push_back({static_cast<int>(node[0]) - 1, static_cast<int>(node[1]) - 1, static_cast<int>(node[2]) - 1,
                                    static_cast<int>(node[3]) - 1});
            }
        }
    } else {
        long long nElems = 0;
        in >> nElems;
        for (long long j = 0; j < nElems; ++j) {
            long long tag = 0, type = 0, nTags = 0, skip = 0, node[27] = {0};
            in >> tag >> type >> nTags;
            for (long long k = 0; k < nTags; ++k) in >> skip;
            const int nn = nodesOf(type);
            if (nn < 0) throw std::runtime_error("loadMsh: unsupported element type in " + filePath);
            for (int k = 0; k < nn; ++k) in >> node[k];
            if (type == 4)
                tets.push_back({static_cast<int>(node[0]) - 1, static_cast<int>(node[1]) - 1, static_cast<int>(node[2]) - 1,
                                static_cast<int>(node[3]) - 1});
        }
    }
    if (!in) throw std::runtime_error("loadMsh: malformed $Elements section in " + filePath);
    // the reference indexes vertices by (node tag - 1): that is only the file order when the tags are 1..n in order
    for (size_t i = 0; i < tags.size(); ++i)
        if (tags[i] != static_cast<long long>(i) + 1) throw std::runtime_error("loadMsh: node tags are not 1..n in file order in " + filePath);
    for (const auto &t : tets)
        for (int c : t)
            if (c < 0 || c >= static_cast<int>(vertices.size())) throw std::runtime_error("loadMsh: node index out of range in " + filePath);
    return {std::move(vertices), std::move(tets)};
}

// frame dump of the headless runs: positions + the triangle faces of the scene
inline void saveObj(const std::string &filePath, const std::vector<Vector3> &vertices, const std::vector<Int3> &faces)
{
    std::ofstream out(filePath);
    if (!out) throw std::runtime_error("saveObj: cannot open " + filePath);
    out.precision(17);
    for (const auto &v : vertices) out << "v " << v.x << ' ' << v.y << ' ' << v.z << '\n';
    for (const auto &f : faces) out << "f " << f[0] + 1 << ' ' << f[1] + 1 << ' ' << f[2] + 1 << '\n';
}
}  // namespace io
}  // namespace spgx
