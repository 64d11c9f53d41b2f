// Mesh helpers of the facade: spg::TriangleMesh (spg/include/spg/geom/triangleMesh.h:9-65) and spg::TetrahedralMesh
// (spg/include/spg/geom/tetrahedralMesh.h:9-58) with the same public surface, the same validation messages and the same
// ORDER of what they derive, because scene builders (apps/solverCompare.cpp:367-498 createCloth, :260-330
// createTetrahedralBeam) turn that order into stencil order, and stencil order decides colours and summation order:
//   edges()                     unique undirected edges, ascending by (min vertex, max vertex)
//   opposingVerticesPerEdge()   per edge, the (triangle, opposite vertex) pairs in face order, corner order within a face
// The reference builds both through a std::map / std::set keyed by the edge; here the half-edges are collected in one
// array and stable-sorted by key, which yields the same order in O(F log F) without a node allocation per edge.
#pragma once
#include <algorithm>
#include <array>
#include <stdexcept>
#include <vector>

#include "io.hpp"

namespace spgx
{
using Int2 = std::array<int, 2>;

class TriangleMesh
{
public:
    struct OpposingVertex {
        int triangleId;
        int vertexId;
    };
    TriangleMesh(const std::vector<Vector3> &vertices, const std::vector<Vector3> &vertices0, const std::vector<Int3> &faces)
        : m_vertices(vertices), m_vertices0(vertices0), m_faces(faces)
    {
        const int nVertices = static_cast<int>(m_vertices.size());
        for (const auto &f : m_faces)
            for (int c : f)
                if (c < 0 || c >= nVertices) throw std::runtime_error("Triangle mesh face contains index to invalid vertex");
        for (const auto &f : m_faces)
            if (f[0] == f[1] || f[1] == f[2] || f[2] == f[0]) throw std::runtime_error("Triangle mesh face contains repeated vertex index");
        struct HalfEdge {
            int lo, hi, face, opposite;
        };
        std::vector<HalfEdge> half;
        half.reserve(3 * m_faces.size());
        for (int f = 0; f < static_cast<int>(m_faces.size()); ++f)
            for (int i = 0; i < 3; ++i) {
                const int a = m_faces[f][i], b = m_faces[f][(i + 1) % 3];
                half.push_back({std::min(a, b), std::max(a, b), f, m_faces[f][(i + 2) % 3]});
            }
        std::stable_sort(half.begin(), half.end(), [](const HalfEdge &p, const HalfEdge &q) { return p.lo != q.lo ? p.lo < q.lo : p.hi < q.hi; });
        for (size_t k = 0; k < half.size(); ++k) {
            if (k == 0 || half[k].lo != half[k - 1].lo || half[k].hi != half[k - 1].hi) {
                m_edges.push_back({half[k].lo, half[k].hi});
                m_opposingVerticesPerEdge.emplace_back();
            }
            m_opposingVerticesPerEdge.back().push_back({half[k].face, half[k].opposite});
        }
    }
    const std::vector<Vector3> &vertices() const { return m_vertices; }
    const std::vector<Vector3> &vertices0() const { return m_vertices0; }
    const std::vector<Int3> &faces() const { return m_faces; }
    const std::vector<Int2> &edges() const { return m_edges; }
    const std::vector<std::vector<OpposingVertex>> &opposingVerticesPerEdge() const { return m_opposingVerticesPerEdge; }

protected:
    std::vector<Vector3> m_vertices, m_vertices0;
    std::vector<Int3> m_faces;
    std::vector<Int2> m_edges;
    std::vector<std::vector<OpposingVertex>> m_opposingVerticesPerEdge;
};

class TetrahedralMesh
{
public:
    TetrahedralMesh(const std::vector<Vector3> &vertices, const std::vector<Vector3> &vertices0, const std::vector<Int4> &tets)
        : m_vertices(vertices), m_vertices0(vertices0), m_tets(tets)
    {
        const int nVertices = static_cast<int>(m_vertices.size());
        for (const auto &t : m_tets)
            for (int c : t)
                if (c < 0 || c >= nVertices) throw std::runtime_error("Tetrahedral mesh cell contains index to invalid vertex");
        for (const auto &t : m_tets)
            for (int i = 0; i < 3; ++i)
                for (int j = i + 1; j < 4; ++j)
                    if (t[i] == t[j]) throw std::runtime_error("Tetrahedral mesh cell contains repeated vertex index");
        m_edges.reserve(6 * m_tets.size());
        for (const auto &t : m_tets)
            for (int i = 0; i < 3; ++i)
                for (int j = i + 1; j < 4; ++j) m_edges.push_back({std::min(t[i], t[j]), std::max(t[i], t[j])});
        std::sort(m_edges.begin(), m_edges.end());
        m_edges.erase(std::unique(m_edges.begin(), m_edges.end()), m_edges.end());
    }
    const std::vector<Vector3> &vertices() const { return m_vertices; }
    const std::vector<Vector3> &vertices0() const { return m_vertices0; }
    const std::vector<Int4> &tets() const { return m_tets; }
    const std::vector<Int2> &edges() const { return m_edges; }

protected:
    std::vector<Vector3> m_vertices, m_vertices0;
    std::vector<Int4> m_tets;
    std::vector<Int2> m_edges;
};
}  // namespace spgx
