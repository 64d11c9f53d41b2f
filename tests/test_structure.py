"""Colouring and incidence are bit-exact with the reference:
spgx_color_vertices / spgx_color_stencils (pure host, linear time) vs coloring::colorVertices / colorStencils
(spg/source/utils/graphColoring.cpp:51-84) executed from the compiled reference."""
import numpy as np
import pytest

from spg_b200 import api, scenes as S


def _sets(sc):
    return [e.stencils for e in sc.energies]


@pytest.mark.parametrize("n", [3, 4, 7, 16, 64, 128])
def test_vertex_colors_cloth(oracle, n):
    sc = S.cloth(n, n)
    mine = api.color_vertices(sc.n_vertices, _sets(sc))
    ref = oracle.color_vertices(sc.n_vertices, _sets(sc))
    assert np.array_equal(mine, ref)
    if n == 64:  # class sizes at 64^2
        assert np.bincount(mine).tolist() == [1024, 1024, 543, 512, 467, 466, 60]


@pytest.mark.parametrize("dims", [(2, 2, 2), (5, 2, 2), (9, 3, 3), (17, 5, 5)])
def test_vertex_colors_beam(oracle, dims):
    sc = S.tet_beam(*dims)
    assert np.array_equal(api.color_vertices(sc.n_vertices, _sets(sc)), oracle.color_vertices(sc.n_vertices, _sets(sc)))


@pytest.mark.parametrize("n", [3, 4, 7, 16, 64])
def test_stencil_colors_cloth(oracle, n):
    sc = S.cloth(n, n, S.K_SPRING, S.K_QUADRATIC_BENDING)
    for e in sc.energies:
        assert np.array_equal(api.color_stencils(e.stencils), oracle.color_stencils(e.stencils))
    if n == 64:  # 18 + 8 + 1 colours
        assert [int(api.color_stencils(e.stencils).max()) + 1 for e in sc.energies] == [18, 8, 1]


def test_stencil_colors_tets(oracle):
    sc = S.tet_beam(9, 3, 3)
    for e in sc.energies:
        assert np.array_equal(api.color_stencils(e.stencils), oracle.color_stencils(e.stencils))


def test_random_graphs(oracle):
    rng = np.random.default_rng(7)
    for trial in range(20):
        nv = int(rng.integers(1, 60))
        sets = []
        for _ in range(int(rng.integers(1, 4))):
            size = int(rng.integers(1, 5))
            n = int(rng.integers(0, 80))
            sets.append(rng.integers(0, nv, size=(n, size)).astype(np.int32))  # repeated vertices allowed
        assert np.array_equal(api.color_vertices(nv, sets), oracle.color_vertices(nv, sets)), trial
        for s in sets:
            if s.shape[0]:
                assert np.array_equal(api.color_stencils(s), oracle.color_stencils(s)), trial


def test_isolated_vertices_and_empty_sets():
    assert api.color_vertices(5, [np.zeros((0, 2), np.int32)]).tolist() == [0, 0, 0, 0, 0]
    assert api.color_vertices(0, [np.zeros((0, 2), np.int32)]).size == 0
    assert api.color_stencils(np.zeros((0, 3), np.int32)).size == 0
    # anchors (S = 1) never create edges (graphColoring.cpp:77)
    assert api.color_vertices(3, [np.array([[0], [1], [2]], np.int32)]).tolist() == [0, 0, 0]


def test_invalid_index_raises():
    with pytest.raises(api.SpgxError, match="invalid node indices"):
        api.color_vertices(2, [np.array([[0, 5]], np.int32)])


def test_colouring_is_linear_time():
    """The reference's greedy colouring is O(N^2) (graphColoring.cpp:29,:33: 2.5 s at 512^2); ours must not be."""
    import time
    sc = S.cloth(512, 512)
    t = time.time()
    c = api.color_vertices(sc.n_vertices, _sets(sc))
    assert time.time() - t < 2.0
    assert c.max() + 1 == 7


# ---- edge cases of the two colourings, against the compiled reference where it can run them ----
@pytest.mark.parametrize("n,sets", [
    (6, [np.array([[0, 1], [1, 2]], np.int32)]),                                   # isolated vertices 3..5
    (6, [np.array([[0, 1], [1, 2]], np.int32), np.empty((0, 2), np.int32)]),       # an energy without stencils
    (5, [np.empty((0, 2), np.int32)]),                                             # no stencil at all
    (1, [np.array([[0]], np.int32)]),                                              # one vertex, one anchor
    (3, [np.array([[0, 1, 2], [2, 1, 0]], np.int32)]),                             # the same triangle twice
    (4, [np.array([[0, 1], [0, 1], [1, 0], [2, 3]], np.int32)]),                   # duplicate and reversed springs
])
def test_vertex_colouring_edge_cases_match_reference(n, sets):
    import oracle_api
    from spg_b200 import api
    if not oracle_api.available("reference"):
        pytest.skip("needs the compiled reference")
    ref = oracle_api.load("reference").color_vertices(n, sets)
    assert np.array_equal(api.color_vertices(n, sets), ref)


@pytest.mark.parametrize("stencils", [
    np.array([[0, 1], [1, 2]], np.int32),
    np.array([[0, 1], [0, 1], [1, 0]], np.int32),          # three stencils on the same two vertices: three colours
    np.array([[0], [1], [2]], np.int32),                   # anchors never conflict: one colour
    np.array([[0, 1, 2, 3]], np.int32),
])
def test_stencil_colouring_edge_cases_match_reference(stencils):
    import oracle_api
    from spg_b200 import api
    if not oracle_api.available("reference"):
        pytest.skip("needs the compiled reference")
    ref = oracle_api.load("reference").color_stencils(stencils)
    assert np.array_equal(api.color_stencils(stencils), ref)


def test_empty_stencil_set_is_defined_here():
    """Graph::greedyColoring writes result[0] unconditionally (graphColoring.cpp:26): on an energy without stencils the
    reference's colorStencils is undefined behaviour (it crashes in this container). Here it is an empty colouring."""
    from spg_b200 import api
    assert api.color_stencils(np.empty((0, 2), np.int32)).size == 0
    assert api.color_vertices(0, [np.empty((0, 2), np.int32)]).size == 0
