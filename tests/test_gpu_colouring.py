"""Colouring built on the device (spg_b200/csrc/colour.cu) == the host colouring == the
reference's Graph::greedyColoring (spg/source/utils/graphColoring.cpp:21-84), bit for bit: cloths, beams, graphs without any
locality, duplicate / reversed stencils, isolated vertices, more than 64 colours, and BASELINE config 3 at full size."""
import numpy as np
import pytest

import oracle_api
from spg_b200 import api, scenes as S

pytestmark = pytest.mark.gpu


def _sets(sc):
    return [e.stencils for e in sc.energies]


@pytest.mark.parametrize("name", ["cloth7", "cloth64x37", "cloth_quadratic", "beam", "beam_stvk"])
def test_device_colouring_equals_reference(ctx, ref_oracle, name):
    sc = {"cloth7": lambda: S.cloth(7, 7), "cloth64x37": lambda: S.cloth(64, 37),
          "cloth_quadratic": lambda: S.cloth(33, 20, S.K_SPRING, S.K_QUADRATIC_BENDING),
          "beam": lambda: S.tet_beam(13, 4, 5, S.K_TET_SNH), "beam_stvk": lambda: S.tet_beam(6, 6, 6, S.K_TET_STVK)}[name]()
    dev, rounds = api.color_vertices_device(ctx, sc.n_vertices, _sets(sc))
    assert np.array_equal(dev, api.color_vertices(sc.n_vertices, _sets(sc)))
    assert np.array_equal(dev, ref_oracle.color_vertices(sc.n_vertices, _sets(sc)))   # the reference's own O(N^2) greedy
    assert 1 <= rounds <= sc.n_vertices


def test_device_colouring_on_graphs_without_locality(ctx):
    rng = np.random.default_rng(12)
    for n, m, S_ in [(500, 900, 2), (1200, 700, 4), (300, 2000, 3), (4000, 9000, 2)]:
        st = np.stack([rng.permutation(n)[:S_] for _ in range(m)]).astype(np.int32)       # distinct vertices per stencil
        st2 = np.concatenate([st[: m // 3], st[: m // 5][:, ::-1]])                            # duplicates and reversed copies
        sets = [st, st2]
        dev, _ = api.color_vertices_device(ctx, n, sets)
        assert np.array_equal(dev, api.color_vertices(n, sets)), (n, m, S_)
    # isolated vertices (no stencil at all) take colour 0; a single vertex; an empty group
    dev, _ = api.color_vertices_device(ctx, 10, [np.array([[2, 7]], np.int32)])
    assert np.array_equal(dev, [0, 0, 0, 0, 0, 0, 0, 1, 0, 0])
    assert np.array_equal(api.color_vertices_device(ctx, 1, [np.zeros((0, 2), np.int32)])[0], [0])
    assert api.color_vertices_device(ctx, 0, [np.zeros((0, 2), np.int32)])[0].size == 0
    with pytest.raises(api.SpgxError, match="invalid node indices"):
        api.color_vertices_device(ctx, 5, [np.array([[1, 9]], np.int32)])


def test_device_colouring_beyond_one_64_colour_window(ctx):
    """a clique of 150 vertices needs 150 colours: the first-fit search walks windows of 64 colours"""
    n = 150
    i, j = np.triu_indices(n, 1)
    st = np.stack([i, j], 1).astype(np.int32)
    dev, rounds = api.color_vertices_device(ctx, n, [st])
    assert np.array_equal(dev, np.arange(n)) and rounds == n
    assert np.array_equal(dev, api.color_vertices(n, [st]))


def test_finalize_uses_the_device_colouring_at_size(ctx, monkeypatch):
    """BASELINE config 3 (1 M vertices): finalize colours on the device, the colours equal the host's, SURVEY 8's class sizes"""
    sc = S.cloth(1024, 1024)
    monkeypatch.delenv("SPGX_DEVICE_COLOURING", raising=False)
    g = api.Group.from_scene(ctx, sc)
    g.finalize(api.SPGX_VBD)
    info = g.kernel_info()
    assert info["colouring_rounds"] > 0
    dev = g.vertex_colors()
    g.close()
    assert np.array_equal(dev, api.color_vertices(sc.n_vertices, _sets(sc)))
    assert np.bincount(dev).tolist() == [262144, 262144, 131583, 131072, 130307, 130306, 1020]
    monkeypatch.setenv("SPGX_DEVICE_COLOURING", "0")
    g = api.Group.from_scene(ctx, S.cloth(40, 40))
    g.finalize(api.SPGX_VBD)
    assert g.kernel_info()["colouring_rounds"] == 0
    g.close()
