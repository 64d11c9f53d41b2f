"""The vectorised numpy scene builders reproduce the reference-side builders bit for bit
(apps/solverCompare.cpp:118-189, :260-330, :332-498 as restated around the reference's own TriangleMesh in
oracle/ref_driver.cpp)."""
import numpy as np
import pytest

from spg_b200 import scenes as S


def _same(a: S.Scene, b: S.Scene):
    assert np.array_equal(a.x, b.x)
    assert np.array_equal(a.x0, b.x0)
    assert np.array_equal(a.mass, b.mass)
    assert len(a.energies) == len(b.energies)
    for ea, eb in zip(a.energies, b.energies):
        assert ea.kind == eb.kind
        assert np.array_equal(ea.stencils, eb.stencils)
        assert np.array_equal(ea.params, eb.params)


@pytest.mark.parametrize("n", [3, 4, 7, 16, 64])
@pytest.mark.parametrize("membrane,bending", [(S.K_MEMBRANE_STVK, S.K_DISCRETE_BENDING),
                                              (S.K_SPRING, S.K_QUADRATIC_BENDING),
                                              (S.K_MEMBRANE_CHOI, S.K_QUADRATIC_BENDING)])
def test_cloth_matches_reference_builder(ref_oracle, n, membrane, bending):
    _same(S.cloth(n, n, membrane, bending), ref_oracle.scene("cloth", n, n, membrane, bending))


def test_cloth_rectangular(ref_oracle):
    _same(S.cloth(5, 9), ref_oracle.scene("cloth", 5, 9, S.K_MEMBRANE_STVK, S.K_DISCRETE_BENDING))


@pytest.mark.parametrize("dims", [(2, 2, 2), (5, 2, 2), (9, 3, 3), (17, 5, 5), (6, 4, 3)])
@pytest.mark.parametrize("kind", [S.K_TET_SNH, S.K_TET_STVK])
def test_beam_matches_reference_builder(ref_oracle, dims, kind):
    _same(S.tet_beam(*dims, kind), ref_oracle.scene("beam", *dims, kind))


def test_counts_match_survey():
    # T = 2(n-1)^2, B = (3n-5)(n-1)
    sc = S.cloth(64, 64)
    assert sc.energies[1].stencils.shape[0] == 7938
    assert sc.energies[0].stencils.shape[0] == 11781
    assert sc.n_vertices == 4096
    assert abs(sc.mass.sum() - 1.0) < 1e-12
