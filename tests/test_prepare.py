"""addStencil + preparePrecomputations restated on the host (spg_b200/csrc/energy_prepare.h) against the
reference's own precomputed members (restThetas, laplacians, inverseReferenceMats, effectiveStiffness /
effectiveCompliance, hydrostaticCorrectionTerms)."""
import numpy as np
import pytest

import oracle_api
from spg_b200 import api, scenes as S

RTOL = 1e-14


def _close(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    assert np.allclose(a, b, rtol=RTOL, atol=1e-300), (a, b)


def _check_energy(og, sc, e):
    en = sc.energies[e]
    rec = api.prepare_energy(en.kind, en.stencils, en.params, sc.x0)
    for i in np.unique(np.linspace(0, en.stencils.shape[0] - 1, 25).astype(int)):
        K, A = og.stiffness(e, i)
        rest = og.rest_data(e, i)
        r = rec[i]
        k = en.kind
        if k in (S.K_SPRING, S.K_SPRING_CONTINUUM, S.K_SPRING_SQUARED_CONTINUUM):
            _close(r, [en.params[i, 0], K[0, 0], A[0, 0]])
        elif k in (S.K_DISCRETE_BENDING, S.K_BENDING_BW):
            _close(r, [rest[0], K[0, 0], A[0, 0]])
        elif k == S.K_QUADRATIC_BENDING:
            L = rest.reshape(3, 12)
            _close(r[:4], [L[0, 0], L[0, 3], L[0, 6], L[0, 9]])
            assert np.array_equal(L, np.kron(np.array([L[0, 0::3]]), np.eye(3)))
            _close(r[4:], [K[0, 0], A[0, 0]])
            _close(K, K[0, 0] * np.eye(3))
        elif k == S.K_MEMBRANE_STVK:
            _close(r[:4], rest)
            _close(r[4:7], [K[0, 0], K[0, 1], K[2, 2]])
            _close(r[7:10], [A[0, 0], A[0, 1], A[2, 2]])
            assert K[0, 0] == K[1, 1] and K[0, 2] == 0 and K[1, 2] == 0
        elif k == S.K_MEMBRANE_CHOI:
            _close(r[:4], rest)
            _close(r[4:8], np.diag(K))
            _close(r[8:12], np.diag(A))
        elif k == S.K_MEMBRANE_BW:
            _close(r[:4], rest)
            _close(r[4:7], np.diag(K))
            _close(r[7:10], np.diag(A))
        elif k in (S.K_TET_STVK, S.K_TET_SSNH):
            _close(r[:9], rest[:9])
            _close(r[9:11], np.diag(K))
            _close(r[11:13], np.diag(A))
        elif k == S.K_TET_SNH:
            _close(r[:9], rest[:9])
            _close(r[9:11], np.diag(K))
            _close(r[11:13], np.diag(A))
            _close(r[13], rest[9])
        elif k == S.K_ANCHOR:
            _close(r, list(en.params[i]) + [A[0, 0]])
        elif k == S.K_STATIC_LOAD:
            _close(r, list(rest[:3]) + [K[0, 0], A[0, 0]])


@pytest.mark.parametrize("membrane,bending", [(S.K_MEMBRANE_STVK, S.K_DISCRETE_BENDING),
                                              (S.K_SPRING, S.K_QUADRATIC_BENDING),
                                              (S.K_MEMBRANE_CHOI, S.K_QUADRATIC_BENDING),
                                              (S.K_MEMBRANE_BW, S.K_DISCRETE_BENDING),
                                              (S.K_SPRING_CONTINUUM, S.K_DISCRETE_BENDING),
                                              (S.K_SPRING_SQUARED_CONTINUUM, S.K_BENDING_BW)])
def test_cloth_records(ref_oracle, membrane, bending):
    sc = S.cloth(9, 7, membrane, bending)
    og = oracle_api.OracleGroup(ref_oracle, sc)
    for e in range(3):
        _check_energy(og, sc, e)


def test_stvk_membrane_with_poisson(ref_oracle):
    sc = S.cloth(5, 5)
    sc.energies[1].params[:, 1] = 0.3
    og = oracle_api.OracleGroup(ref_oracle, sc)
    _check_energy(og, sc, 1)


@pytest.mark.parametrize("kind", [S.K_TET_SNH, S.K_TET_STVK, S.K_TET_SSNH])
def test_beam_records(ref_oracle, kind):
    sc = S.tet_beam(5, 3, 3, kind)
    og = oracle_api.OracleGroup(ref_oracle, sc)
    for e in range(2):
        _check_energy(og, sc, e)


def test_static_load_records(ref_oracle):
    import sys
    from pathlib import Path
    sys.path.insert(0, str(Path(__file__).resolve().parent / "golden"))
    import make_golden as MG
    sc = MG.beam_with_load(S.K_TET_SSNH)
    og = oracle_api.OracleGroup(ref_oracle, sc)
    _check_energy(og, sc, 2)


@pytest.mark.parametrize("kind,stencil,params,x0,msg", [
    (S.K_SPRING, [[0, 1]], [[0.0, 1.0]], [[0, 0, 0], [1, 0, 0]], "rest length"),
    (S.K_SPRING_CONTINUUM, [[0, 1]], [[-1.0, 1.0]], [[0, 0, 0], [1, 0, 0]], "rest length"),
    (S.K_DISCRETE_BENDING, [[0, 1, 2, 3]], [[3.13, 1.0]], [[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, -1, 0]], "restTheta"),
    (S.K_QUADRATIC_BENDING, [[0, 1, 2, 3]], [[0.1, 1.0]], [[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, -1, 0]], "Rest theta"),
    (S.K_MEMBRANE_STVK, [[0, 1, 2]], [[1.0, 0.0]], [[0, 0, 0], [1, 0, 0], [0, 1, 1]], "coordinate z=0"),
    (S.K_MEMBRANE_CHOI, [[0, 2, 1]], [[1.0, 1.0, 1.0]], [[0, 0, 0], [1, 0, 0], [0, 1, 0]], "not positive"),
    (S.K_TET_SNH, [[0, 2, 1, 3]], [[1.0, 0.3]], [[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], "not positive"),
    (S.K_TET_SSNH, [[0, 2, 1, 3]], [[1.0, 0.3]], [[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], "SSNH tetrahedron"),
    (S.K_BENDING_BW, [[0, 1, 2, 3]], [[-3.13, 1.0]], [[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, -1, 0]], "restTheta"),
    (S.K_SPRING_SQUARED_CONTINUUM, [[0, 1]], [[0.0, 1.0]], [[0, 0, 0], [1, 0, 0]], "rest length"),
])
def test_validation_errors_match_reference(ref_oracle, kind, stencil, params, x0, msg):
    """Same std::runtime_error conditions as the reference's addStencil / preparePrecomputations."""
    x0 = np.array(x0, float)
    with pytest.raises(api.SpgxError, match=msg):
        api.prepare_energy(kind, np.array(stencil, np.int32), np.array(params, float), x0)
    sc = S.Scene(x0.copy(), x0.copy(), np.ones(x0.shape[0]))
    sc.energies.append(S.SceneEnergy(kind, np.array(stencil, np.int32), np.array(params, float)))
    with pytest.raises(oracle_api.OracleError, match=msg):
        oracle_api.OracleGroup(ref_oracle, sc)
