"""Per-element parity on the GPU (test plan item 2): every energy kind's vertex force /
3x3 Hessian block (what VBD accumulates per incidence, energy.h:240-258) and constraint values / Jacobian (what the
generic XPBD projection consumes, energy.h:117-153, :289) from the CUDA dual-number evaluators against the
reference's TinyAD path, on perturbed scene stencils and on the energyViewer-style single-stencil fixtures."""
import numpy as np
import pytest

import oracle_api
from spg_b200 import api, scenes as S

pytestmark = pytest.mark.gpu
RTOL = 1e-9   # north_star tolerance; observed agreement is ~1e-13


def _perturb(sc, amp, seed=3):
    rng = np.random.default_rng(seed)
    sc.x = sc.x + amp * rng.standard_normal(sc.x.shape)
    return sc


def _relclose(a, b, rtol=RTOL):
    a, b = np.asarray(a), np.asarray(b)
    scale = max(np.abs(b).max(), 1e-30)
    assert np.abs(a - b).max() <= rtol * scale, (a, b)


def _compare_group(ctx, oracle, sc, solver_kind, n_samples=12):
    g = api.Group.from_scene(ctx, sc)
    g.finalize(solver_kind)
    og = oracle_api.OracleGroup(oracle, sc)
    for e, en in enumerate(sc.energies):
        n = en.stencils.shape[0]
        for i in np.unique(np.linspace(0, n - 1, n_samples).astype(int)):
            for local in range(S.STENCIL_SIZE[en.kind]):
                fr, hr = og.vertex_force_hessian(e, int(i), local)
                # 0: generic dual numbers, 1: the kind's hand-derived block (CSR path), 2: the canonical stream
                # evaluator behind the host-side canonicalisation (what the streaming kernel runs, streams.cuh)
                for hand in (0, 1, 2):
                    f, h = g.eval_vertex_force_hessian(e, int(i), local, hand)
                    _relclose(f, fr)
                    _relclose(h, hr)
            c, gc = g.eval_constraints(e, int(i))
            cr, gr = og.constraints(e, int(i))
            _relclose(c, cr)
            _relclose(gc, gr)
    g.close()


@pytest.mark.parametrize("membrane,bending", [(S.K_MEMBRANE_STVK, S.K_DISCRETE_BENDING),
                                              (S.K_SPRING, S.K_QUADRATIC_BENDING),
                                              (S.K_MEMBRANE_CHOI, S.K_QUADRATIC_BENDING),
                                              (S.K_MEMBRANE_BW, S.K_DISCRETE_BENDING),
                                              (S.K_SPRING_CONTINUUM, S.K_DISCRETE_BENDING),
                                              (S.K_SPRING_SQUARED_CONTINUUM, S.K_BENDING_BW)])
@pytest.mark.parametrize("solver_kind", [api.SPGX_VBD, api.SPGX_XPBD])
def test_cloth_elements(ctx, oracle, membrane, bending, solver_kind):
    sc = _perturb(S.cloth(8, 8, membrane, bending), 0.02)
    _compare_group(ctx, oracle, sc, solver_kind)


@pytest.mark.parametrize("kind", [S.K_TET_SNH, S.K_TET_STVK, S.K_TET_SSNH])
def test_tet_elements(ctx, oracle, kind):
    sc = _perturb(S.tet_beam(5, 3, 3, kind), 0.03)
    _compare_group(ctx, oracle, sc, api.SPGX_VBD)


def test_zero_guards(ctx, oracle):
    """anchor at its target and an undeformed StVK tet: the C(x) forms hit their `val == 0` guards
    (springAnchorEnergy.cpp:16-18, stvkEnergy.cpp:36-38) and must return zero gradients, not NaN."""
    sc = S.tet_beam(3, 2, 2, S.K_TET_STVK)
    g = api.Group.from_scene(ctx, sc)
    g.finalize(api.SPGX_XPBD)
    og = oracle_api.OracleGroup(oracle, sc)
    for e in range(2):
        c, gc = g.eval_constraints(e, 0)
        cr, gr = og.constraints(e, 0)
        assert np.all(np.isfinite(gc))
        np.testing.assert_allclose(c, cr, atol=1e-12)
        np.testing.assert_allclose(gc, gr, atol=1e-9)
    g.close()


def test_energy_viewer_fixtures(ctx, oracle):
    """Single-stencil known-answer inputs in the spirit of apps/energyViewer.cpp:20-137: a non-flat hinge, a
    stretched triangle, a sheared tet."""
    x0 = np.array([[0, 0, 0], [1, 0, 0], [0.3, 0.9, 0], [0.4, -0.8, 0]], float)
    x = np.array([[0, 0, 0.1], [1.1, 0, 0], [0.3, 0.8, 0.4], [0.5, -0.7, 0.3]], float)
    for kind, params in [(S.K_DISCRETE_BENDING, [[0.0, 0.7]]), (S.K_QUADRATIC_BENDING, [[0.0, 0.7]])]:
        sc = S.Scene(x.copy(), x0.copy(), np.ones(4))
        sc.energies.append(S.SceneEnergy(kind, np.array([[0, 1, 2, 3]], np.int32), np.array(params)))
        _compare_group(ctx, oracle, sc, api.SPGX_VBD)
    for kind, params in [(S.K_MEMBRANE_STVK, [[3.0, 0.3]]), (S.K_MEMBRANE_CHOI, [[3.0, 2.0, 1.0]]),
                         (S.K_MEMBRANE_BW, [[3.0, 2.0, 1.0]])]:
        sc = S.Scene(x[:3].copy(), x0[:3].copy(), np.ones(3))
        sc.energies.append(S.SceneEnergy(kind, np.array([[0, 1, 2]], np.int32), np.array(params)))
        _compare_group(ctx, oracle, sc, api.SPGX_VBD)
    t0 = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], float)
    t = t0 + np.array([[0, 0, 0], [0.2, 0.1, 0], [0.1, 0.3, -0.1], [0.05, 0.2, 0.4]])
    for kind in (S.K_TET_SNH, S.K_TET_STVK):
        sc = S.Scene(t.copy(), t0.copy(), np.ones(4))
        sc.energies.append(S.SceneEnergy(kind, np.array([[0, 1, 2, 3]], np.int32), np.array([[10.0, 0.3]])))
        _compare_group(ctx, oracle, sc, api.SPGX_VBD)
    for kind in (S.K_SPRING, S.K_SPRING_CONTINUUM):
        sc = S.Scene(x[:2].copy(), x0[:2].copy(), np.ones(2))
        sc.energies.append(S.SceneEnergy(kind, np.array([[0, 1]], np.int32), np.array([[0.9, 5.0]])))
        _compare_group(ctx, oracle, sc, api.SPGX_VBD)
