"""Rigid bodies: RigidBodyGroup under VBD and XPBD on the device against the compiled reference
(spg/source/sim/simObject/rigidBodyGroup.cpp, the RigidBodyGroup branches of vbd.cpp / xpbd.cpp, energy.h:117-153 and the four
spg/source/sim/energy/rigidBody*.cpp energies) on the reference application's three rigid-body scenes
(apps/solverCompare.cpp:524-660)."""
import numpy as np
import pytest

import oracle_api
from spg_b200 import api, scenes as S

SCENES = {"anchored": lambda: S.rb_anchored(), "articulated": lambda: S.rb_articulated(5), "chain": lambda: S.rb_chain(7)}


def _rel(a, b):
    return float(np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(b).max(), 1e-30))


def _noisy_state(sc, seed, scale=1.0):
    rng = np.random.default_rng(seed)
    n = sc.n_bodies
    return (sc.x + 0.05 * scale * rng.standard_normal((n, 3)), 0.4 * scale * rng.standard_normal((n, 3)),
            0.3 * scale * rng.standard_normal((n, 3)), 0.8 * scale * rng.standard_normal((n, 3)))


def test_rb_scenes_are_consistent():
    """host-only: the scene restatements (solverCompare.cpp:524-660) have the sizes the C ABI expects"""
    for name, make in SCENES.items():
        sc = make()
        assert sc.local_inertia.shape == (sc.n_bodies, 3, 3)
        for e in sc.energies:
            assert e.stencils.shape[1] == api.lib.spgx_rb_kind_stencil_size(e.kind)
            assert e.params.shape == (e.stencils.shape[0], api.lib.spgx_rb_kind_num_model_params(e.kind))
            assert e.stencils.min() >= 0 and e.stencils.max() < sc.n_bodies
    assert api.lib.spgx_rb_kind_stencil_size(4) == -1
    # the chain's rest length is a FLOAT square root widened to double (solverCompare.cpp:654)
    assert SCENES["chain"]().energies[1].params[0, 6] == float(np.sqrt(np.float32(2)))


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(SCENES))
def test_rb_elements_match_reference(ctx, ref_oracle, name):
    """constraint value + Jacobian over the 6 S stencil DOFs (XPBD) and the owner body's force + 6 x 6 block (VBD) on a rotated,
    displaced state"""
    sc = SCENES[name]()
    x, th, v, om = _noisy_state(sc, 3)
    og = oracle_api.OracleRbGroup(ref_oracle, sc)
    og.set_state(x, th, v, om)
    g = api.RbGroup.from_scene(ctx, sc)
    g.finalize(api.SPGX_VBD)
    g.upload(x, th, v, om)
    # setPositions rebuilt the rotation matrices from theta on both sides
    assert _rel(g.download()[4], og.state()[4]) <= 1e-14
    for e, en in enumerate(sc.energies):
        for st in range(en.stencils.shape[0]):
            c, j = g.eval_constraint(e, st)
            cr, jr = og.constraint(e, st)
            assert abs(c - cr) <= 1e-12 * max(abs(cr), 1.0), (e, st)
            assert np.abs(j - jr).max() <= 1e-11 * max(np.abs(jr).max(), 1.0), (e, st)
            for local in range(en.stencils.shape[1]):
                f, h = g.eval_vertex(e, st, local)
                fr, hr = og.vertex_force_hessian(e, st, local)
                assert np.abs(f - fr).max() <= 1e-10 * max(np.abs(fr).max(), 1.0), (e, st, local)
                assert np.abs(h - hr).max() <= 1e-10 * max(np.abs(hr).max(), 1.0), (e, st, local)


@pytest.mark.gpu
@pytest.mark.parametrize("solver,substeps,iters", [(api.SPGX_VBD, 1, 10), (api.SPGX_VBD, 3, 4), (api.SPGX_XPBD, 20, 1),
                                                   (api.SPGX_XPBD_SERIAL, 10, 1)])
@pytest.mark.parametrize("name", sorted(SCENES))
def test_rb_step_matches_reference(ctx, ref_oracle, name, solver, substeps, iters):
    """positions, rotation matrices, axis-angle vectors, linear and angular velocities after one frame from the scene's rest
    state AND from a tumbling state; bounded drift over five more frames"""
    sc = SCENES[name]()
    kind = {api.SPGX_VBD: 0, api.SPGX_XPBD: 1, api.SPGX_XPBD_SERIAL: 2}[solver]
    for start in ("rest", "tumbling"):
        og = oracle_api.OracleRbGroup(ref_oracle, sc)
        osol = oracle_api.OracleRbSolver(og, kind, 0.01, substeps, iters)
        g = api.RbGroup.from_scene(ctx, sc)
        g.finalize(solver)
        if start == "tumbling":
            st = _noisy_state(sc, 11, 0.5)
            osol.set_state(*st)
            g.upload(*st)
        osol.step(1)
        g.step(0.01, substeps, iters)
        got, want = g.download(), osol.state()
        for k, tol in zip(range(5), (1e-9, 1e-8, 1e-7, 1e-7, 1e-9)):   # x, theta, v, omega, R
            assert _rel(got[k], want[k]) <= tol, (start, "x theta v omega R".split()[k], _rel(got[k], want[k]))
        osol.step(5)
        g.step(0.01, substeps, iters, frames=5)
        got, want = g.download(), osol.state()
        assert _rel(got[0], want[0]) <= 1e-6 and _rel(got[4], want[4]) <= 1e-6, start
        assert np.isfinite(got[0]).all()
        g.close()


@pytest.mark.gpu
def test_rb_colours_reset_and_live_edits(ctx, ref_oracle):
    """body colours = the reference's greedy colouring over bodies; reset; updateAnchor and removeEnergy + re-finalize on a
    stepping group (the picking sequence of apps/solverCompare.cpp:728-737, :1475-1502 for rigid bodies)"""
    sc = SCENES["chain"]()
    ref_colors = api.color_vertices(sc.n_bodies, [e.stencils for e in sc.energies])
    g = api.RbGroup.from_scene(ctx, sc)
    g.finalize(api.SPGX_VBD)
    assert np.array_equal(g.body_colors(), ref_colors)   # (spgx colouring == reference colouring: tests/test_structure.py)
    og = oracle_api.OracleRbGroup(ref_oracle, sc)
    osol = oracle_api.OracleRbSolver(og, 0, 0.01, 2, 6)
    g.step(0.01, 2, 6, frames=2)
    osol.step(2)
    # drag the anchor (RBSpringAnchorEnergy::updateAnchor)
    g.update_anchor(0, 0, [0.7, 0.6, 0.1])
    osol.update_anchor(0, 0, [0.7, 0.6, 0.1])
    g.step(0.01, 2, 6, frames=2)
    osol.step(2)
    assert _rel(g.download()[0], osol.state()[0]) <= 1e-8
    # reset returns to the initial pose with zero velocities
    g.reset()
    x, th, v, om, R = g.download()
    assert np.array_equal(x, sc.x) and not v.any() and not om.any() and np.array_equal(R, np.tile(np.eye(3), (sc.n_bodies, 1, 1)))
    # removing the chain's springs leaves free bodies under gravity (except the anchored one); the state carries over
    g.step(0.01, 1, 5)
    before = g.download()
    g.remove_energy(1)
    g.finalize(api.SPGX_VBD)
    after = g.download()
    assert all(np.array_equal(a, b) for a, b in zip(before, after))
    g.step(0.01, 1, 5)
    x2, _, v2, _, _ = g.download()
    assert np.allclose(v2[1:, 1], before[2][1:, 1] - 9.8 * 0.01, rtol=0, atol=1e-12)   # free fall of the released bodies
    # validation messages of the energies' addStencil
    with pytest.raises(api.SpgxError, match="rest length"):
        g.add_energy(api.SPGX_RB_SPRING, [[0, 1]], [[0, 0, 0, 0, 0, 0, 0.0, 1.0]])
    with pytest.raises(api.SpgxError, match="Negative rest length"):
        g.add_energy(api.SPGX_RB_SPRING_SQUARED, [[0, 1]], [[0, 0, 0, 0, 0, 0, -1.0, 1.0]])
    with pytest.raises(api.SpgxError, match="out of range"):
        g.add_energy(api.SPGX_RB_SPRING, [[0, 99]], [[0, 0, 0, 0, 0, 0, 1.0, 1.0]])
