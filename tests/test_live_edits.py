"""Live topology edits: the picking sequence of the reference application -- add a
SpringAnchorEnergy on one particle while the simulation runs, drag its anchor, release it (removeEnergy) -- with the solver
flagged by requirePrecomputationUpdate each time (apps/solverCompare.cpp:717-726, :1446-1455, :1471-1473, :1499-1502, :1524).
The CUDA path goes through spgx_group_add_energy / spgx_group_remove_energy / spgx_group_finalize on a group that is already
stepping: colours, incidence streams and the block-sparse pattern are rebuilt, positions and velocities carried over."""
import numpy as np
import pytest

import oracle_api
from spg_b200 import api, scenes as S

PICK_STIFFNESS = 1e4  # solverCompare.cpp:722


def _rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)


SOLVERS = {
    # name: (product solver kind, oracle solver kind, substeps, iterations, product step, tolerance after the sequence)
    "vbd": (api.SPGX_VBD, 0, 2, 10, lambda g, n: g.vbd_step(0.01, 2, 10, frames=n), 1e-7),
    "xpbd": (api.SPGX_XPBD, 1, 8, 1, lambda g, n: g.xpbd_step(0.01, 8, 1, frames=n), 1e-7),
    "newton_dx": (api.SPGX_NEWTON, 4, 2, 1, lambda g, n: g.implicit_step(api.SPGX_IE_NEWTON_DX, 0.01, 2, frames=n), 1e-6),
}


def _picking_sequence(step, add, drag, release, state):
    """frames free -> pick -> drag twice -> release -> frames free; the states after every phase"""
    out = []
    step(2)
    out.append(state())
    add()
    step(2)
    out.append(state())
    for k in (1, 2):
        drag(k)
        step(1)
        out.append(state())
    release()
    step(2)
    out.append(state())
    return out


# ---- CPU: the restatement follows the compiled reference through the same edits ----
@pytest.mark.parametrize("name", ["vbd", "xpbd", "newton_dx"])
def test_port_matches_reference_through_picking(name):
    if not (oracle_api.available("reference") and oracle_api.available("port")):
        pytest.skip("needs both oracles")
    _, okind, substeps, iters, _, _ = SOLVERS[name]
    sc = S.cloth(10, 10)
    picked = 37
    states = []
    for which in ("reference", "port"):
        o = oracle_api.load(which)
        og = oracle_api.OracleGroup(o, sc)
        sol = oracle_api.OracleSolver(og, okind, 0.01, substeps, iters)
        n_e = len(sc.energies)
        p0 = sc.x[picked].copy()
        states.append(_picking_sequence(
            sol.step,
            lambda: sol.add_energy(S.K_ANCHOR, np.array([[picked]], np.int32), np.array([[*p0, PICK_STIFFNESS]])),
            lambda k: sol.update_anchor(n_e, 0, p0 + np.array([0.02 * k, 0.03 * k, -0.01 * k])),
            lambda: sol.remove_energy(n_e),
            lambda: sol.state()))
    for (xr, vr), (xp, vp) in zip(*states):
        assert _rel(xp, xr) <= 1e-8 and _rel(vp, vr) <= 1e-6


# ---- GPU: the product through the C ABI ----
@pytest.mark.gpu
@pytest.mark.parametrize("name", ["vbd", "xpbd", "newton_dx"])
def test_gpu_picking_sequence_matches_oracle(ctx, oracle, name):
    kind, okind, substeps, iters, pstep, tol = SOLVERS[name]
    sc = S.cloth(12, 12)
    picked = 77
    n_e = len(sc.energies)
    p0 = sc.x[picked].copy()
    og = oracle_api.OracleGroup(oracle, sc)
    sol = oracle_api.OracleSolver(og, okind, 0.01, substeps, iters)
    ref = _picking_sequence(
        sol.step,
        lambda: sol.add_energy(S.K_ANCHOR, np.array([[picked]], np.int32), np.array([[*p0, PICK_STIFFNESS]])),
        lambda k: sol.update_anchor(n_e, 0, p0 + np.array([0.02 * k, 0.03 * k, -0.01 * k])),
        lambda: sol.remove_energy(n_e),
        lambda: sol.state())

    g = api.Group.from_scene(ctx, sc)
    g.finalize(kind)
    ids = {}

    def add():
        ids["pick"] = g.add_energy(S.K_ANCHOR, np.array([[picked]], np.int32), np.array([[*p0, PICK_STIFFNESS]]))
        with pytest.raises(api.SpgxError):
            pstep(g, 1)  # requirePrecomputationUpdate is not optional: an edited group refuses to step
        g.finalize(kind)

    def release():
        g.remove_energy(ids["pick"])
        with pytest.raises(api.SpgxError):
            pstep(g, 1)
        g.finalize(kind)

    got = _picking_sequence(
        lambda n: pstep(g, n), add,
        lambda k: g.update_record(ids["pick"], 0, 0, p0 + np.array([0.02 * k, 0.03 * k, -0.01 * k])),
        release, lambda: g.download())
    assert ids["pick"] == n_e
    for phase, ((x, v), (xr, vr)) in enumerate(zip(got, ref)):
        assert _rel(x, xr) <= tol and _rel(v, vr) <= 100 * tol, (phase, _rel(x, xr), _rel(v, vr))
    # the picked particle really was dragged, and released afterwards
    assert np.linalg.norm(got[3][0][picked] - got[0][0][picked]) > 0.02


@pytest.mark.gpu
def test_gpu_remove_energy_shifts_ids_and_rejects_bad_ids(ctx, oracle):
    """removeEnergy erases from the energy vector (particleGroup.cpp:89-95): later energies move down one place."""
    sc = S.cloth(8, 8)
    g = api.Group.from_scene(ctx, sc)
    g.finalize(api.SPGX_VBD)
    n_e = len(sc.energies)
    with pytest.raises(api.SpgxError):
        g.remove_energy(n_e)
    with pytest.raises(api.SpgxError):
        g.remove_energy(-1)
    # drop the bending energy: the scene is now membrane + anchors, as if it had been built that way
    bend = [i for i, e in enumerate(sc.energies) if e.kind == S.K_DISCRETE_BENDING][0]
    g.remove_energy(bend)
    g.finalize(api.SPGX_VBD)
    g.vbd_step(0.01, 1, 10, frames=2)
    sc2 = S.Scene(sc.x.copy(), sc.x0.copy(), sc.mass.copy())
    sc2.energies = [e for i, e in enumerate(sc.energies) if i != bend]
    og = oracle_api.OracleGroup(oracle, sc2)
    sol = oracle_api.OracleSolver(og, 0, 0.01, 1, 10)
    sol.step(2)
    x, v = g.download()
    xr, vr = sol.state()
    assert _rel(x, xr) <= 1e-9 and _rel(v, vr) <= 1e-7
