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


# ---- record edits on kinds that run through the streaming colour pass ----
@pytest.mark.gpu
@pytest.mark.parametrize("scene,eid", [("cloth", 0), ("cloth", 1), ("beam", 1)])
def test_gpu_record_edit_reaches_the_streams(ctx, oracle, scene, eid):
    """spgx_energy_update_record on discrete bending / StVK membrane / SNH tets of a group finalized for VBD: the colour
    pass reads those kinds from the canonical incidence streams, so the edit must be patched into them.  A group edited
    after finalize == a group built with the edited parameters from the start (bit for bit), and both follow the oracle."""
    sc = S.cloth(13, 11) if scene == "cloth" else S.tet_beam(6, 3, 3, S.K_TET_SNH)
    e = sc.energies[eid]

    def edited(p):   # bending: {restTheta + 0.05, 3 k}; membrane / tets: {3 young, poisson 0.3}
        return np.array([p[0] + 0.05, 3.0 * p[1]]) if e.kind == S.K_DISCRETE_BENDING else np.array([3.0 * p[0], 0.3])
    edits = {int(i): edited(e.params[i]) for i in (0, e.stencils.shape[0] // 2, e.stencils.shape[0] - 1)}
    edited_scene = S.Scene(sc.x.copy(), sc.x0.copy(), sc.mass.copy())
    edited_scene.energies = [S.SceneEnergy(q.kind, q.stencils.copy(), q.params.copy()) for q in sc.energies]
    for i, p in edits.items():
        edited_scene.energies[eid].params[i] = p
    ga = api.Group.from_scene(ctx, sc)
    ga.finalize(api.SPGX_VBD)
    assert ga.streamed()
    ga.vbd_step(0.01, 1, 5)                    # the edit lands on a group that is already stepping
    gb = api.Group.from_scene(ctx, edited_scene)
    gb.finalize(api.SPGX_VBD)
    rec2 = gb.records(eid)
    for i in edits:
        ga.update_record(eid, i, 0, rec2[i])
    assert np.array_equal(ga.records(eid), rec2)
    xa, va = ga.download()
    gb.upload(xa, va)
    ga.vbd_step(0.01, 1, 5, frames=2)
    gb.vbd_step(0.01, 1, 5, frames=2)
    xa2, va2 = ga.download()
    xb2, vb2 = gb.download()
    assert np.array_equal(xa2, xb2) and np.array_equal(va2, vb2)
    # and the edited scene follows the reference from the same state
    og = oracle_api.OracleGroup(oracle, edited_scene)
    osol = oracle_api.OracleSolver(og, 0, 0.01, 1, 5)
    osol.set_state(xa, va)
    osol.step(2)
    xr, vr = osol.state()
    assert _rel(xb2, xr) <= 1e-8 and _rel(vb2, vr) <= 1e-6
    # the edit must matter (otherwise the test proves nothing)
    gc = api.Group.from_scene(ctx, sc)
    gc.finalize(api.SPGX_VBD)
    gc.upload(xa, va)
    gc.vbd_step(0.01, 1, 5, frames=2)
    assert not np.array_equal(gc.download()[0], xa2)
    for g in (ga, gb, gc):
        g.close()


@pytest.mark.gpu
@pytest.mark.parametrize("solver", ["vbd", "xpbd"])
def test_gpu_async_transfers_equal_blocking_ones(ctx, solver):
    """spgx_upload_async / spgx_download_async around steps enqueued without synchronising == the blocking calls."""
    sc = S.cloth(21, 17) if solver == "vbd" else S.cloth(21, 17, S.K_SPRING, S.K_QUADRATIC_BENDING)

    def step(g, sync):
        if solver == "vbd":
            g.vbd_step(0.01, 1, 4, sync=sync)
        else:
            g.xpbd_step(0.01, 4, 1, sync=sync)
    g = api.Group.from_scene(ctx, sc)
    g.finalize(api.SPGX_VBD if solver == "vbd" else api.SPGX_XPBD)
    step(g, True)
    x1, v1 = g.download()
    # blocking: one frame from (x1, v1)
    g.upload(x1, v1)
    step(g, True)
    xs, vs = g.download()
    # pipelined: same inputs from pinned memory every frame, outputs to alternating pinned buffers
    xin, vin = api.pinned_empty(x1.shape), api.pinned_empty(v1.shape)
    xin[:], vin[:] = x1, v1
    xo = [api.pinned_empty(x1.shape) for _ in range(2)]
    vo = [api.pinned_empty(x1.shape) for _ in range(2)]
    for k in range(5):
        g.upload_async(xin, vin)
        step(g, False)
        g.download_async(xo[k & 1], vo[k & 1])
    ctx.sync()
    for k in range(2):
        assert np.array_equal(xo[k], xs) and np.array_equal(vo[k], vs)
    g.transfer_wait()
    g.close()
