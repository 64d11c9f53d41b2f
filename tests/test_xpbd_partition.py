"""The XPBD domain-decomposition plan (spg_b200/csrc/xpbd_partition.h, spgx_xpbd_partition_plan; SURVEY.md section 8e "next")
checked on the CPU by SIMULATING the ranks with the oracle's own stencil projection (ref_project_position =
StencilBlockEnergy::projectPosition, energy.h:117-153): every simulated rank keeps a private copy of the positions, runs
only the stencils the plan says it holds, pass by pass, and after each pass copies exactly the vertices the plan's send /
receive lists name from their owners.  If the plan is right, every rank's owned vertices end the frame bit-identical to
the single-process XPBD::step (xpbd.cpp:16-131); a plan that misses one halo update does not."""
import numpy as np
import pytest

import oracle_api
from spg_b200 import api, scenes as S

DT, SUBSTEPS, GRAVITY = 0.01, 4, np.array([0.0, -9.8, 0.0])


def _plans(sc, nranks, owner):
    sets = [e.stencils for e in sc.energies]
    return [api.xpbd_partition_plan(sc.n_vertices, sets, r, nranks, owner) for r in range(nranks)]


def _owner_of(sc, nranks, owner):
    if owner is not None:
        return np.asarray(owner, np.int32)
    return (np.arange(sc.n_vertices, dtype=np.int64) * nranks // sc.n_vertices).astype(np.int32)


def _colours(sc):
    return [api.color_stencils(e.stencils) for e in sc.energies]


def _simulate(oracle, sc, nranks, owner, frames, drop=None):
    """drop = (rank, pass): that rank ignores what it should receive in that pass (to show the exchange is needed)"""
    own = _owner_of(sc, nranks, owner)
    plans = _plans(sc, nranks, owner)
    groups = [oracle_api.OracleGroup(oracle, sc) for _ in range(nranks)]
    dt = DT / SUBSTEPS
    xs = [sc.x.copy() for _ in range(nranks)]
    vs = [np.zeros_like(sc.x) for _ in range(nranks)]
    pass_energy = []
    for e, col in enumerate(_colours(sc)):
        pass_energy += [e] * (int(col.max()) + 1 if col.size else 0)
    assert all(p["n_passes"] == len(pass_energy) for p in plans)
    for _ in range(frames * SUBSTEPS):
        x_old = [x.copy() for x in xs]
        for r in range(nranks):  # predict on every local copy (particleGroup.cpp:57-63)
            vs[r] += dt * GRAVITY
            xs[r] += vs[r] * dt
        for p, e in enumerate(pass_energy):
            for r in range(nranks):
                pl = plans[r]
                groups[r].set_state(xs[r], vs[r])
                for st in pl["held"][pl["held_offsets"][p]:pl["held_offsets"][p + 1]]:
                    groups[r].project(e, int(st), dt)
                xs[r] = groups[r].state()[0]
            for r in range(nranks):  # the exchange of pass p, exactly as the lists say
                pl = plans[r]
                npeers = len(pl["peers"])
                for i, q in enumerate(pl["peers"]):
                    cell = p * npeers + i
                    want = pl["recv_verts"][pl["recv_offsets"][cell]:pl["recv_offsets"][cell + 1]]
                    pq = plans[q]
                    j = list(pq["peers"]).index(r)
                    cq = p * len(pq["peers"]) + j
                    sent = pq["send_verts"][pq["send_offsets"][cq]:pq["send_offsets"][cq + 1]]
                    assert np.array_equal(want, sent)          # both sides agree on the message
                    assert np.all(own[sent] == q)              # only owners send
                    if drop == (r, p):
                        continue
                    xs[r][want] = xs[q][sent]
        for r in range(nranks):  # particleGroup.cpp:65-72
            vs[r] = (xs[r] - x_old[r]) * (1.0 / dt)
    x = np.empty_like(sc.x)
    v = np.empty_like(sc.x)
    for r in range(nranks):
        x[own == r] = xs[r][own == r]
        v[own == r] = vs[r][own == r]
    return x, v, plans, (xs, own)


def _reference(oracle, sc, frames):
    og = oracle_api.OracleGroup(oracle, sc)
    sol = oracle_api.OracleSolver(og, 1, DT, SUBSTEPS, 1, tuple(GRAVITY))
    sol.step(frames)
    return sol.state()


SCENES = {
    "cloth_springs": lambda: S.cloth(9, 9, S.K_SPRING, S.K_QUADRATIC_BENDING),       # BASELINE config 2's energies
    "cloth_stvk": lambda: S.cloth(8, 8, S.K_MEMBRANE_STVK, S.K_DISCRETE_BENDING),
    "beam": lambda: S.tet_beam(5, 3, 3),
}


@pytest.mark.parametrize("name", sorted(SCENES))
@pytest.mark.parametrize("nranks", [2, 3])
def test_partitioned_xpbd_equals_single_process(oracle, name, nranks):
    sc = SCENES[name]()
    x, v, plans, (xs, own) = _simulate(oracle, sc, nranks, None, frames=2)
    xr, vr = _reference(oracle, sc, 2)
    assert np.array_equal(x, xr) and np.array_equal(v, vr)
    # halo copies are current too (the next frame starts from them)
    for r, pl in enumerate(plans):
        assert np.array_equal(xs[r][pl["halo"]], xr[pl["halo"]])
        assert pl["n_owned"] == int((own == r).sum())
    # the exchange is really used, and much smaller than the vertex set
    sent = sum(p["send_verts"].size for p in plans)
    assert 0 < sent
    per_pass_boundary = max(np.diff(p["send_offsets"]).max() for p in plans)
    assert per_pass_boundary <= sc.n_vertices // 2


def test_single_rank_plan_is_trivial(oracle):
    sc = SCENES["cloth_springs"]()
    pl = api.xpbd_partition_plan(sc.n_vertices, [e.stencils for e in sc.energies], 0, 1)
    assert pl["n_owned"] == sc.n_vertices and pl["halo"].size == 0 and pl["peers"].size == 0
    assert pl["send_verts"].size == 0 and pl["recv_verts"].size == 0
    assert pl["held"].size == sum(e.stencils.shape[0] for e in sc.energies)
    x, v, _, _ = _simulate(oracle, sc, 1, None, frames=1)
    xr, vr = _reference(oracle, sc, 1)
    assert np.array_equal(x, xr) and np.array_equal(v, vr)   # the simulation itself is the reference's step


def test_irregular_ownership_and_dropped_message(oracle):
    sc = SCENES["cloth_stvk"]()
    rng = np.random.default_rng(5)
    owner = rng.integers(0, 3, sc.n_vertices).astype(np.int32)   # a partition with no locality at all
    x, v, plans, _ = _simulate(oracle, sc, 3, owner, frames=1)
    xr, vr = _reference(oracle, sc, 1)
    assert np.array_equal(x, xr) and np.array_equal(v, vr)
    # losing one non-empty message changes the result: the lists are necessary, not just sufficient
    pl = plans[0]
    cells = np.flatnonzero(np.diff(pl["recv_offsets"]) > 0)
    first_pass = int(cells[0] // len(pl["peers"]))
    x2, _, _, _ = _simulate(oracle, sc, 3, owner, frames=1, drop=(0, first_pass))
    assert not np.array_equal(x2, xr)


def test_bad_arguments():
    sc = SCENES["cloth_springs"]()
    sets = [e.stencils for e in sc.energies]
    with pytest.raises(api.SpgxError):
        api.xpbd_partition_plan(sc.n_vertices, sets, 2, 2)
    with pytest.raises(api.SpgxError):
        api.xpbd_partition_plan(sc.n_vertices, sets, 0, 65)
    with pytest.raises(api.SpgxError):
        api.xpbd_partition_plan(sc.n_vertices, sets, 0, 2, owner=np.full(sc.n_vertices, 7, np.int32))
