"""Pins oracle/spg_port.cpp (the CPU restatement, TEST INFRASTRUCTURE): against the golden vectors generated from the
unmodified reference (tests/golden/golden_v1.npz: element known-answers, colourings, whole-step trajectories), and --
where oracle/_ref/libspgref.so exists -- against the compiled reference itself on larger and randomised inputs."""
import sys
from pathlib import Path

import numpy as np
import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent / "golden"))
import make_golden as MG  # noqa: E402
import oracle_api  # noqa: E402
from spg_b200 import scenes as S  # noqa: E402

G = np.load(Path(__file__).resolve().parent / "golden" / "golden_v1.npz")
RTOL = 1e-9  # the gate north_star states; the port sits at ~1e-13 (summation order differs from Eigen / TinyAD)


def _rel(a, b):
    return np.abs(np.asarray(a) - b).max() / max(np.abs(b).max(), 1e-30)


@pytest.fixture(scope="module")
def port():
    if not oracle_api.available("port"):
        pytest.skip("oracle/libspgport.so not built (make -C oracle port)")
    return oracle_api.load("port")


def test_port_elements_match_golden(port):
    assert port.kind == "port"
    for i, (kind, x, x0, st, p) in enumerate(MG.fixtures()):
        og = oracle_api.OracleGroup(port, MG.fixture_scene(kind, x, x0, st, p))
        for local in range(len(st)):
            f, h = og.vertex_force_hessian(0, 0, local)
            assert _rel(f, G[f"fix{i}_k{kind}_f{local}"]) <= RTOL, (kind, local)
            assert _rel(h, G[f"fix{i}_k{kind}_h{local}"]) <= RTOL, (kind, local)
        if kind == S.K_STATIC_LOAD:  # energy-only kind: the constraint form throws, as in the reference
            with pytest.raises(oracle_api.OracleError, match="Constraint-based load energy not implemented"):
                og.constraints(0, 0)
            continue
        c, gc = og.constraints(0, 0)
        ref_c = G[f"fix{i}_k{kind}_c"]
        assert np.abs(c - ref_c).max() <= RTOL * max(np.abs(ref_c).max(), 1.0), kind
        assert _rel(gc, G[f"fix{i}_k{kind}_gc"]) <= RTOL, kind


@pytest.mark.parametrize("name", sorted(MG.COLOR_CASES))
def test_port_colouring_matches_golden(port, name):
    sc = MG.COLOR_CASES[name]()
    sets = [e.stencils for e in sc.energies]
    assert np.array_equal(port.color_vertices(sc.n_vertices, sets), G[f"vcol_{name}"])
    for e, s in enumerate(sets):
        assert np.array_equal(port.color_stencils(s), G[f"scol_{name}_{e}"])


@pytest.mark.parametrize("name", sorted(MG.STEP_CASES))
def test_port_steps_match_golden(port, name):
    build, solver, nsub, iters, frames = MG.STEP_CASES[name]
    og = oracle_api.OracleGroup(port, build())
    osol = oracle_api.OracleSolver(og, solver, 0.01, nsub, iters)
    osol.step(frames)
    x, v = osol.state()
    assert _rel(x, G[f"step_{name}_x"]) <= RTOL * (10 if frames > 1 else 1)
    assert _rel(v, G[f"step_{name}_v"]) <= 1e-7


def test_port_errors_follow_the_reference(port):
    sc = MG.fixture_scene(S.K_SPRING, np.zeros((2, 3)), np.zeros((2, 3)), [0, 1], [0.0, 1.0])
    with pytest.raises(oracle_api.OracleError, match="Zero/negative rest length invalid"):
        oracle_api.OracleGroup(port, sc)
    x0 = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0.5]], float)
    sc = MG.fixture_scene(S.K_MEMBRANE_STVK, x0, x0, [0, 1, 2], [1.0, 0.3])
    with pytest.raises(oracle_api.OracleError, match="must have coordinate z=0"):
        oracle_api.OracleGroup(port, sc)
    t0 = np.array([[0, 0, 0], [0, 1, 0], [1, 0, 0], [0, 0, 1]], float)  # inverted tet
    sc = MG.fixture_scene(S.K_TET_SNH, t0, t0, [0, 1, 2, 3], [10.0, 0.3])
    with pytest.raises(oracle_api.OracleError, match="rest volume is not positive"):
        oracle_api.OracleGroup(port, sc)


# ---- against the compiled reference itself (build container; skipped where oracle/_ref cannot exist) ----
def test_port_vs_reference_random_elements(port, ref_oracle):
    rng = np.random.default_rng(7)
    for kind, x, x0, st, p in MG.fixtures():
        for _ in range(5):
            xr = np.asarray(x, float) + 0.05 * rng.standard_normal(np.shape(x))
            sc = MG.fixture_scene(kind, xr, x0, st, p)
            a, b = oracle_api.OracleGroup(port, sc), oracle_api.OracleGroup(ref_oracle, sc)
            assert abs(a.energy(0, 0) - b.energy(0, 0)) <= RTOL * max(abs(b.energy(0, 0)), 1e-12)
            ga, ha = a.grad_hess(0, 0)
            gb, hb = b.grad_hess(0, 0)
            assert _rel(ga, gb) <= RTOL and _rel(ha, hb) <= RTOL, kind
            ka, aa = a.stiffness(0, 0)
            kb, ab = b.stiffness(0, 0)
            assert _rel(ka, kb) <= 1e-12 and _rel(aa, ab) <= 1e-12, kind
            assert np.allclose(a.rest_data(0, 0), b.rest_data(0, 0), rtol=1e-12, atol=1e-15), kind
            if kind == S.K_STATIC_LOAD:
                continue
            a.project(0, 0, 0.01)
            b.project(0, 0, 0.01)
            assert _rel(a.state()[0], b.state()[0]) <= RTOL, kind


@pytest.mark.parametrize("case", ["vbd_cloth", "xpbd_cloth", "vbd_beam", "xpbd_serial_beam"])
def test_port_vs_reference_steps(port, ref_oracle, case):
    build, solver, nsub, iters = {
        "vbd_cloth": (lambda: S.cloth(24, 17), 0, 1, 10),
        "xpbd_cloth": (lambda: S.cloth(20, 20, S.K_SPRING, S.K_QUADRATIC_BENDING), 1, 20, 1),
        "vbd_beam": (lambda: S.tet_beam(7, 3, 4, S.K_TET_SNH), 0, 1, 20),
        "xpbd_serial_beam": (lambda: S.tet_beam(6, 3, 3, S.K_TET_STVK), 2, 10, 1),
    }[case]
    sc = build()
    res = []
    for o in (port, ref_oracle):
        osol = oracle_api.OracleSolver(oracle_api.OracleGroup(o, sc), solver, 0.01, nsub, iters)
        if solver == 0:
            colors, n = osol.vbd_colors()
            rowptr, entries = osol.vbd_incidence()
            res.append((colors, n, rowptr, entries))
        osol.step(1)
        res.append(osol.state())
    if solver == 0:
        (ca, na, ra, ea), sa, (cb, nb, rb, eb), sb = res
        assert na == nb and np.array_equal(ca, cb) and np.array_equal(ra, rb) and np.array_equal(ea, eb)  # bit-exact
    else:
        sa, sb = res
    assert _rel(sa[0], sb[0]) <= RTOL and _rel(sa[1], sb[1]) <= 1e-7
