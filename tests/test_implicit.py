"""The implicit-Euler family and symplectic Euler (reference:
spg/source/sim/solver/implicitEulerBase.cpp, implicitEulerBaraffWitkin.cpp, implicitEulerNewtonDx.cpp,
implicitEulerNewtonDv.cpp, simplecticEuler.cpp).

CPU part (-m "not gpu"): the oracle that travels (oracle/libspgport.so, or the compiled reference where present)
reproduces the golden vectors generated from the unmodified reference (tests/golden/make_golden_implicit.py), and the
port equals the reference on the system force, the stiffness matrix, the CG iteration counts and whole steps.
GPU part (-m gpu): the CUDA path through the C ABI against the same golden vectors and against the oracle:
  * element Hessians, system force, block-sparse stiffness matrix: <= 1e-9 relative (observed ~1e-13);
  * the linear solve: same CG iteration count as the reference's Eigen::ConjugateGradient on the same system, residual
    under the 1e-5 tolerance, solution within 1e-7 of the oracle's;
  * whole steps: CG iteration counts equal to the reference's, positions / velocities within 1e-7 relative.  The step
    gate is looser than the 1e-9 of VBD / XPBD on purpose: the reference stops its CG at |r| <= 1e-5 |b|
    (implicitEulerBase.h:22), so its own answer carries an error of that order and rounding differences inside the
    Krylov recurrences are amplified by the condition number of M - dt^2 K (anchors 1e5 vs bending 1e-2).
"""
import sys
from pathlib import Path

import numpy as np
import pytest
import scipy.sparse as sp

sys.path.insert(0, str(Path(__file__).resolve().parent / "golden"))
import make_golden as MG  # noqa: E402
import make_golden_implicit as MI  # noqa: E402
import oracle_api  # noqa: E402
from spg_b200 import scenes as S  # noqa: E402

G = np.load(Path(__file__).resolve().parent / "golden" / "golden_v2_implicit.npz")
RTOL = 1e-9
STEP_RTOL = 1e-7


def _rel(a, b):
    return np.abs(np.asarray(a) - b).max() / max(np.abs(b).max(), 1e-30)


def _golden_K(name):
    n = 3 * G[f"sys_{name}_x"].shape[0]
    return sp.csr_matrix((G[f"sys_{name}_Kval"], (G[f"sys_{name}_Krow"], G[f"sys_{name}_Kcol"])), shape=(n, n))


# ---------------------------------------------------------------------------------------------------------------------
# CPU: the oracle is pinned
# ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", sorted(MI.SCENES))
def test_oracle_system_matches_golden(oracle, name):
    og = oracle_api.OracleGroup(oracle, MI.SCENES[name]())
    sol = oracle_api.OracleSolver(og, MI.NDX, 0.01, 1, 1)
    sol.set_state(G[f"sys_{name}_x"], G[f"sys_{name}_v"])
    assert _rel(sol.system_force(), G[f"sys_{name}_f"]) <= 1e-11
    K, Kg = sol.system_matrix(0), _golden_K(name)
    assert abs(K - Kg).max() <= 1e-11 * abs(Kg).max()
    assert abs(K - K.T).max() <= 1e-11 * abs(Kg).max()


@pytest.mark.parametrize("name", sorted(MI.STEP_CASES))
def test_oracle_steps_match_golden(oracle, name):
    scn, kind, nsub, frames = MI.STEP_CASES[name]
    og = oracle_api.OracleGroup(oracle, MI.SCENES[scn]())
    sol = oracle_api.OracleSolver(og, kind, 0.01, nsub, 1)
    sol.step(frames)
    x, v = sol.state()
    if kind != MI.SYMPLECTIC:
        # the compiled reference reproduces its own counts; the independent restatement (different summation order in the
        # products) may cross the |r| <= 1e-5 |b| threshold one iteration earlier or later
        slack = 0 if oracle.kind == "reference" else 1
        assert np.abs(sol.cg_iterations() - G[f"step_{name}_cg"]).max() <= slack
    # symplectic Euler at 200-400 substeps accumulates its omp-atomic summation order: 1e-9 still holds
    assert _rel(x, G[f"step_{name}_x"]) <= RTOL and _rel(v, G[f"step_{name}_v"]) <= 1e-8


def test_port_equals_reference_implicit(ref_oracle):
    """the restatement (accumulateForces, negativeHessianTriplets, Eigen's CG) against the reference itself"""
    port = oracle_api.load("port")
    if port.kind != "port":
        pytest.skip("port oracle not built")
    rng = np.random.default_rng(5)
    for sc in (S.cloth(8, 11, S.K_MEMBRANE_BW, S.K_DISCRETE_BENDING), S.tet_beam(4, 3, 3, S.K_TET_SSNH)):
        x = sc.x + 1e-3 * rng.standard_normal(sc.x.shape)
        v = 0.1 * rng.standard_normal(sc.x.shape)
        out = []
        for o in (ref_oracle, port):
            og = oracle_api.OracleGroup(o, sc)
            sol = oracle_api.OracleSolver(og, MI.BW, 0.01, 2, 1)
            sol.set_state(x, v)
            f, K = sol.system_force(), sol.system_matrix(0)
            b = 0.01 * f.reshape(-1)
            A = sp.diags(np.repeat(sc.mass, 3)) - 1e-4 * K  # M - dt^2 K, dt f: the system the Newton dv solver builds
            xs, it = sol.linear_solve(A, b)
            sol.step(2)
            out.append((f, K, xs, it, b, sol.state(), sol.cg_iterations()))
        (f0, K0, x0, it0, _, s0, cg0), (f1, K1, x1, it1, _, s1, cg1) = out
        assert _rel(f1, f0) <= 1e-12 and abs(K1 - K0).max() <= 1e-12 * abs(K0).max()
        # on a noisy state, rounding differences between two correct CG implementations grow with the iteration count
        # (loss of orthogonality) up to the order of the stopping tolerance 1e-5; observed 2e-9 .. 5e-6
        assert abs(it0 - it1) <= 1 and _rel(x1, x0) <= 1e-4, (it0, it1, _rel(x1, x0))
        assert np.abs(cg0 - cg1).max() <= 1
        assert _rel(s1[0], s0[0]) <= 1e-6 and _rel(s1[1], s0[1]) <= 1e-4


# ---------------------------------------------------------------------------------------------------------------------
# GPU: the CUDA path through the C ABI
# ---------------------------------------------------------------------------------------------------------------------
def _bsr_to_csr(rowptr, cols, blocks, n):
    return sp.bsr_matrix((blocks, cols, rowptr), shape=(3 * n, 3 * n)).tocsr()


@pytest.mark.gpu
def test_gpu_element_hessians_match_golden(ctx):
    from spg_b200 import api
    for i, (kind, x, x0, st, p) in enumerate(MG.fixtures()):
        g = api.Group.from_scene(ctx, MG.fixture_scene(kind, x, x0, st, p))
        g.finalize(api.SPGX_NEWTON)
        h = g.element_hessian(0, 0)
        assert _rel(h, G[f"fix{i}_k{kind}_hess"]) <= RTOL, kind
        f = g.system_force((0.0, 0.0, 0.0))
        assert _rel(-f.reshape(-1), G[f"fix{i}_k{kind}_grad"]) <= RTOL, kind
        g.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(MI.SCENES))
def test_gpu_system_matches_golden(ctx, name):
    from spg_b200 import api
    sc = MI.SCENES[name]()
    g = api.Group.from_scene(ctx, sc)
    g.finalize(api.SPGX_NEWTON)
    g.upload(G[f"sys_{name}_x"], G[f"sys_{name}_v"])
    assert _rel(g.system_force(), G[f"sys_{name}_f"]) <= RTOL
    rowptr, cols, blocks = g.system_stiffness()
    K, Kg = _bsr_to_csr(rowptr, cols, blocks, sc.n_vertices), _golden_K(name)
    assert abs(K - Kg).max() <= RTOL * abs(Kg).max()
    # reproducible bit for bit: no atomics anywhere in the assembly
    _, _, blocks2 = g.system_stiffness()
    assert np.array_equal(blocks, blocks2)
    # the pattern holds exactly the pairs of vertices that share a stencil
    nb = set()
    for e in sc.energies:
        for st in e.stencils:
            nb.update((int(a), int(b)) for a in st for b in st)
    nb.update((v, v) for v in range(sc.n_vertices))
    rows = np.repeat(np.arange(sc.n_vertices), np.diff(rowptr))
    assert set(zip(rows.tolist(), cols.tolist())) == nb
    g.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["cloth12", "beam_snh"])
def test_gpu_linear_solve_vs_oracle(ctx, oracle, name):
    """solveLinearSystem on the same system: iteration count, residual, solution"""
    from spg_b200 import api
    sc = MI.SCENES[name]()
    g = api.Group.from_scene(ctx, sc)
    g.finalize(api.SPGX_NEWTON)
    g.upload(G[f"sys_{name}_x"], G[f"sys_{name}_v"])
    rowptr, cols, blocks = g.system_stiffness()
    K = _bsr_to_csr(rowptr, cols, blocks, sc.n_vertices)
    M = sp.diags(np.repeat(sc.mass, 3))
    dt = 0.01
    b = dt * g.system_force().reshape(-1)
    og = oracle_api.OracleGroup(oracle, sc)
    osol = oracle_api.OracleSolver(og, MI.NDV, dt, 1, 1)
    for cM, cH in ((1.0, dt * dt), (1.0 / (dt * dt), 1.0)):
        A = (cM * M - cH * K).tocsr()  # K = -Hessian
        x, it = g.linear_solve(cM, cH, b)
        xr, itr = osol.linear_solve(A, b)
        assert it == itr, (it, itr)
        assert np.linalg.norm(A @ x.reshape(-1) - b) <= 1e-5 * np.linalg.norm(b) * (1 + 1e-6)
        assert _rel(x.reshape(-1), xr) <= STEP_RTOL
    # a zero right-hand side returns zero without iterating (ConjugateGradient.h:44-49)
    x, it = g.linear_solve(1.0, dt * dt, np.zeros_like(b))
    assert it == 0 and not x.any()
    # an iteration cap below convergence stops there (cg.setMaxIterations)
    x, it = g.linear_solve(1.0, dt * dt, b, max_iterations=3)
    assert it == 3
    g.close()


_GPU_KIND = {MI.BW: 0, MI.NDX: 1, MI.NDV: 2, MI.BDF2: 3, MI.STATIC: 4, MI.QUASI: 5, MI.NROBUST: 6, MI.QROBUST: 7}


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(MI.STEP_CASES))
def test_gpu_steps_match_golden(ctx, name):
    from spg_b200 import api
    scn, kind, nsub, frames = MI.STEP_CASES[name]
    g = api.Group.from_scene(ctx, MI.SCENES[scn]())
    g.finalize(api.SPGX_NEWTON)
    if kind == MI.SYMPLECTIC:
        g.symplectic_step(0.01, nsub, frames=frames)
        tol = RTOL
    else:
        g.implicit_step(_GPU_KIND[kind], 0.01, nsub, frames=frames)
        assert np.array_equal(g.cg_iterations(), G[f"step_{name}_cg"]), (g.cg_iterations(), G[f"step_{name}_cg"])
        tol = STEP_RTOL
    x, v = g.download()
    assert _rel(x, G[f"step_{name}_x"]) <= tol, _rel(x, G[f"step_{name}_x"])
    assert _rel(v, G[f"step_{name}_v"]) <= 10 * tol, _rel(v, G[f"step_{name}_v"])
    g.close()


@pytest.mark.gpu
def test_gpu_implicit_vs_oracle_cloth32(ctx, oracle):
    """a larger scene than the golden ones, every member of the family, 3 frames each"""
    from spg_b200 import api
    sc = S.cloth(32, 32)
    for kind in (MI.BW, MI.NDX, MI.NDV):
        g = api.Group.from_scene(ctx, sc)
        g.finalize(api.SPGX_NEWTON)
        og = oracle_api.OracleGroup(oracle, sc)
        osol = oracle_api.OracleSolver(og, kind, 0.01, 1, 1)
        g.implicit_step(_GPU_KIND[kind], 0.01, 1, frames=3)
        osol.step(3)
        x, v = g.download()
        xr, vr = osol.state()
        assert np.abs(g.cg_iterations() - osol.cg_iterations()).max() <= 1, (g.cg_iterations(), osol.cg_iterations())
        assert _rel(x, xr) <= 1e-6 and _rel(v, vr) <= 1e-4, (kind, _rel(x, xr), _rel(v, vr))
        g.close()


@pytest.mark.gpu
def test_gpu_newton_state_errors(ctx):
    from spg_b200 import api
    g = api.Group.from_scene(ctx, S.cloth(5, 5))
    g.finalize(api.SPGX_VBD)
    with pytest.raises(api.SpgxError):
        g.implicit_step(0, 0.01)
    g.finalize(api.SPGX_NEWTON)
    with pytest.raises(api.SpgxError):
        g.vbd_step(0.01, 1, 1)
    with pytest.raises(api.SpgxError):
        g.xpbd_step(0.01, 1, 1)
    with pytest.raises(api.SpgxError):
        g.linear_solve(1.0, 1.0, np.zeros(75))  # nothing assembled yet
    with pytest.raises(api.SpgxError):
        g.implicit_step(99, 0.01)
    g.implicit_step(1, 0.01)
    x, _ = g.download()
    assert np.isfinite(x).all()
    g.close()


@pytest.mark.gpu
def test_gpu_bdf2_reset_forgets_previous_state(ctx):
    """BDF2::reset (bdf2.cpp:78-83) drops x-1 / v-1: a run after reset repeats the first run bit for bit"""
    from spg_b200 import api
    g = api.Group.from_scene(ctx, S.cloth(9, 9))
    g.finalize(api.SPGX_NEWTON)
    g.implicit_step(api.SPGX_IE_BDF2, 0.01, 2, frames=2)
    first = g.download()
    g.reset()
    g.implicit_step(api.SPGX_IE_BDF2, 0.01, 2, frames=2)
    second = g.download()
    assert np.array_equal(first[0], second[0]) and np.array_equal(first[1], second[1])
    g.close()


@pytest.mark.gpu
@pytest.mark.parametrize("scene", ["cloth", "beam"])
@pytest.mark.parametrize("kind", [MI.NROBUST, MI.QROBUST])
def test_gpu_robust_newton_iterations_vs_oracle(ctx, oracle, scene, kind):
    """a state that needs several Newton iterations per substep (random velocities, dt = 0.03): same number of linear
    solves as the reference (i.e. the same Newton / line-search / roll-back decisions), CG counts within 2 -- within 3 % for
    the long solves: the reference scatters its forces under `omp atomic` (energy.h:185-212), so on a noisy state its own
    count of a 140-iteration solve moves by a few iterations with the host's thread count and from run to run (observed on the
    pool: 141 against the device's 138 on one box, equal counts on the others)"""
    from spg_b200 import api
    sc = S.cloth(12, 12) if scene == "cloth" else S.tet_beam(5, 3, 3, S.K_TET_SNH)
    v0 = 0.5 * np.random.default_rng(3).standard_normal(sc.x.shape)
    g = api.Group.from_scene(ctx, sc)
    g.finalize(api.SPGX_NEWTON)
    g.upload(sc.x, v0)
    og = oracle_api.OracleGroup(oracle, sc)
    osol = oracle_api.OracleSolver(og, kind, 0.03, 1, 1)
    osol.set_state(sc.x, v0)
    g.implicit_step(_GPU_KIND[kind], 0.03, 1, frames=5)
    osol.step(5)
    cg, cgr = g.cg_iterations(), osol.cg_iterations()
    assert cg.size == cgr.size and cg.size > 5, (cg, cgr)
    assert (np.abs(cg - cgr) <= np.maximum(2, np.ceil(0.03 * cgr))).all(), (cg, cgr)
    x, v = g.download()
    xr, vr = osol.state()
    assert _rel(x, xr) <= 1e-6 and _rel(v, vr) <= 1e-4, (_rel(x, xr), _rel(v, vr))
    g.close()



@pytest.mark.gpu
def test_gpu_hand_derived_membrane_hessian_equals_the_generic_one(ctx, oracle, monkeypatch):
    """membraneStvkHessianKernel (SPGX_MEMBRANE_HAND=1: the 9 x 9 StVK membrane Hessian derived by hand, staged stores)
    against the generic mixed-dual path and the oracle: element Hessians of a deformed cloth, the assembled matrix, one
    Newton dx step with the same CG iteration count"""
    from spg_b200 import api
    sc = S.cloth(13, 9)
    rng = np.random.default_rng(4)
    x = sc.x + 2e-2 * rng.standard_normal(sc.x.shape)
    out = {}
    for hand in ("0", "1"):
        monkeypatch.setenv("SPGX_MEMBRANE_HAND", hand)
        g = api.Group.from_scene(ctx, sc)
        g.finalize(api.SPGX_NEWTON)
        g.upload(x, np.zeros_like(x))
        eid = [i for i, e in enumerate(sc.energies) if e.kind == S.K_MEMBRANE_STVK][0]
        hs = np.stack([g.element_hessian(eid, st) for st in range(0, sc.energies[eid].stencils.shape[0], 7)])
        _, _, blocks = g.system_stiffness()
        g.implicit_step(api.SPGX_IE_NEWTON_DX, 0.01, 1)
        out[hand] = (hs, blocks, g.download()[0], g.cg_iterations())
        g.close()
    og = oracle_api.OracleGroup(oracle, sc)
    og.set_state(x, np.zeros_like(x))
    for k, st in enumerate(range(0, sc.energies[eid].stencils.shape[0], 7)):
        _, href = og.grad_hess(eid, st)
        assert _rel(out["1"][0][k], href) <= 1e-11
    assert _rel(out["1"][0], out["0"][0]) <= 1e-12 and _rel(out["1"][1], out["0"][1]) <= 1e-12
    # the two matrices differ in the last bits, and on a noisy state the CG (stopped at |r| <= 1e-5 |b|) may cross its threshold
    # an iteration or two apart (observed 123 vs 121): this is why the hand-derived kernel stays opt-in -- the golden step cases
    # hold the DEFAULT path to the reference's exact iteration counts
    # (the two solutions are two CG runs stopped at 1e-5 |b| on an ill-conditioned system: they agree to that order, observed 6e-5)
    assert np.abs(out["1"][3] - out["0"][3]).max() <= 3 and _rel(out["1"][2], out["0"][2]) <= 1e-3


@pytest.mark.gpu
def test_gpu_newton_dx_consecutive_frames_follow_the_reference(ctx, ref_oracle):
    """bench config 6 restarts every frame because the reference's Newton dx formulation (no PSD projection, CG capped at 300
    iterations) does not survive consecutive frames of a fine cloth from rest.  Pinned here at a size the reference runs in
    seconds: frame by frame the device path does what the reference does -- same CG iteration counts while the solves converge,
    the cap hit at the same frame, the states together within the CG tolerance while they are finite, and non-finite (if at
    all) at the same frame -- so the restart is a property of the formulation, not of the device path."""
    from spg_b200 import api
    sc = S.cloth(96, 96)
    og = oracle_api.OracleGroup(ref_oracle, sc)
    osol = oracle_api.OracleSolver(og, MI.NDX, 0.01, 1, 1)
    g = api.Group.from_scene(ctx, sc)
    g.finalize(api.SPGX_NEWTON)
    ref_cg, dev_cg, ref_finite, dev_finite, err = [], [], [], [], []
    for frame in range(6):
        osol.step(1)
        g.implicit_step(api.SPGX_IE_NEWTON_DX, 0.01, 1)
        xr, _ = osol.state()
        x, _ = g.download()
        ref_cg.append(int(osol.cg_iterations()[-1]))
        dev_cg.append(int(g.cg_iterations()[-1]))
        ref_finite.append(bool(np.isfinite(xr).all()))
        dev_finite.append(bool(np.isfinite(x).all()))
        err.append(_rel(x, xr) if ref_finite[-1] and dev_finite[-1] else np.nan)
    report = f"cg ref {ref_cg} dev {dev_cg}; finite ref {ref_finite} dev {dev_finite}; rel err {err}"
    print(report)
    assert ref_finite == dev_finite, report
    assert [c >= 300 for c in ref_cg] == [c >= 300 for c in dev_cg], report   # the iteration cap is hit at the same frames
    first_capped = next((i for i, c in enumerate(ref_cg) if c >= 300), len(ref_cg))
    for i in range(first_capped):                                              # converged solves: the reference's counts
        assert abs(ref_cg[i] - dev_cg[i]) <= 2 and err[i] <= 1e-5, report
