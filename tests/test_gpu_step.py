"""Step parity on the GPU (SURVEY.md section 4 items 3-4, section 8d gates): colours + incidence bit-exact, positions and
velocities within 1e-9 relative after one substep, bounded drift over N frames -- CUDA path through the C ABI vs
the reference's VBD::step / XPBD::step (vbd.cpp:12-195, xpbd.cpp:16-131)."""
import numpy as np
import pytest

import oracle_api
from spg_b200 import api, scenes as S

pytestmark = pytest.mark.gpu
RTOL = 1e-9


def _rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)


def _vbd_pair(ctx, oracle, sc, substeps, iterations, dt=0.01):
    g = api.Group.from_scene(ctx, sc)
    g.finalize(api.SPGX_VBD)
    og = oracle_api.OracleGroup(oracle, sc)
    osol = oracle_api.OracleSolver(og, 0, dt, substeps, iterations)
    return g, og, osol


@pytest.mark.parametrize("n", [4, 7, 16, 64])
def test_vbd_structure_bit_exact(ctx, oracle, n):
    sc = S.cloth(n, n)
    g, og, osol = _vbd_pair(ctx, oracle, sc, 1, 1)
    colors, ncol = osol.vbd_colors()
    assert np.array_equal(g.vertex_colors(), colors)
    assert g.num_colors() == ncol
    rp, en = g.vertex_incidence()
    rpr, enr = osol.vbd_incidence()
    assert np.array_equal(rp, rpr) and np.array_equal(en, enr)


@pytest.mark.parametrize("membrane,bending", [(S.K_MEMBRANE_STVK, S.K_DISCRETE_BENDING),
                                              (S.K_MEMBRANE_CHOI, S.K_QUADRATIC_BENDING),
                                              (S.K_MEMBRANE_BW, S.K_DISCRETE_BENDING),
                                              (S.K_SPRING, S.K_QUADRATIC_BENDING),
                                              (S.K_SPRING_CONTINUUM, S.K_DISCRETE_BENDING),
                                              (S.K_SPRING_SQUARED_CONTINUUM, S.K_BENDING_BW),
                                              (S.K_MEMBRANE_STVK, S.K_BENDING_BW)])
def test_vbd_cloth_one_substep(ctx, oracle, membrane, bending):
    sc = S.cloth(12, 12, membrane, bending)
    g, og, osol = _vbd_pair(ctx, oracle, sc, 1, 10)
    g.vbd_step(0.01, 1, 10)
    osol.step(1)
    x, v = g.download()
    xr, vr = osol.state()
    assert _rel(x, xr) <= RTOL and _rel(v, vr) <= RTOL


def test_vbd_config1_cloth64(ctx, oracle):
    """BASELINE config 1: 64x64 cloth, StVK membrane + discrete bending, VBD 10 iterations, 1 substep, 3 frames."""
    sc = S.cloth(64, 64)
    g, og, osol = _vbd_pair(ctx, oracle, sc, 1, 10)
    drift = []
    for frame in range(3):
        g.vbd_step(0.01, 1, 10)
        osol.step(1)
        x, v = g.download()
        xr, vr = osol.state()
        drift.append((_rel(x, xr), _rel(v, vr)))
    assert drift[0][0] <= RTOL and drift[0][1] <= RTOL, drift
    assert drift[-1][0] <= 1e-7, drift   # bounded drift over N frames


@pytest.mark.parametrize("kind", [S.K_TET_SNH, S.K_TET_STVK, S.K_TET_SSNH])
def test_vbd_tet_beam(ctx, oracle, kind):
    sc = S.tet_beam(9, 3, 3, kind)
    g, og, osol = _vbd_pair(ctx, oracle, sc, 2, 20)
    g.vbd_step(0.01, 2, 20, frames=2)
    osol.step(2)
    x, v = g.download()
    xr, vr = osol.state()
    assert _rel(x, xr) <= 1e-8 and _rel(v, vr) <= 1e-7


def test_vbd_graph_and_plain_launch_agree(ctx):
    sc = S.cloth(16, 16)
    out = []
    for mode in (True, False):
        g = api.Group.from_scene(ctx, sc)
        g.finalize(api.SPGX_VBD)
        g.set_graph_mode(mode)
        g.vbd_step(0.01, 2, 5, frames=3)
        out.append(g.download())
        t = g.timings()
        assert t.kernel_launches == 3 * 2 * (2 + 5 * g.num_colors())
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1])


@pytest.mark.parametrize("iters", [1, 3])
def test_xpbd_graph_and_plain_launch_agree(ctx, iters, monkeypatch):
    """XPBD frames as a replayed CUDA graph (the velocity update of a substep and the prediction of the next in one launch),
    as plain launches with the phase timer (two launches), and with the fusion switched off: bit-identical state over
    consecutive frames, and the launch count that goes with each form."""
    sc = S.cloth(24, 19, S.K_SPRING, S.K_QUADRATIC_BENDING)
    out = {}
    for name, graph, fuse in (("graph", True, "1"), ("plain", False, "1"), ("graph_unfused", True, "0")):
        monkeypatch.setenv("SPGX_XPBD_FUSE_STATE", fuse)
        g = api.Group.from_scene(ctx, sc)
        g.finalize(api.SPGX_XPBD)
        g.set_graph_mode(graph)
        g.xpbd_step(0.01, 5, iters, frames=3)
        out[name] = g.download()
        passes = sum(g.num_colors(e) for e in range(len(sc.energies)))
        n_state = 5 + 1 if name == "graph" else 10   # predict + 4 fused + finalize, or predict and finalize per substep
        assert g.timings().kernel_launches == 3 * (5 * iters * passes + n_state), name
        g.close()
    monkeypatch.delenv("SPGX_XPBD_FUSE_STATE", raising=False)
    for name in ("plain", "graph_unfused"):
        assert np.array_equal(out["graph"][0], out[name][0]) and np.array_equal(out["graph"][1], out[name][1]), name


def test_vbd_reset_upload_roundtrip(ctx):
    sc = S.cloth(9, 9)
    g = api.Group.from_scene(ctx, sc)
    g.finalize(api.SPGX_VBD)
    x0, v0 = g.download()
    assert np.array_equal(x0, sc.x) and not v0.any()
    g.vbd_step(0.01, 1, 3)
    x1, v1 = g.download()
    assert not np.array_equal(x1, sc.x)
    g.reset()
    xr, vr = g.download()
    assert np.array_equal(xr, sc.x) and not vr.any()
    g.upload(x1, v1)
    x2, v2 = g.download()
    assert np.array_equal(x1, x2) and np.array_equal(v1, v2)


def _xpbd_pair(ctx, oracle, sc, substeps, serial=False):
    g = api.Group.from_scene(ctx, sc)
    g.finalize(api.SPGX_XPBD_SERIAL if serial else api.SPGX_XPBD)
    og = oracle_api.OracleGroup(oracle, sc)
    osol = oracle_api.OracleSolver(og, 2 if serial else 1, 0.01, substeps, 1)
    return g, og, osol


@pytest.mark.parametrize("n", [4, 16, 64])
def test_xpbd_structure_bit_exact(ctx, oracle, n):
    sc = S.cloth(n, n, S.K_SPRING, S.K_QUADRATIC_BENDING)
    g, og, osol = _xpbd_pair(ctx, oracle, sc, 1)
    for e in range(3):
        colors, ncol = osol.xpbd_colors(e)
        assert np.array_equal(g.stencil_colors(e), colors)
        assert g.num_colors(e) == ncol


@pytest.mark.parametrize("membrane,bending", [(S.K_SPRING, S.K_QUADRATIC_BENDING),
                                              (S.K_MEMBRANE_STVK, S.K_DISCRETE_BENDING),
                                              (S.K_MEMBRANE_CHOI, S.K_QUADRATIC_BENDING),
                                              (S.K_MEMBRANE_BW, S.K_DISCRETE_BENDING),
                                              (S.K_SPRING_CONTINUUM, S.K_QUADRATIC_BENDING),
                                              (S.K_SPRING_SQUARED_CONTINUUM, S.K_BENDING_BW)])
def test_xpbd_cloth(ctx, oracle, membrane, bending):
    """config-2 style: 20 small steps per frame, one projection pass each."""
    sc = S.cloth(16, 16, membrane, bending)
    g, og, osol = _xpbd_pair(ctx, oracle, sc, 20)
    g.xpbd_step(0.01, 20, 1)
    osol.step(1)
    x, v = g.download()
    xr, vr = osol.state()
    assert _rel(x, xr) <= RTOL and _rel(v, vr) <= 1e-8


@pytest.mark.parametrize("kind", [S.K_TET_SNH, S.K_TET_STVK, S.K_TET_SSNH])
def test_xpbd_tets(ctx, oracle, kind):
    sc = S.tet_beam(5, 3, 3, kind)
    g, og, osol = _xpbd_pair(ctx, oracle, sc, 10)
    g.xpbd_step(0.01, 10, 1)
    osol.step(1)
    x, v = g.download()
    xr, vr = osol.state()
    assert _rel(x, xr) <= RTOL and _rel(v, vr) <= 1e-8


def test_xpbd_serial_gauss_seidel(ctx, oracle):
    sc = S.cloth(6, 6, S.K_SPRING, S.K_QUADRATIC_BENDING)
    g, og, osol = _xpbd_pair(ctx, oracle, sc, 5, serial=True)
    g.xpbd_step(0.01, 5, 1)
    osol.step(1)
    x, v = g.download()
    xr, vr = osol.state()
    assert _rel(x, xr) <= RTOL


def test_xpbd_lambda_iterations_equal_reference_at_one(ctx):
    """iterations > 1 keeps lambda on device; with a single pass the accumulated-lambda kernel must reproduce the
    small-steps form exactly (lambda starts at 0 each substep)."""
    sc = S.cloth(10, 10, S.K_SPRING, S.K_QUADRATIC_BENDING)
    g = api.Group.from_scene(ctx, sc)
    g.finalize(api.SPGX_XPBD)
    g.xpbd_step(0.01, 4, 1)
    x1, _ = g.download()
    g.reset()
    g.xpbd_step(0.01, 4, 3)
    x3, _ = g.download()
    assert np.isfinite(x3).all() and not np.array_equal(x1, x3)
    assert np.abs(x1 - x3).max() < 1e-2


@pytest.mark.parametrize("scene", ["cloth", "snh", "tstvk"])
def test_vbd_streaming_kernel_variants_agree(ctx, oracle, scene, monkeypatch):
    """The streaming kernel (owner-centric incidence streams, every lanes-per-vertex variant) against the
    generic CSR kernel and the oracle on the same scene: all within the 1e-9 gate of each other."""
    sc = {"cloth": lambda: S.cloth(14, 260), "snh": lambda: S.tet_beam(9, 4, 3, S.K_TET_SNH),
          "tstvk": lambda: S.tet_beam(7, 3, 4, S.K_TET_STVK)}[scene]()
    og = oracle_api.OracleGroup(oracle, sc)
    osol = oracle_api.OracleSolver(og, 0, 0.01, 2, 6)
    osol.step(1)
    xr, vr = osol.state()
    results = {}
    cases = [("csr", {"SPGX_NO_STREAMS": "1"})]
    cases += [(f"g{l}v{var}", {"SPGX_VBD_LANES": str(l), "SPGX_VBD_VARIANT": str(var)}) for l in (2, 4) for var in (4, 14, 15)]
    cases += [("g2v4packed", {"SPGX_VBD_LANES": "2", "SPGX_VBD_VARIANT": "4", "SPGX_VBD_NBRFIXED": "0"})]
    # single-incidence records only (no wedges: streams.cuh ST_WEDGE), register pipeline and ring
    cases += [(f"g{l}v{var}single", {"SPGX_VBD_LANES": str(l), "SPGX_VBD_VARIANT": str(var), "SPGX_VBD_WEDGE": "0"})
              for l, var in ((2, 4), (4, 15))]
    knobs = ("SPGX_NO_STREAMS", "SPGX_VBD_LANES", "SPGX_VBD_VARIANT", "SPGX_VBD_NBRFIXED", "SPGX_VBD_WEDGE")
    for name, env in cases:
        for k in knobs:
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        g = api.Group.from_scene(ctx, sc)
        g.finalize(api.SPGX_VBD)
        g.vbd_step(0.01, 2, 6)
        results[name] = g.download()
        info = g.kernel_info()
        if name != "csr":
            assert info["variant"] == int(name.split("v")[1].replace("packed", "").replace("single", ""))
            # a cloth's membrane and bending incidences pair up into wedges (three incidences a record) unless switched off:
            # every triangle corner whose two hinges exist, i.e. all but the corners next to the mesh boundary
            if scene == "cloth" and "single" not in name:
                n_corners = 3 * sc.energies[1].stencils.shape[0]
                assert 0.8 * n_corners < info["wedge_records"] <= n_corners, info
            else:
                assert info["wedge_records"] == 0, info
        g.close()
    for k in knobs:
        monkeypatch.delenv(k, raising=False)
    for name, (x, v) in results.items():
        assert _rel(x, xr) <= RTOL and _rel(v, vr) <= 1e-8, name
        assert _rel(x, results["csr"][0]) <= 1e-11, name
    for l in (2, 4):   # the ring path evaluates the same records in the same order; the occupancy point changes nothing
        assert np.array_equal(results[f"g{l}v15"][0], results[f"g{l}v4"][0]), l
        assert np.array_equal(results[f"g{l}v14"][0], results[f"g{l}v4"][0]), l
    assert np.array_equal(results["g2v4packed"][0], results["g2v4"][0])   # the layout of the neighbour table changes nothing


@pytest.mark.parametrize("case", ["shuffled", "holes", "strip"])
def test_vbd_wedge_records_on_irregular_input(ctx, oracle, case, monkeypatch):
    """Wedge records (streams.cuh: ST_WEDGE -- a fan triangle's membrane and its two hinges as one record of the colour pass) on
    input that does not pair up cleanly: triangle stencils rotated and hinge stencils reordered at random (the canonicalisation
    moves the order into the sign of the rest angle), a third of the hinges and a tenth of the triangles missing (what completes
    no wedge stays a single-incidence record), a two-row strip (every vertex on the boundary).  Against the generic CSR kernel
    (1e-11) and the reference (1e-9) after one frame."""
    rng = np.random.default_rng(7)
    sc = S.cloth(2, 40) if case == "strip" else S.cloth(18, 23)
    bend, mem = sc.energies[0], sc.energies[1]
    if case in ("shuffled", "holes"):
        rot = rng.integers(0, 3, size=mem.stencils.shape[0])
        mem.stencils = np.stack([np.roll(t, -r) for t, r in zip(mem.stencils, rot)]).astype(np.int32)
        flip = rng.random(bend.stencils.shape[0]) < 0.5
        swap = rng.random(bend.stencils.shape[0]) < 0.5
        st = bend.stencils.copy()
        st[flip, 0], st[flip, 1] = bend.stencils[flip, 1], bend.stencils[flip, 0]
        a, b = st[swap, 2].copy(), st[swap, 3].copy()
        st[swap, 2], st[swap, 3] = b, a
        bend.stencils = st
    if case == "holes":
        keep = rng.random(bend.stencils.shape[0]) > 0.33
        bend.stencils, bend.params = bend.stencils[keep], bend.params[keep]
        keep = rng.random(mem.stencils.shape[0]) > 0.1
        mem.stencils, mem.params = mem.stencils[keep], mem.params[keep]
    og = oracle_api.OracleGroup(oracle, sc)
    osol = oracle_api.OracleSolver(og, 0, 0.01, 2, 6)
    osol.step(1)
    xr, vr = osol.state()
    out = {}
    for name, env in (("csr", {"SPGX_NO_STREAMS": "1"}), ("wedge", {}), ("single", {"SPGX_VBD_WEDGE": "0"})):
        for k in ("SPGX_NO_STREAMS", "SPGX_VBD_WEDGE"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        g = api.Group.from_scene(ctx, sc)
        g.finalize(api.SPGX_VBD)
        info = g.kernel_info()
        if name == "wedge":
            n_corners = 3 * mem.stencils.shape[0]
            assert 0 < info["wedge_records"] < n_corners if case != "strip" else info["wedge_records"] >= 0, info
        g.vbd_step(0.01, 2, 6)
        out[name] = g.download()
        g.close()
    monkeypatch.delenv("SPGX_VBD_WEDGE", raising=False)
    for name in ("wedge", "single"):
        assert _rel(out[name][0], out["csr"][0]) <= 1e-11, name
        assert _rel(out[name][0], xr) <= RTOL and _rel(out[name][1], vr) <= 1e-8, name


def test_static_load_vbd_and_xpbd_refusal(ctx, oracle):
    """StaticLoadEnergy is energy-only: VBD matches the reference, XPBD throws the reference's message
    (staticLoadEnergy.cpp:8-13) through the C ABI."""
    import sys
    from pathlib import Path
    sys.path.insert(0, str(Path(__file__).resolve().parent / "golden"))
    import make_golden as MG
    sc = MG.beam_with_load(S.K_TET_SSNH)
    g, og, osol = _vbd_pair(ctx, oracle, sc, 1, 15)
    g.vbd_step(0.01, 1, 15)
    osol.step(1)
    x, v = g.download()
    xr, vr = osol.state()
    assert _rel(x, xr) <= RTOL and _rel(v, vr) <= 1e-8
    g.close()
    g2 = api.Group.from_scene(ctx, sc)
    g2.finalize(api.SPGX_XPBD)
    with pytest.raises(api.SpgxError, match="Constraint-based load energy not implemented"):
        g2.xpbd_step(0.01, 4, 1)
    g2.close()


@pytest.mark.parametrize("solver", [api.SPGX_VBD, api.SPGX_XPBD])
def test_debug_scan_finds_non_finite_state(ctx, solver):
    """spgx_check_finite (what Verbosity::Debug of the facade runs after every step): clean after a step, and a NaN planted
    in one vertex is found, counted and located in the caller's vertex numbering (VBD keeps its state in colour order)"""
    sc = S.cloth(12, 9) if solver == api.SPGX_VBD else S.cloth(12, 9, S.K_SPRING, S.K_QUADRATIC_BENDING)
    g = api.Group.from_scene(ctx, sc)
    g.finalize(solver)
    (g.vbd_step if solver == api.SPGX_VBD else g.xpbd_step)(0.01, 2, 3)
    assert g.check_finite() == (0, 0, -1)
    x, v = g.download()
    x[41, 1] = np.nan
    v[77] = np.inf
    g.upload(x, v)
    assert g.check_finite() == (1, 3, 41)
