"""The BASELINE configs AT THEIR FULL SIZE against the reference itself (SURVEY.md section 8c/8d gates).

tests/golden/atsize_c{2,3,4}.npz hold what the UNMODIFIED reference (oracle/_ref) computes for one frame of
  config 2: cloth 256 x 256, springs + quadratic bending + 2 anchors, XPBD 20 substeps x 1 pass (the reference coloured it),
  config 3: cloth 1024 x 1024, StVK membrane + discrete bending + 2 anchors, VBD 10 iterations (the reference coloured it),
  config 4: stable Neo-Hookean beam 149 x 38 x 38 (1 013 060 tets), VBD 20 iterations (the reference coloured it),
written by tests/golden/make_golden_atsize.py (committed; ~10 min of host time for config 3): sampled vertices, block sums
over every vertex, and a hash of the reference's own colouring.  The GPU path (through the C ABI) must reproduce
  * the colouring bit for bit (sha256 of the int32 colours), * positions and velocities of the sampled vertices within
  1e-9 relative (tets 1e-8, as at small size), * the block sums of (x - x_initial) and v -- every vertex is in one -- within
  the bound the per-vertex tolerance implies for a block.
`SPGX_ATSIZE_LIVE=1` additionally steps the oracle library live on the host cores and compares EVERY vertex (minutes).
The fixture-only half of this file (`-m "not gpu"`) checks the fixtures themselves: shapes, the sampled indices, and that
the hash they carry equals the linear-time host colouring (no device needed)."""
import hashlib
import os
import sys
from pathlib import Path

import numpy as np
import pytest

GOLD = Path(__file__).resolve().parent / "golden"
sys.path.insert(0, str(GOLD))
import make_golden_atsize as MG  # noqa: E402
from spg_b200 import api, scenes as S  # noqa: E402

# config: (position tolerance, velocity tolerance), relative to the largest |x| / |v| of the reference state
TOL = {2: (1e-9, 1e-9), 3: (1e-9, 1e-9), 4: (1e-8, 1e-7), 5: (1e-9, 1e-9)}


def _fixture(cfg):
    f = GOLD / f"atsize_c{cfg}.npz"
    if not f.exists():
        pytest.skip(f"{f.name} not generated (tests/golden/make_golden_atsize.py --config {cfg})")
    return np.load(f)


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.int32).tobytes()).hexdigest()


def _scene(cfg):
    return MG.CONFIGS[cfg][0](S)


@pytest.mark.parametrize("cfg", [2, 3, 4, 5])
def test_fixture_is_consistent_with_the_host_colouring(cfg):
    """No device: the fixture's colour hash (the reference's own colouring) equals the linear-time restatement's at full
    size, the sample is the one the generator defines, and the block partition covers every vertex once."""
    gold = _fixture(cfg)
    if cfg == 5 and not os.environ.get("SPGX_ATSIZE_C5_CPU"):
        pytest.skip("16.8 M-vertex host colouring takes a minute: set SPGX_ATSIZE_C5_CPU=1")
    sc = _scene(cfg)
    kind, substeps, iterations = (int(k) for k in gold["solver"])
    assert sc.n_vertices == int(gold["n_vertices"])
    assert np.array_equal(gold["idx"], MG.CONFIGS[cfg][5]())
    edges = gold["blk_edges"]
    assert edges[0] == 0 and edges[-1] == sc.n_vertices and (np.diff(edges) > 0).all()
    if kind == 0:
        assert _sha(api.color_vertices(sc.n_vertices, [e.stencils for e in sc.energies])) == str(gold["colors_sha256"][0])
    else:
        for e, h in zip(sc.energies, gold["colors_sha256"]):
            assert _sha(api.color_stencils(e.stencils)) == str(h)
    assert np.isfinite(gold["x"]).all() and np.isfinite(gold["v"]).all()


def _step_gpu(ctx, sc, gold):
    kind, substeps, iterations = (int(k) for k in gold["solver"])
    dt = float(gold["dt"])
    g = api.Group.from_scene(ctx, sc)
    if kind == 0:
        g.finalize(api.SPGX_VBD)
        assert _sha(g.vertex_colors()) == str(gold["colors_sha256"][0])   # bit-exact colouring at full size
        assert g.num_colors() == int(gold["ncolors"][0])
        g.vbd_step(dt, substeps, iterations)
    else:
        g.finalize(api.SPGX_XPBD)
        for e in range(len(sc.energies)):
            assert _sha(g.stencil_colors(e)) == str(gold["colors_sha256"][e])
            assert g.num_colors(e) == int(gold["ncolors"][e])
        g.xpbd_step(dt, substeps, iterations)
    x, v = g.download()
    g.close()
    return x, v


def _compare(cfg, sc, gold, x, v):
    tx, tv = TOL[cfg]
    xs, vs = float(gold["x_absmax"]), float(gold["v_absmax"])
    idx = gold["idx"]
    ex = np.abs(x[idx] - gold["x"]).max() / xs
    ev = np.abs(v[idx] - gold["v"]).max() / vs
    assert ex <= tx and ev <= tv, (ex, ev)
    # every vertex: block sums.  A block of n vertices each within tol * scale is within n * tol * scale; the observed
    # differences are three orders below (rounding, not bias), so a single vertex off by 1e-6 already fails
    edges = gold["blk_edges"]
    n = np.diff(edges)[:, None]
    bdx = np.add.reduceat(x - sc.x, edges[:-1], axis=0)
    bv = np.add.reduceat(v, edges[:-1], axis=0)
    edx = (np.abs(bdx - gold["blk_dx"]) / (n * xs)).max()
    edv = (np.abs(bv - gold["blk_v"]) / (n * vs)).max()
    assert edx <= tx and edv <= tv, (edx, edv)
    return ex, ev, edx, edv


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [2, 3, 4])
def test_config_matches_reference_at_full_size(ctx, cfg):
    gold = _fixture(cfg)
    sc = _scene(cfg)
    x, v = _step_gpu(ctx, sc, gold)
    assert np.isfinite(x).all() and np.isfinite(v).all()
    errs = _compare(cfg, sc, gold, x, v)
    print(f"config {cfg} at full size vs the reference: sample x {errs[0]:.2e} v {errs[1]:.2e}; block sums dx {errs[2]:.2e} v {errs[3]:.2e}")
    if cfg == 3:
        # the two pinned corners, as the round-1 review quotes them from its own run of the reference
        assert np.allclose(x[0], [-1.791162e-4, -4.825037e-4, 1.791162e-4], rtol=2e-7)
        assert np.allclose(x[1023], [1.0000855856, -7.3066227632e-4, 8.5585374163e-5], rtol=2e-7)


@pytest.mark.gpu
def test_config5_matches_reference_at_full_size(ctx):
    """cloth 4096 x 4096 (16.8 M vertices) on one GPU against the reference run (colours injected there: its own quadratic
    colouring is infeasible at this size; the injected colours equal the reference's at 1024 x 1024, see config 3)."""
    gold = _fixture(5)
    sc = _scene(5)
    x, v = _step_gpu(ctx, sc, gold)
    _compare(5, sc, gold, x, v)


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [2, 3, 4])
def test_config_matches_live_oracle_everywhere(ctx, oracle, cfg):
    """Opt-in (minutes of host time): the oracle library stepped live, every vertex compared."""
    if not os.environ.get("SPGX_ATSIZE_LIVE"):
        pytest.skip("set SPGX_ATSIZE_LIVE=1 to step the oracle at full size on the host cores")
    import oracle_api
    gold = _fixture(cfg)
    sc = _scene(cfg)
    kind, substeps, iterations = (int(k) for k in gold["solver"])
    x, v = _step_gpu(ctx, sc, gold)
    oracle.set_threads(len(os.sched_getaffinity(0)))
    osol = oracle_api.OracleSolver(oracle_api.OracleGroup(oracle, sc), kind, float(gold["dt"]), substeps, iterations)
    if kind == 0:
        osol.inject_vbd_colors(api.color_vertices(sc.n_vertices, [e.stencils for e in sc.energies]))
    osol.step(1)
    xr, vr = osol.state()
    tx, tv = TOL[cfg]
    assert np.abs(x - xr).max() <= tx * np.abs(xr).max() and np.abs(v - vr).max() <= tv * np.abs(vr).max()
