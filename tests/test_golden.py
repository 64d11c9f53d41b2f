"""Golden vectors generated from the unmodified reference (tests/golden/make_golden.py -> golden_v1.npz).
CPU part: the oracle library that travels with the snapshot reproduces them exactly (the oracle is pinned), and the
product's host-side colouring matches them bit for bit.  GPU part: the CUDA path through the C ABI matches them within
the 1e-9 gate without needing the oracle library at all."""
import sys
from pathlib import Path

import numpy as np
import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent / "golden"))
import make_golden as MG  # noqa: E402
from spg_b200 import api, scenes as S  # noqa: E402

G = np.load(Path(__file__).resolve().parent / "golden" / "golden_v1.npz")
RTOL = 1e-9


def _rel(a, b):
    return np.abs(np.asarray(a) - b).max() / max(np.abs(b).max(), 1e-30)


def test_oracle_reproduces_golden(oracle):
    import oracle_api
    for i, (kind, x, x0, st, p) in enumerate(MG.fixtures()):
        og = oracle_api.OracleGroup(oracle, MG.fixture_scene(kind, x, x0, st, p))
        for local in range(len(st)):
            f, h = og.vertex_force_hessian(0, 0, local)
            assert _rel(f, G[f"fix{i}_k{kind}_f{local}"]) <= 1e-12 and _rel(h, G[f"fix{i}_k{kind}_h{local}"]) <= 1e-12
    for name in ("vbd_cloth9x13_2sub", "xpbd_snh_beam"):
        build, solver, nsub, iters, frames = MG.STEP_CASES[name]
        og = oracle_api.OracleGroup(oracle, build())
        osol = oracle_api.OracleSolver(og, solver, 0.01, nsub, iters)
        osol.step(frames)
        x, v = osol.state()
        if oracle.kind == "reference":
            assert np.array_equal(x, G[f"step_{name}_x"]) and np.array_equal(v, G[f"step_{name}_v"])
        else:
            assert _rel(x, G[f"step_{name}_x"]) <= RTOL


@pytest.mark.parametrize("name", sorted(MG.COLOR_CASES))
def test_host_colouring_matches_golden(name):
    sc = MG.COLOR_CASES[name]()
    sets = [e.stencils for e in sc.energies]
    assert np.array_equal(api.color_vertices(sc.n_vertices, sets), G[f"vcol_{name}"])
    for e, s in enumerate(sets):
        assert np.array_equal(api.color_stencils(s), G[f"scol_{name}_{e}"])


@pytest.mark.gpu
def test_gpu_elements_match_golden(ctx):
    for i, (kind, x, x0, st, p) in enumerate(MG.fixtures()):
        g = api.Group.from_scene(ctx, MG.fixture_scene(kind, x, x0, st, p))
        g.finalize(api.SPGX_VBD)
        for local in range(len(st)):
            for mode in (0, 1, 2):
                f, h = g.eval_vertex_force_hessian(0, 0, local, mode)
                assert _rel(f, G[f"fix{i}_k{kind}_f{local}"]) <= RTOL, (kind, local, mode)
                assert _rel(h, G[f"fix{i}_k{kind}_h{local}"]) <= RTOL, (kind, local, mode)
        if kind != S.K_STATIC_LOAD:  # energy-only kind: no constraint form (staticLoadEnergy.cpp:8-13)
            c, gc = g.eval_constraints(0, 0)
            assert np.abs(c - G[f"fix{i}_k{kind}_c"]).max() <= RTOL * max(np.abs(G[f"fix{i}_k{kind}_c"]).max(), 1.0)
            assert _rel(gc, G[f"fix{i}_k{kind}_gc"]) <= RTOL
        g.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(MG.STEP_CASES))
def test_gpu_steps_match_golden(ctx, name):
    build, solver, nsub, iters, frames = MG.STEP_CASES[name]
    sc = build()
    g = api.Group.from_scene(ctx, sc)
    g.finalize(api.SPGX_VBD if solver == 0 else api.SPGX_XPBD)
    if solver == 0:
        g.vbd_step(0.01, nsub, iters, frames=frames)
    else:
        g.xpbd_step(0.01, nsub, iters, frames=frames)
    x, v = g.download()
    assert _rel(x, G[f"step_{name}_x"]) <= RTOL * (10 if frames > 1 else 1)
    assert _rel(v, G[f"step_{name}_v"]) <= 1e-7
    g.close()
