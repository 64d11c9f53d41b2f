"""Generates tests/golden/golden_v1.npz from the UNMODIFIED reference (oracle/_ref/libspgref.so, built from
/root/reference by oracle/Makefile).  Run in the build container:  python tests/golden/make_golden.py
The reference ships no golden vectors of its own, so these pin the oracle AND the CUDA path to
outputs of the reference itself; the GPU box never needs /root/reference."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "tests")]
import oracle_api  # noqa: E402
from spg_b200 import scenes as S  # noqa: E402


def fixtures():
    """single-stencil known-answer inputs in the spirit of apps/energyViewer.cpp:20-137"""
    x0 = np.array([[0, 0, 0], [1, 0, 0], [0.3, 0.9, 0], [0.4, -0.8, 0]], float)
    x = np.array([[0, 0, 0.1], [1.1, 0, 0], [0.3, 0.8, 0.4], [0.5, -0.7, 0.3]], float)
    t0 = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], float)
    t = t0 + np.array([[0, 0, 0], [0.2, 0.1, 0], [0.1, 0.3, -0.1], [0.05, 0.2, 0.4]])
    out = []
    for kind, p in [(S.K_DISCRETE_BENDING, [0.1, 0.7]), (S.K_QUADRATIC_BENDING, [0.0, 0.7])]:
        out.append((kind, x, x0, [0, 1, 2, 3], p))
    for kind, p in [(S.K_MEMBRANE_STVK, [3.0, 0.3]), (S.K_MEMBRANE_CHOI, [3.0, 2.0, 1.0]), (S.K_MEMBRANE_BW, [3.0, 2.0, 1.0])]:
        out.append((kind, x[:3], x0[:3], [0, 1, 2], p))
    for kind in (S.K_TET_SNH, S.K_TET_STVK):
        out.append((kind, t, t0, [0, 1, 2, 3], [10.0, 0.3]))
    for kind in (S.K_SPRING, S.K_SPRING_CONTINUUM):
        out.append((kind, x[:2], x0[:2], [0, 1], [0.9, 5.0]))
    out.append((S.K_ANCHOR, x[:1], x0[:1], [0], [0.2, -0.1, 0.3, 50.0]))
    # SURVEY.md 8f rank 1 (appended: the indices of the fixtures above are part of the golden keys)
    out.append((S.K_BENDING_BW, x, x0, [0, 1, 2, 3], [0.1, 0.7]))
    out.append((S.K_SPRING_SQUARED_CONTINUUM, x[:2], x0[:2], [0, 1], [0.9, 5.0]))
    out.append((S.K_TET_SSNH, t, t0, [0, 1, 2, 3], [10.0, 0.3]))
    out.append((S.K_STATIC_LOAD, x[:1], x0[:1], [0], [0.3, -2.0, 0.5, 7.0]))
    return out


def fixture_scene(kind, x, x0, st, p):
    sc = S.Scene(np.array(x, float).copy(), np.array(x0, float).copy(), np.ones(len(x)))
    sc.energies.append(S.SceneEnergy(kind, np.array([st], np.int32), np.array([p], float)))
    return sc


def beam_with_load(kind):
    """a tet beam whose free end face is pulled sideways by a StaticLoadEnergy (energy-only kind: VBD)"""
    sc = S.tet_beam(5, 3, 3, kind)
    ids = np.arange(sc.n_vertices - 9, sc.n_vertices, dtype=np.int32)
    p = np.tile(np.array([0.2, 0.0, 1.0, 0.05]), (ids.size, 1))
    sc.energies.append(S.SceneEnergy(S.K_STATIC_LOAD, ids[:, None].copy(), p))
    return sc


STEP_CASES = {
    # name: (scene builder, solver kind (0 VBD, 1 XPBD), substeps, iterations, frames)
    "vbd_cloth16": (lambda: S.cloth(16, 16), 0, 1, 10, 2),
    "vbd_cloth9x13_2sub": (lambda: S.cloth(9, 13), 0, 2, 5, 1),
    "xpbd_cloth16": (lambda: S.cloth(16, 16, S.K_SPRING, S.K_QUADRATIC_BENDING), 1, 20, 1, 1),
    "vbd_snh_beam": (lambda: S.tet_beam(5, 3, 3, S.K_TET_SNH), 0, 1, 20, 1),
    "vbd_stvk_beam": (lambda: S.tet_beam(5, 3, 3, S.K_TET_STVK), 0, 1, 20, 1),
    "xpbd_snh_beam": (lambda: S.tet_beam(5, 3, 3, S.K_TET_SNH), 1, 10, 1, 1),
    "vbd_cloth_bw": (lambda: S.cloth(11, 9, S.K_MEMBRANE_BW, S.K_BENDING_BW), 0, 1, 8, 1),
    "xpbd_cloth_bw_sqspring": (lambda: S.cloth(11, 9, S.K_SPRING_SQUARED_CONTINUUM, S.K_BENDING_BW), 1, 10, 1, 1),
    "vbd_ssnh_beam_load": (lambda: beam_with_load(S.K_TET_SSNH), 0, 1, 15, 1),
    "xpbd_ssnh_beam": (lambda: S.tet_beam(5, 3, 3, S.K_TET_SSNH), 1, 10, 1, 1),
}
COLOR_CASES = {
    "cloth16": lambda: S.cloth(16, 16),
    "cloth7x5": lambda: S.cloth(7, 5),
    "cloth16_xpbd": lambda: S.cloth(16, 16, S.K_SPRING, S.K_QUADRATIC_BENDING),
    "beam5x3x3": lambda: S.tet_beam(5, 3, 3),
}


def main():
    o = oracle_api.load("reference")
    assert o.kind == "reference"
    g = {}
    for i, (kind, x, x0, st, p) in enumerate(fixtures()):
        og = oracle_api.OracleGroup(o, fixture_scene(kind, x, x0, st, p))
        for local in range(len(st)):
            f, h = og.vertex_force_hessian(0, 0, local)
            g[f"fix{i}_k{kind}_f{local}"] = f
            g[f"fix{i}_k{kind}_h{local}"] = h
        if kind != S.K_STATIC_LOAD:  # its constraint form throws (staticLoadEnergy.cpp:8-13)
            c, gc = og.constraints(0, 0)
            g[f"fix{i}_k{kind}_c"] = c
            g[f"fix{i}_k{kind}_gc"] = gc
    for name, build in COLOR_CASES.items():
        sc = build()
        sets = [e.stencils for e in sc.energies]
        g[f"vcol_{name}"] = o.color_vertices(sc.n_vertices, sets)
        for e, s in enumerate(sets):
            g[f"scol_{name}_{e}"] = o.color_stencils(s)
    for name, (build, solver, nsub, iters, frames) in STEP_CASES.items():
        sc = build()
        og = oracle_api.OracleGroup(o, sc)
        osol = oracle_api.OracleSolver(og, solver, 0.01, nsub, iters)
        osol.step(frames)
        x, v = osol.state()
        g[f"step_{name}_x"] = x
        g[f"step_{name}_v"] = v
    np.savez_compressed(Path(__file__).with_name("golden_v1.npz"), **g)
    print("wrote", len(g), "arrays")


if __name__ == "__main__":
    main()
