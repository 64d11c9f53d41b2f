"""Generates tests/golden/golden_v2_implicit.npz from the UNMODIFIED reference (oracle/_ref/libspgref.so): the
implicit-Euler family and symplectic Euler.  Run in the build container:
    python tests/golden/make_golden_implicit.py
Holds, per case: the system force (getSystemForce) and stiffness matrix (getSystemStiffnessMatrix, COO) after one
reference frame, positions / velocities / CG iteration counts after the stated frames of each solver, and the full
element Hessians (energyHessian) of the single-stencil fixtures of make_golden.py."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "tests"), str(ROOT / "tests" / "golden")]
import oracle_api  # noqa: E402
import make_golden as MG  # noqa: E402
from spg_b200 import scenes as S  # noqa: E402

# oracle solver kinds (oracle/ref_driver.cpp: ref_solver_create)
BW, NDX, NDV, SYMPLECTIC, BDF2, STATIC, QUASI, NROBUST, QROBUST = 3, 4, 5, 6, 7, 8, 9, 10, 11
SCENES = {
    "cloth12": lambda: S.cloth(12, 12),
    "cloth10_quad": lambda: S.cloth(10, 10, S.K_SPRING, S.K_QUADRATIC_BENDING),
    "cloth9x7_choi_bw": lambda: S.cloth(9, 7, S.K_MEMBRANE_CHOI, S.K_BENDING_BW),
    "beam_snh": lambda: S.tet_beam(5, 3, 3, S.K_TET_SNH),
    "beam_stvk": lambda: S.tet_beam(5, 3, 3, S.K_TET_STVK),
}
# name: (scene, solver kind, substeps, frames)
STEP_CASES = {}
for _sc in SCENES:
    for _k, _kn in ((BW, "bw"), (NDX, "ndx"), (NDV, "ndv")):
        STEP_CASES[f"{_kn}_{_sc}"] = (_sc, _k, 1, 2)
STEP_CASES["ndx_cloth12_3sub"] = ("cloth12", NDX, 3, 1)
# BDF2 (keeps the previous state across substeps and frames), quasi-static and static Newton (the static solve needs a
# scene without zero-energy modes: the anchored beams)
for _sc in ("cloth12", "cloth10_quad", "beam_snh"):
    STEP_CASES[f"bdf2_{_sc}"] = (_sc, BDF2, 2, 2)
    STEP_CASES[f"quasi_{_sc}"] = (_sc, QUASI, 1, 3)
for _sc in ("beam_snh", "beam_stvk"):
    STEP_CASES[f"static_{_sc}"] = (_sc, STATIC, 1, 3)
# the adaptive solvers: 7 frames so that the "five good substeps in a row" branch is reached
for _sc in ("cloth12", "cloth10_quad", "beam_snh"):
    STEP_CASES[f"nrobust_{_sc}"] = (_sc, NROBUST, 1, 7)
    STEP_CASES[f"qrobust_{_sc}"] = (_sc, QROBUST, 2, 4)
STEP_CASES["symplectic_cloth10_quad"] = ("cloth10_quad", SYMPLECTIC, 200, 1)
STEP_CASES["symplectic_beam_snh"] = ("beam_snh", SYMPLECTIC, 400, 1)


def main():
    o = oracle_api.load("reference")
    assert o.kind == "reference"
    g = {}
    for i, (kind, x, x0, st, p) in enumerate(MG.fixtures()):
        og = oracle_api.OracleGroup(o, MG.fixture_scene(kind, x, x0, st, p))
        grad, hess = og.grad_hess(0, 0)
        g[f"fix{i}_k{kind}_grad"] = grad
        g[f"fix{i}_k{kind}_hess"] = hess
    for name, build in SCENES.items():
        sc = build()
        og = oracle_api.OracleGroup(o, sc)
        sol = oracle_api.OracleSolver(og, NDX, 0.01, 1, 1)
        sol.step(1)  # a deformed, moving state
        x, v = sol.state()
        g[f"sys_{name}_x"], g[f"sys_{name}_v"] = x, v
        g[f"sys_{name}_f"] = sol.system_force()
        K = sol.system_matrix(0).tocoo()
        g[f"sys_{name}_Krow"], g[f"sys_{name}_Kcol"], g[f"sys_{name}_Kval"] = K.row.astype(np.int32), K.col.astype(np.int32), K.data
    for name, (scn, kind, nsub, frames) in STEP_CASES.items():
        og = oracle_api.OracleGroup(o, SCENES[scn]())
        sol = oracle_api.OracleSolver(og, kind, 0.01, nsub, 1)
        sol.step(frames)
        x, v = sol.state()
        g[f"step_{name}_x"], g[f"step_{name}_v"] = x, v
        if kind != SYMPLECTIC:
            g[f"step_{name}_cg"] = sol.cg_iterations()
    np.savez_compressed(Path(__file__).with_name("golden_v2_implicit.npz"), **g)
    print("wrote", len(g), "arrays")


if __name__ == "__main__":
    main()
