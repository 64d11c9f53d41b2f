"""The C++ facade (include/spgx/spgx.hpp) driven by examples/rope_facade.cpp, against the oracle on the same rope
(apps/solverCompare.cpp:56-102 style scene): VBD and XPBD, with a live anchor edit half way
(SpringAnchorEnergy::updateAnchor, solverCompare.cpp:1524)."""
import subprocess
from pathlib import Path

import numpy as np
import pytest

import oracle_api
from spg_b200 import scenes as S

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def _rope_scene(n):
    dl = 1.0 / (n - 1)
    x = np.zeros((n, 3))
    x[:, 0] = np.arange(n) * dl
    sc = S.Scene(x.copy(), x.copy(), np.full(n, 1.0 / n))
    st = np.stack([np.arange(n - 1), np.arange(1, n)], 1).astype(np.int32)
    sc.energies.append(S.SceneEnergy(S.K_SPRING, st, np.tile([dl, 1e3], (n - 1, 1))))
    sc.energies.append(S.SceneEnergy(S.K_ANCHOR, np.array([[0]], np.int32), np.array([[0.0, 0.0, 0.0, 1e5]])))
    return sc


@pytest.mark.parametrize("solver,substeps,iters", [("vbd", 2, 8), ("xpbd", 10, 1)])
def test_rope_facade_matches_oracle(tmp_path, oracle, solver, substeps, iters):
    exe = tmp_path / "rope_facade"
    subprocess.run(["g++", "-std=c++17", "-O1", "-I", str(ROOT / "include"), str(ROOT / "examples" / "rope_facade.cpp"),
                    "-L", str(ROOT / "spg_b200" / "lib"), "-lspgx", f"-Wl,-rpath,{ROOT / 'spg_b200' / 'lib'}", "-o", str(exe)],
                   check=True)
    n, frames = 24, 6
    res = subprocess.run([str(exe), solver, str(n), str(frames), str(substeps), str(iters)], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    x = np.array([[float(a) for a in l.split()] for l in res.stdout.strip().splitlines()])
    sc = _rope_scene(n)
    og = oracle_api.OracleGroup(oracle, sc)
    osol = oracle_api.OracleSolver(og, 0 if solver == "vbd" else 1, 0.01, substeps, iters)
    osol.step(frames // 2)
    og.update_anchor(1, 0, [0.0, 0.05, 0.0])
    osol.step(frames - frames // 2)
    xr, _ = osol.state()
    assert np.abs(x - xr).max() <= 1e-9 * np.abs(xr).max()


@pytest.mark.parametrize("solver,substeps,iters", [("vbd", 2, 8), ("xpbd", 10, 1)])
def test_rope_facade_picking_matches_oracle(tmp_path, oracle, solver, substeps, iters):
    """Picking through the facade (addEnergy / removeEnergy on the solver's copy of the group + requirePrecomputationUpdate:
    apps/solverCompare.cpp:717-726, :1446-1455, :1499-1502) against the same edits on the reference solver."""
    exe = tmp_path / "rope_facade"
    subprocess.run(["g++", "-std=c++17", "-O1", "-I", str(ROOT / "include"), str(ROOT / "examples" / "rope_facade.cpp"),
                    "-L", str(ROOT / "spg_b200" / "lib"), "-lspgx", f"-Wl,-rpath,{ROOT / 'spg_b200' / 'lib'}", "-o", str(exe)],
                   check=True)
    n, frames = 24, 9
    res = subprocess.run([str(exe), solver, str(n), str(frames), str(substeps), str(iters), "pick"], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    x = np.array([[float(a) for a in l.split()] for l in res.stdout.strip().splitlines()])
    sc = _rope_scene(n)
    og = oracle_api.OracleGroup(oracle, sc)
    osol = oracle_api.OracleSolver(og, 0 if solver == "vbd" else 1, 0.01, substeps, iters)
    picked = False
    for f in range(frames):
        if f == frames // 3:
            xs, _ = osol.state()
            osol.add_energy(S.K_ANCHOR, np.array([[n - 1]], np.int32), np.array([[*xs[n - 1], 1e4]]))
            picked = True
        if picked and f > frames // 3:
            osol.update_anchor(2, 0, [1.0, 0.01 * f, 0.02])
        if f == 2 * frames // 3:
            osol.remove_energy(2)
            picked = False
        osol.step(1)
    xr, _ = osol.state()
    assert np.abs(x - xr).max() <= 1e-8 * np.abs(xr).max()


@pytest.mark.parametrize("solver,substeps,iters", [("vbd", 2, 6), ("xpbd", 10, 1)])
def test_rigid_body_chain_facade_matches_reference(tmp_path, ref_oracle, solver, substeps, iters):
    """RigidBodyGroup + RBSpringEnergy + RBSpringAnchorEnergy through the facade (examples/rb_chain_facade.cpp builds the
    reference application's chain scene, apps/solverCompare.cpp:605-660, and drags the anchor half way) against the compiled
    reference on the same scene and the same edit"""
    exe = tmp_path / "rb_chain_facade"
    subprocess.run(["g++", "-std=c++17", "-O1", "-I", str(ROOT / "include"), str(ROOT / "examples" / "rb_chain_facade.cpp"),
                    "-L", str(ROOT / "spg_b200" / "lib"), "-lspgx", f"-Wl,-rpath,{ROOT / 'spg_b200' / 'lib'}", "-o", str(exe)],
                   check=True)
    n, frames = 6, 6
    res = subprocess.run([str(exe), solver, str(n), str(frames), str(substeps), str(iters)], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    out = np.array([[float(a) for a in l.split()] for l in res.stdout.strip().splitlines()])
    sc = S.rb_chain(n)
    og = oracle_api.OracleRbGroup(ref_oracle, sc)
    osol = oracle_api.OracleRbSolver(og, 0 if solver == "vbd" else 1, 0.01, substeps, iters)
    osol.step(frames // 2)
    osol.update_anchor(0, 0, [0.7, 0.6, 0.1])
    osol.step(frames - frames // 2)
    xr, thr = osol.state()[:2]
    assert np.abs(out[:, :3] - xr).max() <= 1e-8 * np.abs(xr).max()
    assert np.abs(out[:, 3:] - thr).max() <= 1e-7 * max(np.abs(thr).max(), 1e-30)
