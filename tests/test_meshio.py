"""Mesh ingestion and the headless runner: spgx::io::loadObj / loadMsh (include/spgx/io.hpp;
reference: spg/source/geom/io/obj.cpp:10-35, msh.cpp:10-41) and examples/headless.cpp, the window-less counterpart of
apps/solverCompare.cpp.  CPU tests cover the two parsers; GPU tests run a loaded mesh through the facade against the oracle."""
import subprocess
from pathlib import Path

import numpy as np
import pytest

import oracle_api
from spg_b200 import scenes as S

ROOT = Path(__file__).resolve().parent.parent


def _compile(tmp_path, source, name):
    exe = tmp_path / name
    subprocess.run(["g++", "-std=c++17", "-O1", "-I", str(ROOT / "include"), str(source), "-L", str(ROOT / "spg_b200" / "lib"),
                    "-lspgx", f"-Wl,-rpath,{ROOT / 'spg_b200' / 'lib'}", "-o", str(exe)], check=True)
    return exe


def _dump(exe, path):
    res = subprocess.run([str(exe), str(path)], capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(res.stderr.strip())
    lines = res.stdout.strip().splitlines()
    nv = int(lines[0].split()[1])
    verts = np.array([[float(a) for a in l.split()] for l in lines[1:1 + nv]]).reshape(nv, 3)
    _, ne, k = lines[1 + nv].split()
    elems = np.array([[int(a) for a in l.split()] for l in lines[2 + nv:2 + nv + int(ne)]], dtype=np.int32).reshape(int(ne), int(k))
    return verts, elems


def _write_obj(path, verts, faces):
    with open(path, "w") as f:
        f.write("# test mesh\n")
        for v in verts:
            f.write("v %.17g %.17g %.17g\n" % tuple(v))
        for t in faces:
            f.write("f %d %d %d\n" % tuple(t + 1))


def _write_msh41(path, verts, tets, extra_triangles=()):
    with open(path, "w") as f:
        f.write("$MeshFormat\n4.1 0 8\n$EndMeshFormat\n")
        f.write("$Nodes\n1 %d 1 %d\n3 1 0 %d\n" % (len(verts), len(verts), len(verts)))
        for i in range(len(verts)):
            f.write("%d\n" % (i + 1))
        for v in verts:
            f.write("%.17g %.17g %.17g\n" % tuple(v))
        f.write("$EndNodes\n")
        nblocks = 1 + (1 if len(extra_triangles) else 0)
        f.write("$Elements\n%d %d 1 %d\n" % (nblocks, len(tets) + len(extra_triangles), len(tets) + len(extra_triangles)))
        tag = 1
        if len(extra_triangles):
            f.write("2 1 2 %d\n" % len(extra_triangles))
            for t in extra_triangles:
                f.write("%d %d %d %d\n" % (tag, *(np.asarray(t) + 1)))
                tag += 1
        f.write("3 1 4 %d\n" % len(tets))
        for t in tets:
            f.write("%d %d %d %d %d\n" % (tag, *(t + 1)))
            tag += 1
        f.write("$EndElements\n")


def _write_msh22(path, verts, tets):
    with open(path, "w") as f:
        f.write("$MeshFormat\n2.2 0 8\n$EndMeshFormat\n$Nodes\n%d\n" % len(verts))
        for i, v in enumerate(verts):
            f.write("%d %.17g %.17g %.17g\n" % (i + 1, *v))
        f.write("$EndNodes\n$Elements\n%d\n" % (len(tets) + 1))
        f.write("1 2 2 0 1 1 2 3\n")  # a boundary triangle: skipped
        for i, t in enumerate(tets):
            f.write("%d 4 2 0 1 %d %d %d %d\n" % (i + 2, *(t + 1)))
        f.write("$EndElements\n")


# ---------------------------------------------------------------- CPU: the parsers
def test_load_obj_keeps_what_the_reference_keeps(tmp_path):
    exe = _compile(tmp_path, ROOT / "tests" / "native" / "io_dump.cpp", "io_dump")
    verts, _, faces = S.triangle_square_mesh(4, 5)
    p = tmp_path / "grid.obj"
    _write_obj(p, verts, faces)
    v, f = _dump(exe, p)
    assert np.array_equal(v, verts) and np.array_equal(f, faces)
    # texture / normal indices, a quad (fan-triangulated), relative indices, and a second shape that is ignored
    q = tmp_path / "mixed.obj"
    q.write_text("o first\nv 0 0 0\nv 1 0 0\nvt 0 0\nvn 0 0 1\nv 1 1 0\nv 0 1 0.5\n"
                 "f 1/1/1 2/1/1 3/1/1 4/1/1\nf -4 -3 -1\n"
                 "o second\nv 5 5 5\nf 1 2 5\n")
    v, f = _dump(exe, q)
    assert v.shape == (5, 3) and v[3, 2] == 0.5 and v[4, 0] == 5.0   # every vertex of the file (obj.cpp:21-24)
    assert f.tolist() == [[0, 1, 2], [0, 2, 3], [0, 1, 3]]            # the first shape's faces only (obj.cpp:26-31)
    bad = tmp_path / "bad.obj"
    bad.write_text("v 0 0 0\nv 1 0 0\nf 1 2 7\n")
    with pytest.raises(RuntimeError, match="out of range"):
        _dump(exe, bad)
    with pytest.raises(RuntimeError, match="cannot open"):
        _dump(exe, tmp_path / "missing.obj")


def test_load_msh_ascii_41_and_22(tmp_path):
    exe = _compile(tmp_path, ROOT / "tests" / "native" / "io_dump.cpp", "io_dump")
    verts, tets = S.tet_beam_mesh(3, 2, 2)
    tets = tets.astype(np.int32)
    p = tmp_path / "beam41.msh"
    _write_msh41(p, verts, tets, extra_triangles=[(0, 1, 2), (1, 2, 3)])
    v, t = _dump(exe, p)
    assert np.array_equal(v, verts) and np.array_equal(t, tets)   # 0-based node ids of the 4-node elements (msh.cpp:30-35)
    p = tmp_path / "beam22.msh"
    _write_msh22(p, verts, tets)
    v, t = _dump(exe, p)
    assert np.array_equal(v, verts) and np.array_equal(t, tets)
    b = tmp_path / "binary.msh"
    b.write_text("$MeshFormat\n4.1 1 8\n$EndMeshFormat\n")
    with pytest.raises(RuntimeError, match="binary"):
        _dump(exe, b)
    s = tmp_path / "sparse.msh"   # node tags that are not 1..n: (tag - 1) would not be the vertex index
    s.write_text("$MeshFormat\n2.2 0 8\n$EndMeshFormat\n$Nodes\n4\n1 0 0 0\n2 1 0 0\n3 0 1 0\n9 0 0 1\n$EndNodes\n"
                 "$Elements\n1\n1 4 2 0 1 1 2 3 9\n$EndElements\n")
    with pytest.raises(RuntimeError, match="node tags"):
        _dump(exe, s)


def test_triangle_and_tetrahedral_mesh_tables(tmp_path):
    """spgx::TriangleMesh / TetrahedralMesh (include/spgx/geom.hpp): edge order and opposing vertices as the reference's
    classes give them (triangleMesh.h:36-50, tetrahedralMesh.h:32-46); scenes.triangle_edges is the same table, checked
    against the compiled reference's createCloth stencils in test_scenes.py."""
    exe = _compile(tmp_path, ROOT / "tests" / "native" / "geom_dump.cpp", "geom_dump")
    res = subprocess.run([str(exe), "selftest"], capture_output=True, text=True)
    assert res.returncode == 0 and res.stdout.strip() == "ok", res.stdout + res.stderr
    _, verts0, faces = S.triangle_square_mesh(5, 7)
    obj = tmp_path / "grid.obj"
    _write_obj(obj, verts0, faces)
    lines = subprocess.run([str(exe), str(obj)], capture_output=True, text=True, check=True).stdout.strip().splitlines()
    edges, count, opps = S.triangle_edges(faces)
    assert int(lines[0].split()[1]) == len(edges)
    for row, e, c, o in zip(lines[1:], edges, count, opps):
        t = [int(a) for a in row.split()]
        assert t[0] == e[0] and t[1] == e[1] and t[2] == c
        assert [t[4 + 2 * k] for k in range(c)] == [int(a) for a in o[:c]]
        for k in range(c):   # the triangle named really holds the edge and the opposite vertex
            assert set(faces[t[3 + 2 * k]]) == {e[0], e[1], t[4 + 2 * k]}
    verts, tets = S.tet_beam_mesh(3, 2, 3)
    msh = tmp_path / "beam.msh"
    _write_msh41(msh, verts, tets.astype(np.int32))
    lines = subprocess.run([str(exe), str(msh)], capture_output=True, text=True, check=True).stdout.strip().splitlines()
    got = np.array([[int(a) for a in l.split()] for l in lines[1:]])
    pairs = np.concatenate([np.sort(tets[:, [i, j]], axis=1) for i in range(3) for j in range(i + 1, 4)])
    assert np.array_equal(got, np.unique(pairs, axis=0))


# ---------------------------------------------------------------- GPU: a loaded mesh through the facade vs the oracle
def _run_headless(exe, *args):
    res = subprocess.run([str(exe), *map(str, args)], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    return np.array([[float(a) for a in l.split()] for l in res.stdout.strip().splitlines()])


@pytest.mark.gpu
@pytest.mark.parametrize("solver,okind,substeps,iters,tol", [("vbd", 0, 2, 10, 1e-8), ("xpbd", 1, 10, 1, 1e-8),
                                                             ("newton_dx", 4, 2, 1, 1e-6)])
def test_headless_obj_cloth_matches_oracle(tmp_path, oracle, solver, okind, substeps, iters, tol):
    exe = _compile(tmp_path, ROOT / "examples" / "headless.cpp", "headless")
    n = 12
    ref_scene = S.cloth(n, n)   # createCloth: rest pose in the z = 0 plane, initial pose laid flat, anchors on vertices 0, n - 1
    _, verts0, faces = S.triangle_square_mesh(n, n)
    obj = tmp_path / "cloth.obj"
    _write_obj(obj, verts0, faces)
    anchors = [0, n - 1]
    frames = 4
    x = _run_headless(exe, obj, solver, frames, substeps, iters, ",".join(map(str, anchors)), tmp_path / "frame")
    og = oracle_api.OracleGroup(oracle, ref_scene)
    sol = oracle_api.OracleSolver(og, okind, 0.01, substeps, iters)
    sol.step(frames)
    xr, _ = sol.state()
    assert np.abs(x - xr).max() <= tol * np.abs(xr).max()
    # the frame dump: one .obj per frame, the last one holding the final positions and the mesh's faces
    exe_dump = _compile(tmp_path, ROOT / "tests" / "native" / "io_dump.cpp", "io_dump")
    assert sorted(p.name for p in tmp_path.glob("frame*.obj")) == ["frame%04d.obj" % f for f in range(frames)]
    v, f = _dump(exe_dump, tmp_path / ("frame%04d.obj" % (frames - 1)))
    assert np.array_equal(v, x) and np.array_equal(f, faces)


@pytest.mark.gpu
@pytest.mark.parametrize("solver,okind,substeps,iters,tol", [("vbd", 0, 2, 20, 1e-8), ("baraff_witkin", 3, 1, 1, 1e-6)])
def test_headless_msh_beam_matches_oracle(tmp_path, oracle, solver, okind, substeps, iters, tol):
    exe = _compile(tmp_path, ROOT / "examples" / "headless.cpp", "headless")
    sc = S.tet_beam(6, 3, 3)
    tets = sc.energies[1].stencils.astype(np.int32)
    anchors = sc.energies[0].stencils.ravel()
    msh = tmp_path / "beam.msh"
    _write_msh41(msh, sc.x, tets)
    frames = 3
    x = _run_headless(exe, msh, solver, frames, substeps, iters, ",".join(map(str, anchors)))
    og = oracle_api.OracleGroup(oracle, sc)
    sol = oracle_api.OracleSolver(og, okind, 0.01, substeps, iters)
    sol.step(frames)
    xr, _ = sol.state()
    assert np.abs(x - xr).max() <= tol * np.abs(xr).max()
