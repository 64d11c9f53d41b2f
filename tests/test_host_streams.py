"""Host-side check of the canonical stream evaluators (spg_b200/csrc/streams.cuh: the hand-derived owner-vertex blocks
the streaming VBD kernel runs, including the host canonicalisation of discrete-bending stencils) against the generic
second-order dual-number evaluation of the same energy (the restatement of energy.h:240-258), on 2000 random stencils
per kind and every owner position.  The evaluators are __host__ __device__, so this compiles them with g++."""
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_stream_evaluators_match_generic_duals(tmp_path):
    exe = tmp_path / "host_streams"
    subprocess.run(["g++", "-std=c++17", "-O2", "-o", str(exe), str(ROOT / "tests" / "native" / "host_streams.cpp")],
                   check=True)
    res = subprocess.run([str(exe)], capture_output=True, text=True)
    assert res.returncode == 0, res.stdout
    lines = dict(l.split(" max_rel ") for l in res.stdout.strip().splitlines())
    # "wedge": Stream<ST_WEDGE> (a fan triangle's membrane + the two hinges sharing its normal) against the sum of the three generic
    # evaluations, with the three stencils in every vertex order makeWedgeRecord has to undo
    assert set(lines) == {"discrete_bending", "membrane_stvk", "tet_snh", "tet_stvk", "wedge"}
    assert all(float(v) < 1e-10 for v in lines.values()), lines


def test_hand_derived_hinge_hessian_matches_generic_duals(tmp_path):
    """spg_b200/csrc/hinge_hessian.cuh (the full 12 x 12 discrete-bending Hessian the block-sparse assembly uses) against
    Dual<12, true> on the same energy: 2000 random hinges, flat, nearly flat and folded."""
    exe = tmp_path / "host_hinge_hessian"
    subprocess.run(["g++", "-std=c++17", "-O2", "-o", str(exe), str(ROOT / "tests" / "native" / "host_hinge_hessian.cpp")],
                   check=True)
    res = subprocess.run([str(exe)], capture_output=True, text=True)
    assert res.returncode == 0, res.stdout
    assert float(res.stdout.split("max_rel")[1]) < 1e-11, res.stdout


def test_hand_derived_membrane_hessian_matches_generic_duals(tmp_path):
    """spg_b200/csrc/membrane_hessian.cuh (the full 9 x 9 StVK membrane Hessian; opt-in in the assembly until measured on the
    device) against Dual<9, true> on the same energy: 2000 random triangles."""
    exe = tmp_path / "host_membrane_hessian"
    subprocess.run(["g++", "-std=c++17", "-O2", "-o", str(exe), str(ROOT / "tests" / "native" / "host_membrane_hessian.cpp")],
                   check=True)
    res = subprocess.run([str(exe)], capture_output=True, text=True)
    assert res.returncode == 0, res.stdout
    assert float(res.stdout.split("max_rel")[1]) < 1e-11, res.stdout


def test_stream_records_of_whole_meshes_replay_to_the_generic_sums(tmp_path):
    """spg_b200/csrc/stream_records.h (which incidence of a vertex becomes which record of the streaming colour pass: wedges
    where a fan triangle's membrane and two hinges pair up, single-incidence records otherwise) on four triangulated grids:
    the records of every vertex, replayed with the canonical evaluators, give the generic per-vertex gradient and block; no
    incidence is lost or doubled; interior vertices pair up completely."""
    exe = tmp_path / "host_stream_records"
    subprocess.run(["g++", "-std=c++17", "-O2", "-o", str(exe), str(ROOT / "tests" / "native" / "host_stream_records.cpp")],
                   check=True)
    res = subprocess.run([str(exe)], capture_output=True, text=True)
    assert res.returncode == 0, res.stdout
    lines = res.stdout.strip().splitlines()
    assert len(lines) == 4
    for l in lines:
        f = l.split()
        assert float(f[2]) < 1e-10 and int(f[4]) > 0 and int(f[8]) == 0 and int(f[10]) == 0, l
