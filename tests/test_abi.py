"""The C-ABI library loads and exports every symbol include/spgx.h declares (no compute calls without a GPU), the
ctypes signature table covers the header exactly, and the C++ facade + example compile against it."""
import ctypes
import re
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
HEADER = (ROOT / "include" / "spgx.h").read_text()


def _declared():
    body = re.sub(r"/\*.*?\*/", "", HEADER, flags=re.S)
    return sorted(set(re.findall(r"\b(spgx_[a-z0-9_]+)\s*\(", body)))


def test_library_exports_every_declared_symbol():
    from spg_b200 import api
    names = _declared()
    assert len(names) >= 30
    for n in names:
        assert hasattr(api.lib, n), f"{n} declared in include/spgx.h but not exported"


def test_ctypes_table_matches_header():
    from spg_b200 import api
    assert sorted(api.SIGNATURES) == _declared()


def test_no_torch_or_cxx_types_in_header():
    body = re.sub(r"/\*.*?\*/", "", HEADER, flags=re.S)  # comments may cite the reference's C++ names
    assert "std::" not in body and "torch" not in body and "at::" not in body
    # compiles as plain C
    subprocess.run(["gcc", "-std=c99", "-fsyntax-only", "-x", "c", str(ROOT / "include" / "spgx.h")], check=True)


def test_host_only_entry_points_work_without_gpu():
    from spg_b200 import api
    assert api.lib.spgx_kind_stencil_size(4) == 3 and api.lib.spgx_kind_num_constraints(5) == 4
    assert api.lib.spgx_kind_stencil_size(99) == -1
    assert b"sm_100a" in api.lib.spgx_version()


def test_compute_fails_loudly_without_gpu():
    """No CPU fallback: without a CUDA device the context cannot be created."""
    from spg_b200 import api
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(api.SpgxError, match="no CUDA device"):
        api.Context(0)


def test_facade_example_compiles(tmp_path):
    exe = tmp_path / "rope_facade"
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-I", str(ROOT / "include"), str(ROOT / "examples" / "rope_facade.cpp"),
                    "-L", str(ROOT / "spg_b200" / "lib"), "-lspgx", f"-Wl,-rpath,{ROOT / 'spg_b200' / 'lib'}", "-o", str(exe)],
                   check=True)
    import torch
    if not torch.cuda.is_available():
        res = subprocess.run([str(exe), "vbd", "8", "2", "1", "5"], capture_output=True, text=True)
        assert res.returncode == 1 and "no CUDA device" in res.stderr  # std::runtime_error from the facade


def test_integration_notes_name_every_entry_point():
    """INTEGRATION.md section C is the map from the C ABI to what it replaces in the reference: no declared entry point
    may be missing from it."""
    import re
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    names = set(re.findall(r"\b(spgx_[a-z0-9_]+)\s*\(", (root / "include" / "spgx.h").read_text()))
    doc = (root / "INTEGRATION.md").read_text()
    assert not [n for n in sorted(names) if n not in doc]
