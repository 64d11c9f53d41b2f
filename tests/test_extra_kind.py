"""The authoring contract (north_star: "any energy written as C(x) or E(x) still plugs in unchanged"; reference:
StencilBlockEnergy<S, C>, spg/include/spg/sim/energy/energy.h:75-76, :305-347).  A 15th, throw-away kind is defined OUTSIDE
the library sources -- one EnergyKind<> specialisation (tests/native/extra_kind.cuh: C = |x - y|^2 - L0^2 as one template
body, its record layout and prepare()) + one registry line (tests/native/extra_kind.def) -- and the unchanged library sources
are compiled with it (spg_b200/lib/libspgx_extra.so).  VBD, XPBD and Newton dx then run it against the REFERENCE running
the same energy written against its own StencilBlockEnergy<2> base class (tests/native/extra_energy_ref.h)."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "spg_b200" / "lib" / "libspgx_extra.so"
REF = ROOT / "oracle" / "_ref" / "libspgref_extra.so"


def _run(mode):
    if not LIB.exists():
        pytest.fail(f"{LIB} is missing: __graft_entry__.build() builds it")
    env = dict(os.environ, SPGX_LIB=str(LIB))
    if mode == "gpu":
        env["SPGX_ORACLE_REF_LIB"] = str(REF)
    res = subprocess.run([sys.executable, str(ROOT / "tests" / "extra_kind_worker.py"), mode], capture_output=True, text=True,
                         timeout=600, env=env)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    return json.loads(res.stdout.strip().splitlines()[-1])


def test_extra_kind_registers_through_the_registry():
    out = _run("cpu")
    assert out["sizes"] == [2, 1, 2, 3]      # S, NC, NMODEL, NP straight from the EnergyKind<> traits
    assert out["beyond"] == -1
    assert out["records_ok"]
    assert "rest length" in out["validation"]  # the kind's own prepare() validation surfaces as the C ABI error message
    # the stock library knows nothing of it
    from spg_b200 import api
    assert api.lib.spgx_kind_stencil_size(14) == -1


@pytest.mark.gpu
def test_extra_kind_matches_the_reference_running_the_same_lambda():
    if not REF.exists():
        pytest.skip("oracle/_ref/libspgref_extra.so not built (needs /root/reference at build time)")
    out = _run("gpu")
    assert out["oracle"] == "reference"
    assert max(out["element"]) <= 1e-9, out["element"]
    assert out["vbd_colors_equal"]
    assert max(out["vbd"]) <= 1e-9, out["vbd"]
    assert max(out["xpbd"]) <= 1e-9, out["xpbd"]
    assert out["newton_cg"][0] == out["newton_cg"][1], out["newton_cg"]
    assert out["newton_dx"][0] <= 1e-7 and out["newton_dx"][1] <= 1e-6, out["newton_dx"]
