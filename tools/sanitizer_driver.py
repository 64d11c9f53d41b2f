"""Small runs of every solver family for compute-sanitizer (tools/sanitize.sh): python tools/sanitizer_driver.py <what>
what: vbd (streaming colour pass + CSR anchors), vbd_generic (CSR colour pass), vbd_tets, xpbd, newton."""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from spg_b200 import api, scenes as S  # noqa: E402

what = sys.argv[1]
ctx = api.Context(0)
if what == "vbd":
    g = api.Group.from_scene(ctx, S.cloth(40, 33))
    g.finalize(api.SPGX_VBD)
    assert g.streamed()
    for graph in (True, False):
        g.set_graph_mode(graph)
        g.vbd_step(0.01, 2, 4, frames=2)
elif what == "vbd_generic":
    g = api.Group.from_scene(ctx, S.cloth(24, 19, S.K_MEMBRANE_CHOI, S.K_QUADRATIC_BENDING))
    g.finalize(api.SPGX_VBD)
    g.vbd_step(0.01, 1, 3, frames=2)
elif what == "vbd_tets":
    g = api.Group.from_scene(ctx, S.tet_beam(17, 5, 5, S.K_TET_SNH))
    g.finalize(api.SPGX_VBD)
    g.vbd_step(0.01, 1, 5, frames=2)
elif what == "xpbd":
    g = api.Group.from_scene(ctx, S.cloth(32, 32, S.K_SPRING, S.K_QUADRATIC_BENDING))
    g.finalize(api.SPGX_XPBD)
    g.xpbd_step(0.01, 5, 1, frames=2)
    g.xpbd_step(0.01, 2, 3, frames=1)
elif what == "newton":
    g = api.Group.from_scene(ctx, S.cloth(20, 17))
    g.finalize(api.SPGX_NEWTON)
    g.implicit_step(api.SPGX_IE_NEWTON_DX, 0.01, 1)
    g.implicit_step(api.SPGX_IE_BARAFF_WITKIN, 0.01, 1)
else:
    raise SystemExit("unknown run")
x, v = g.download()
assert np.isfinite(x).all() and np.isfinite(v).all()
assert g.check_finite() == (0, 0, -1)
print("ok", what, g.timings().kernel_launches)
