#!/bin/bash
mkdir -p gpurun_out
timeout 600 python - > gpurun_out/dbg27.log 2>&1 <<'PY'
import numpy as np
from spg_b200 import api, scenes as S
ctx = api.Context(0)
for n in (256, 1024):
    sc = S.cloth(n, n)
    g = api.Group.from_scene(ctx, sc)
    g.finalize(api.SPGX_NEWTON)
    for f in range(14):
        g.implicit_step(1, 0.01, 1, frames=1)
        x, v = g.download()
        fin = bool(np.isfinite(x).all())
        print(n, f, g.cg_iterations(), fin, float(np.nanmax(np.abs(x))), float(np.nanmax(np.abs(v))), flush=True)
        if not fin:
            fo = g.system_force()
            rp, cols, blk = g.system_stiffness()
            print(' force finite', bool(np.isfinite(fo).all()), 'H finite', bool(np.isfinite(blk).all()), 'bad x rows', np.argwhere(~np.isfinite(x).all(axis=1))[:10].ravel())
            break
    g.close()
PY
tail -40 gpurun_out/dbg27.log
