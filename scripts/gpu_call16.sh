#!/bin/bash
# GPU call 16 (1 GPU): re-verify HEAD after the container rebuild (full parity suite, smoke, config-3 bench) + first run
# of the implicit-Euler family: parity tests and a first timing
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_implicit.py -m gpu -x -q > gpurun_out/pytest16_implicit.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest16_implicit.log
tail -25 gpurun_out/pytest16_implicit.log
timeout 1500 python -m pytest tests -m gpu -q --deselect tests/test_implicit.py > gpurun_out/pytest16.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest16.log
tail -8 gpurun_out/pytest16.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke16.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke16.log; tail -2 gpurun_out/smoke16.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/b16_c3.json 2>gpurun_out/b16_c3.err; tail -c 1500 gpurun_out/b16_c3.json
timeout 600 python - > gpurun_out/implicit16.log 2>&1 <<'PY'
import time, numpy as np
from spg_b200 import api, scenes as S
ctx = api.Context(0)
for n in (64, 256, 1024):
    sc = S.cloth(n, n)
    g = api.Group.from_scene(ctx, sc)
    t = time.time(); g.finalize(api.SPGX_NEWTON); tf = time.time() - t
    for kind in (0, 1, 2):
        g.reset()
        g.implicit_step(kind, 0.01, 1, frames=1)
        g.implicit_step(kind, 0.01, 1, frames=3)
        tm = g.timings()
        x, v = g.download()
        print(n, "kind", kind, "finalize_s", round(tf, 2), "ms/frame", round(tm.total_ms / 3, 3), "cg", g.cg_iterations(), "launches", tm.kernel_launches, "finite", bool(np.isfinite(x).all()), flush=True)
    g.close()
PY
tail -12 gpurun_out/implicit16.log
