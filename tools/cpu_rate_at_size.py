"""The reference's own CPU rate at BASELINE config 3's FULL size beside its rate at the bounded sample the bench arm times
('the CPU rate is size-independent'; VERDICT round 1 asked for the measured figure in the repo).
  python tools/cpu_rate_at_size.py > profiles/<round>_cpu_reference_rate.txt
Colours are injected at 1024^2 (the reference's O(N^2) colouring would take ~40 s there and is not part of step());
TEST INFRASTRUCTURE: loads oracle/_ref."""
import os
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import oracle_api  # noqa: E402
from spg_b200 import api, scenes as S  # noqa: E402

oracle = oracle_api.load("reference")
threads = len(os.sched_getaffinity(0))
oracle.set_threads(threads)
print(f"oracle {oracle.kind}, {threads} host threads, VBD 10 iterations, 1 substep, dt 0.01", flush=True)
for n, frames in ((64, 20), (256, 3), (1024, 1)):
    sc = S.cloth(n, n)
    og = oracle_api.OracleGroup(oracle, sc)
    osol = oracle_api.OracleSolver(og, 0, 0.01, 1, 10)
    osol.inject_vbd_colors(api.color_vertices(sc.n_vertices, [e.stencils for e in sc.energies]))
    if n < 1024:
        osol.step(1)
    t = time.perf_counter()
    osol.step(frames)
    el = time.perf_counter() - t
    print(f"cloth {n}x{n}: {frames} frame(s) in {el:.2f} s -> {sc.n_vertices * 10 * frames / el:.4g} vertex-iterations/s", flush=True)
