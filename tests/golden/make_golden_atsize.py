"""Generates tests/golden/atsize_c{2,3,4,5}.npz: the UNMODIFIED reference (oracle/_ref/libspgref.so, built from
/root/reference by oracle/Makefile) stepped once on the BASELINE configs AT THEIR FULL SIZE.

    python tests/golden/make_golden_atsize.py --config 3        # ~10 min on 8 cores
    python tests/golden/make_golden_atsize.py --config 2 4      # seconds / minutes
    python tests/golden/make_golden_atsize.py --config 5        # hours (16.8 M vertices), colours injected

The full state of a 1 M-vertex scene is 50 MB, so a fixture keeps
  * `colors_sha256`: hash of the reference's OWN colouring (VBD vertex colours / XPBD stencil colours per energy) as int32
    bytes -- the reference coloured the scene itself (quadratic, 40 s at 1 M) unless `injected` says the linear
    restatement's colours were injected (config 5, where the quadratic colouring is infeasible);
  * `idx`, `x`, `v`: a few thousand sampled vertices (mesh borders, the rows next to the anchors, a regular sub-grid);
  * `blk_dx`, `blk_v`: sums of (x - x_initial) and v over NBLK contiguous index blocks -- every vertex is in one, so a
    single wrong vertex anywhere shows up (a checksum of checksums; compared with an absolute tolerance).
The GPU box never needs /root/reference: tests/test_gpu_atsize.py reads these files only."""
import argparse
import hashlib
import sys
import time
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "tests")]

NBLK = 4096


def cloth_sample(n):
    """rows 0-7 and the last row, the first / last column, every 16th vertex both ways, the two pinned corners"""
    idx = set(range(8 * n)) | set(range((n - 1) * n, n * n))
    idx |= set(range(0, n * n, n)) | set(range(n - 1, n * n, n))
    g = np.arange(0, n, 16)
    idx |= set((g[:, None] * n + g[None, :]).ravel().tolist())
    return np.array(sorted(idx), dtype=np.int64)


def beam_sample(nx, ny, nz):
    """the anchored face and the two layers next to it, the free end face, every 53rd vertex"""
    layer = ny * nz
    idx = set(range(3 * layer)) | set(range((nx - 1) * layer, nx * layer)) | set(range(0, nx * layer, 53))
    return np.array(sorted(idx), dtype=np.int64)


def block_sums(a, nblk=NBLK):
    """sums over nblk contiguous index blocks (the last one ragged); float64 pairwise sums of numpy"""
    n = a.shape[0]
    edges = np.linspace(0, n, nblk + 1).astype(np.int64)
    return np.add.reduceat(a, edges[:-1], axis=0), edges


CONFIGS = {
    # name: (scene builder, solver kind (0 VBD, 1 XPBD), dt, substeps, iterations, sampler, inject colours)
    2: (lambda S: S.cloth(256, 256, S.K_SPRING, S.K_QUADRATIC_BENDING), 1, 0.01, 20, 1, lambda: cloth_sample(256), False),
    3: (lambda S: S.cloth(1024, 1024), 0, 0.01, 1, 10, lambda: cloth_sample(1024), False),
    4: (lambda S: S.tet_beam(149, 38, 38, S.K_TET_SNH), 0, 0.01, 1, 20, lambda: beam_sample(149, 38, 38), False),
    5: (lambda S: S.cloth(4096, 4096), 0, 0.01, 1, 10, lambda: cloth_sample(4096), True),
}


def make(cfg: int, threads: int | None):
    import oracle_api
    from spg_b200 import api, scenes as S
    oracle = oracle_api.load("reference")
    assert oracle.kind == "reference", "golden vectors come from the compiled reference only"
    if threads:
        oracle.set_threads(threads)
    builder, kind, dt, substeps, iterations, sampler, inject = CONFIGS[cfg]
    t0 = time.time()
    sc = builder(S)
    og = oracle_api.OracleGroup(oracle, sc)
    osol = oracle_api.OracleSolver(og, kind, dt, substeps, iterations)
    sets = [e.stencils for e in sc.energies]
    if kind == 0 and inject:
        osol.inject_vbd_colors(api.color_vertices(sc.n_vertices, sets))
    print(f"config {cfg}: {sc.n_vertices} vertices, scene {time.time() - t0:.1f} s", flush=True)
    t0 = time.time()
    osol.step(1)
    t_step = time.time() - t0
    print(f"config {cfg}: reference step {t_step:.1f} s on {oracle.max_threads()} threads", flush=True)
    x, v = osol.state()
    assert np.isfinite(x).all() and np.isfinite(v).all()
    hashes = []
    if kind == 0:
        colors, ncol = osol.vbd_colors()
        mine = api.color_vertices(sc.n_vertices, sets)
        assert np.array_equal(colors, mine), "linear-time colouring differs from the reference's at full size"
        hashes.append(hashlib.sha256(colors.astype(np.int32).tobytes()).hexdigest())
        ncolors = [ncol]
    else:
        ncolors = []
        for e in range(len(sc.energies)):
            colors, ncol = osol.xpbd_colors(e)
            mine = api.color_stencils(sc.energies[e].stencils)
            assert np.array_equal(colors, mine), f"stencil colouring of energy {e} differs from the reference's at full size"
            hashes.append(hashlib.sha256(colors.astype(np.int32).tobytes()).hexdigest())
            ncolors.append(ncol)
    idx = sampler()
    blk_dx, edges = block_sums(x - sc.x)
    blk_v, _ = block_sums(v)
    out = HERE / f"atsize_c{cfg}.npz"
    np.savez_compressed(out, idx=idx, x=x[idx], v=v[idx], blk_dx=blk_dx, blk_v=blk_v, blk_edges=edges,
                        colors_sha256=np.array(hashes), ncolors=np.array(ncolors), injected=np.array(inject),
                        x_absmax=np.abs(x).max(), v_absmax=np.abs(v).max(), n_vertices=sc.n_vertices,
                        solver=np.array([kind, substeps, iterations]), dt=dt, ref_step_s=t_step,
                        ref_threads=oracle.max_threads())
    print(f"config {cfg}: wrote {out} ({out.stat().st_size / 1e6:.2f} MB), {idx.size} sampled vertices, colours {ncolors}", flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, nargs="+", default=[2, 4, 3])
    ap.add_argument("--threads", type=int, default=0)
    a = ap.parse_args()
    for c in a.config:
        make(c, a.threads or None)
