"""Worker of tests/test_extra_kind.py.  Runs with SPGX_LIB = spg_b200/lib/libspgx_extra.so (the library sources + the
throw-away kind of tests/native/extra_kind.{def,cuh}) and, for the GPU part, SPGX_ORACLE_REF_LIB = oracle/_ref/
libspgref_extra.so (the reference + the same energy written against StencilBlockEnergy<2>, tests/native/extra_energy_ref.h).
Prints one JSON line."""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
sys.path.insert(0, str(ROOT / "tests" / "golden"))

K_TEST = 14


def rope_scene(S, n=24):
    """a zig-zag rope hanging from an anchored end: the new energy on every segment + the reference's anchor"""
    t = np.arange(n, dtype=np.float64)
    x = np.stack([0.05 * t, 0.01 * np.sin(1.7 * t), 0.02 * np.cos(0.9 * t)], axis=1)
    x0 = x.copy()
    mass = np.full(n, 1.0 / n)
    seg = np.stack([np.arange(n - 1), np.arange(1, n)], axis=1).astype(np.int32)
    rest = 0.97 * np.linalg.norm(x[seg[:, 0]] - x[seg[:, 1]], axis=1)  # slightly pre-stretched
    params = np.stack([rest, np.full(n - 1, 40.0)], axis=1)
    sc = S.Scene(x=x, x0=x0, mass=mass, name="rope with the throw-away kind")
    sc.energies.append(S.SceneEnergy(K_TEST, seg, params))
    sc.energies.append(S.SceneEnergy(S.K_ANCHOR, np.array([[0]], np.int32), np.array([[*x[0], 1e4]])))
    return sc


def main():
    mode = sys.argv[1]
    from spg_b200 import api, scenes as S
    out = {"sizes": [api.lib.spgx_kind_stencil_size(K_TEST), api.lib.spgx_kind_num_constraints(K_TEST),
                     api.lib.spgx_kind_num_model_params(K_TEST), api.lib.spgx_kind_num_record_params(K_TEST)],
           "beyond": api.lib.spgx_kind_stencil_size(K_TEST + 1)}
    sc = rope_scene(S)
    rec = api.prepare_energy(K_TEST, sc.energies[0].stencils, sc.energies[0].params, sc.x0)
    out["records_ok"] = bool(np.array_equal(rec[:, 0], sc.energies[0].params[:, 0]) and np.array_equal(rec[:, 1], sc.energies[0].params[:, 1])
                             and np.array_equal(rec[:, 2], 1.0 / sc.energies[0].params[:, 1]))
    try:
        bad = sc.energies[0].params.copy()
        bad[3, 0] = -1.0
        api.prepare_energy(K_TEST, sc.energies[0].stencils, bad, sc.x0)
        out["validation"] = "no error"
    except api.SpgxError as e:
        out["validation"] = str(e)
    if mode == "gpu":
        import oracle_api
        oracle_api.STENCIL_SIZE.append(2)
        oracle_api.NUM_CONSTRAINTS.append(1)
        oracle = oracle_api.load()
        out["oracle"] = oracle.kind
        ctx = api.Context(0)
        rel = lambda a, b: float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))
        # per element: constraint value + Jacobian (XPBD), owner-vertex force + 3x3 block (VBD)
        g = api.Group.from_scene(ctx, sc)
        g.finalize(api.SPGX_XPBD)
        og = oracle_api.OracleGroup(oracle, sc)
        ec = ej = ef = eh = 0.0
        for st in (0, 7, 22):
            c, j = g.eval_constraints(0, st)
            cr, jr = og.constraints(0, st)
            ec, ej = max(ec, rel(c, cr)), max(ej, rel(j, jr))
        g.close()
        g = api.Group.from_scene(ctx, sc)
        g.finalize(api.SPGX_VBD)
        for st in (0, 7, 22):
            for local in (0, 1):
                f, h = g.eval_vertex_force_hessian(0, st, local, 0)
                fr, hr = og.vertex_force_hessian(0, st, local)
                ef, eh = max(ef, rel(f, fr)), max(eh, rel(h, hr))
        out["element"] = [ec, ej, ef, eh]
        # whole steps: VBD 10 iterations; XPBD 20 substeps x 1 pass (reference semantics); Newton dx
        g.vbd_step(0.01, 1, 10, frames=2)
        x, v = g.download()
        osol = oracle_api.OracleSolver(og, 0, 0.01, 1, 10)
        out["vbd_colors_equal"] = bool(np.array_equal(g.vertex_colors(), osol.vbd_colors()[0]))
        osol.step(2)
        xr, vr = osol.state()
        out["vbd"] = [rel(x, xr), rel(v, vr)]
        g.close()
        g = api.Group.from_scene(ctx, sc)
        g.finalize(api.SPGX_XPBD)
        g.xpbd_step(0.01, 20, 1, frames=2)
        x, v = g.download()
        og2 = oracle_api.OracleGroup(oracle, sc)
        osol2 = oracle_api.OracleSolver(og2, 1, 0.01, 20, 1)
        osol2.step(2)
        xr, vr = osol2.state()
        out["xpbd"] = [rel(x, xr), rel(v, vr)]
        g.close()
        g = api.Group.from_scene(ctx, sc)
        g.finalize(api.SPGX_NEWTON)
        g.implicit_step(api.SPGX_IE_NEWTON_DX, 0.01, 1)
        x, v = g.download()
        og3 = oracle_api.OracleGroup(oracle, sc)
        import make_golden_implicit as MI
        osol3 = oracle_api.OracleSolver(og3, MI.NDX, 0.01, 1, 1)
        osol3.step(1)
        xr, vr = osol3.state()
        out["newton_dx"] = [rel(x, xr), rel(v, vr)]
        out["newton_cg"] = [g.cg_iterations().tolist(), osol3.cg_iterations().tolist()]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
