"""Worker of tests/test_gpu_multi.py, launched with torchrun (one process per GPU): steps a partitioned scene, gathers
the owned vertices of every rank on rank 0 and writes them next to the single-domain result of rank 0's own GPU."""
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    import torch
    import torch.distributed as dist
    from spg_b200 import api, scenes as S
    scene_name, out = sys.argv[1], sys.argv[2]
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    sc = {"cloth": lambda: S.cloth(37, 21), "beam": lambda: S.tet_beam(13, 4, 3, S.K_TET_SNH),
          "cloth_edit": lambda: S.cloth(37, 21),
          # domain-decomposed XPBD (SURVEY 8e "XPBD"): the config-2 scene (springs + quadratic bending + anchors) and a beam
          "xpbd_cloth": lambda: S.cloth(37, 21, S.K_SPRING, S.K_QUADRATIC_BENDING),
          "xpbd_cloth_lambda": lambda: S.cloth(29, 17, S.K_MEMBRANE_STVK, S.K_DISCRETE_BENDING),
          "xpbd_beam": lambda: S.tet_beam(13, 4, 3, S.K_TET_SNH)}[scene_name]()
    xpbd = scene_name.startswith("xpbd")
    solver_kind = api.SPGX_XPBD if xpbd else api.SPGX_VBD

    def run(g, reconnect=None, partitioned=False):
        """the stepping sequence under test; "cloth_edit": a live edit in the middle (the picking anchor of
        apps/solverCompare.cpp:717-726 on a vertex next to a strip boundary, then re-finalize: positions AND velocities of
        owned and halo vertices must survive it -- ADVICE round 1)"""
        if xpbd:
            # reference semantics (20 substeps x 1 pass) and, for one scene, 3 iterations with lambda kept on the device
            g.xpbd_step(0.01, 20, 3 if scene_name == "xpbd_cloth_lambda" else 1, frames=2)
            return
        g.vbd_step(0.01, 2, 5, frames=3)
        if scene_name == "cloth_edit":
            picked = 21 * (37 // world) + 3   # first row of rank 1's strip
            anchor = np.concatenate([sc.x[picked] + np.array([0.0, 0.05, 0.0]), [1e4]])
            g.add_energy(S.K_ANCHOR, np.array([[picked]], np.int32), anchor[None, :])
            if partitioned:
                g.disconnect_peers(dist)   # collective: nobody re-allocates an array a neighbour still maps
            g.finalize(api.SPGX_VBD)
            if reconnect:
                reconnect()
            g.vbd_step(0.01, 2, 5, frames=2)

    ctx = api.Context(lr)
    uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        uid.copy_(torch.from_numpy(api.comm_unique_id()))
    dist.broadcast(uid, 0)
    ctx.init_comm(rank, world, uid.cpu().numpy())
    results = {}
    for mode in ("nccl", "p2p"):
        for graph in (True, False):
            g = api.Group.from_scene(ctx, sc)
            g.set_partition(rank, world)
            g.finalize(solver_kind)
            assert g.halo_mode() == api.SPGX_HALO_NCCL
            if mode == "p2p":
                assert g.connect_peers(dist) == api.SPGX_HALO_P2P
            g.set_graph_mode(graph)
            run(g, (lambda: g.connect_peers(dist)) if mode == "p2p" else None, partitioned=True)
            x = np.full((sc.n_vertices, 3), np.nan)
            v = np.full((sc.n_vertices, 3), np.nan)
            g.download_into(x, v)
            info = g.partition_info()
            lo, hi = info["owned_range"]
            assert np.isfinite(x[lo:hi]).all() and np.isnan(x[:lo]).all() and np.isnan(x[hi:]).all()
            xs, vs = np.nan_to_num(x), np.nan_to_num(v)
            tx = torch.from_numpy(xs).cuda()
            tv = torch.from_numpy(vs).cuda()
            dist.all_reduce(tx)  # every vertex is owned exactly once: the sum assembles the global state
            dist.all_reduce(tv)
            results[(mode, graph)] = (tx.cpu().numpy(), tv.cpu().numpy())
            g.disconnect_peers(dist)  # nobody frees an array a neighbour still maps or stores into
            g.close()
    if rank == 0:
        g1 = api.Group.from_scene(ctx, sc)
        g1.finalize(solver_kind)
        run(g1)
        x1, v1 = g1.download()
        np.savez(out, x1=x1, v1=v1, **{f"x_{m}_{int(gr)}": r[0] for (m, gr), r in results.items()},
                 **{f"v_{m}_{int(gr)}": r[1] for (m, gr), r in results.items()})
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
