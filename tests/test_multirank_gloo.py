"""World-size-2 run of the halo-exchange schedule on the CPU (gloo): each rank builds its own partition plan through
the C ABI, sweeps its owned vertices colour by colour with a stand-in vertex update, and after every colour sends /
receives exactly the per-(peer, colour) ranges of the plan with torch.distributed point-to-point calls -- the same
schedule comm.cu issues with ncclSend / ncclRecv.  The gathered result must equal the single-domain sweep bit for bit."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, str(ROOT))
    import torch
    import torch.distributed as dist
    from spg_b200 import api, scenes as S
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc = S.cloth(10, 8)
    sets = [e.stencils for e in sc.energies]
    V = sc.n_vertices
    colors = api.color_vertices(V, sets)
    p = api.partition_plan(V, sets, rank, world)
    nbr = [set() for _ in range(V)]
    for st in sets:
        for row in st:
            for a in row:
                nbr[a] |= set(int(b) for b in row if b != a)
    g2l = {int(g): s for s, g in enumerate(p["slot2vert"])}
    x = torch.from_numpy(sc.x[p["slot2vert"]].copy())
    npeers = len(p["peers"])
    for it in range(2):
        for c in range(p["n_colors"]):
            for s in range(p["n_owned"]):
                v = int(p["slot2vert"][s])
                if colors[v] != c:
                    continue
                xs = torch.stack([x[g2l[u]] for u in sorted(nbr[v])])
                x[s] = 0.5 * x[s] + 0.5 * xs.mean(dim=0) + 1e-3 * torch.sin(x[s] * 3.0)
            reqs, bufs = [], []
            for ip, peer in enumerate(p["peers"]):
                q = c * npeers + ip
                sent = p["send_verts"][p["send_offsets"][q]:p["send_offsets"][q + 1]]
                if len(sent):
                    buf = torch.stack([x[g2l[int(g)]] for g in sent]).contiguous()
                    reqs.append(dist.isend(buf, int(peer)))
                cnt = int(p["recv_count"][q])
                if cnt:
                    rb = torch.empty((cnt, 3), dtype=torch.float64)
                    bufs.append((int(p["recv_slot"][q]), cnt, rb))
                    reqs.append(dist.irecv(rb, int(peer)))
            for r in reqs:
                r.wait()
            for lo, cnt, rb in bufs:
                x[lo:lo + cnt] = rb
    np.save(Path(out_dir) / f"x{rank}.npy", x[:p["n_owned"]].numpy())
    np.save(Path(out_dir) / f"v{rank}.npy", p["slot2vert"][:p["n_owned"]])
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_halo_exchange_matches_single_domain(tmp_path):
    import torch.multiprocessing as mp
    from spg_b200 import api, scenes as S
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    sc = S.cloth(10, 8)
    sets = [e.stencils for e in sc.energies]
    V = sc.n_vertices
    colors = api.color_vertices(V, sets)
    nbr = [set() for _ in range(V)]
    for st in sets:
        for row in st:
            for a in row:
                nbr[a] |= set(int(b) for b in row if b != a)
    import torch
    x = torch.from_numpy(sc.x.copy())
    for it in range(2):
        for c in range(int(colors.max()) + 1):
            for v in np.nonzero(colors == c)[0]:
                xs = torch.stack([x[u] for u in sorted(nbr[v])])
                x[v] = 0.5 * x[v] + 0.5 * xs.mean(dim=0) + 1e-3 * torch.sin(x[v] * 3.0)
    got = np.full((V, 3), np.nan)
    for r in range(world):
        got[np.load(tmp_path / f"v{r}.npy")] = np.load(tmp_path / f"x{r}.npy")
    assert np.array_equal(got, x.numpy())
