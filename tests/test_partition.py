"""Domain decomposition plan (spg_b200/csrc/partition.h through spgx_partition_plan): pure host logic, checked on the
CPU.  Properties: the owned sets partition the mesh; every stencil touching an owned vertex is fully held (owned +
halo); colours are contiguous owned-slot ranges; what rank a sends to rank b after colour c is exactly, and in the
same order, what b expects to receive from a into one contiguous halo range; and a CPU emulation of the colour-pass
schedule with per-colour halo exchange reproduces the single-domain Gauss-Seidel sweep bit for bit."""
import numpy as np
import pytest

from spg_b200 import api, scenes as S


def _plans(sc, nranks, owner=None):
    sets = [e.stencils for e in sc.energies]
    return [api.partition_plan(sc.n_vertices, sets, r, nranks, owner) for r in range(nranks)]


def _check_consistency(sc, nranks, owner=None):
    sets = [e.stencils for e in sc.energies]
    colors = api.color_vertices(sc.n_vertices, sets)
    plans = _plans(sc, nranks, owner)
    V = sc.n_vertices
    own = np.arange(V, dtype=np.int64) * nranks // V if owner is None else np.asarray(owner)
    seen = np.zeros(V, int)
    for r, p in enumerate(plans):
        owned = p["slot2vert"][:p["n_owned"]]
        halo = p["slot2vert"][p["n_owned"]:]
        seen[owned] += 1
        assert np.all(own[owned] == r) and np.all(own[halo] != r)
        # owned slots sorted by (colour, boundary first, id); halo by (owner, colour, id)
        boundary = np.zeros(V, bool)
        if len(p["send_verts"]):
            boundary[p["send_verts"]] = True
        key = (colors[owned].astype(np.int64) * 2 + (~boundary[owned])) * V + owned
        assert np.all(np.diff(key) > 0)
        hkey = (own[halo] * (colors.max() + 1) + colors[halo]) * V + halo
        assert np.all(np.diff(hkey) > 0)
        held = set(p["slot2vert"].tolist())
        for st in sets:
            touch = np.isin(st, owned).any(axis=1)
            assert set(st[touch].ravel().tolist()) <= held
        # no halo vertex that is not needed
        need = set()
        for st in sets:
            touch = np.isin(st, owned).any(axis=1)
            need |= set(st[touch].ravel().tolist())
        assert need == held or p["n_owned"] == 0
    assert np.all(seen == 1)
    # send / receive tables match pairwise
    for a, pa in enumerate(plans):
        npa = len(pa["peers"])
        for ia, b in enumerate(pa["peers"]):
            pb = plans[b]
            assert a in pb["peers"].tolist()
            ib = pb["peers"].tolist().index(a)
            npb = len(pb["peers"])
            for c in range(pa["n_colors"]):
                q = c * npa + ia
                sent = pa["send_verts"][pa["send_offsets"][q]:pa["send_offsets"][q + 1]]
                qb = c * npb + ib
                lo = pb["recv_slot"][qb]
                expected = pb["slot2vert"][lo:lo + pb["recv_count"][qb]]
                assert np.array_equal(sent, expected), (a, b, c)
                assert np.all(colors[sent] == c)
    return plans, colors


@pytest.mark.parametrize("n,nranks", [(8, 2), (13, 3), (16, 4), (9, 8)])
def test_cloth_strips(n, nranks):
    plans, _ = _check_consistency(S.cloth(n, n), nranks)
    if n >= 2 * nranks + 4:
        # row strips: at most two neighbours, halo of two rows each side
        assert max(len(p["peers"]) for p in plans) <= 2
        assert max(p["n_local"] - p["n_owned"] for p in plans) <= 4 * n


def test_tet_beam_slabs():
    sc = S.tet_beam(9, 3, 3)
    plans, _ = _check_consistency(sc, 3)
    assert max(len(p["peers"]) for p in plans) <= 2


def test_scattered_owner_and_empty_rank():
    sc = S.cloth(7, 6)
    rng = np.random.default_rng(5)
    owner = rng.integers(0, 3, sc.n_vertices).astype(np.int32)
    _check_consistency(sc, 3, owner)
    owner[:] = 1  # ranks 0 and 2 own nothing
    plans, _ = _check_consistency(sc, 3, owner)
    assert plans[0]["n_local"] == 0 and plans[1]["n_owned"] == sc.n_vertices and len(plans[1]["peers"]) == 0


def test_single_rank_is_the_unpartitioned_order():
    sc = S.cloth(10, 10)
    p = _plans(sc, 1)[0]
    colors = api.color_vertices(sc.n_vertices, [e.stencils for e in sc.energies])
    order = np.lexsort((np.arange(sc.n_vertices), colors))
    assert np.array_equal(p["slot2vert"], order) and p["n_local"] == p["n_owned"] == sc.n_vertices


def _toy_update(x, v, nbrs):
    # any deterministic vertex update that reads the current neighbour values (stands in for the 3x3 Newton step)
    return 0.5 * x[v] + 0.5 * np.mean(x[nbrs], axis=0) + 1e-3 * np.sin(x[v] * 3.0)


def test_schedule_emulation_matches_single_domain():
    """Colour passes + per-colour halo exchange on 3 emulated ranks == the single-domain coloured Gauss-Seidel sweep,
    bit for bit (every vertex update is a deterministic function of its stencil neighbours)."""
    sc = S.cloth(9, 7)
    sets = [e.stencils for e in sc.energies]
    V = sc.n_vertices
    nbr = [set() for _ in range(V)]
    for st in sets:
        for row in st:
            for a in row:
                nbr[a] |= set(int(b) for b in row if b != a)
    nbr = [np.array(sorted(s), int) for s in nbr]
    nranks = 3
    plans, colors = _check_consistency(sc, nranks)
    ncol = int(colors.max()) + 1
    x_ref = sc.x.copy()
    for it in range(3):
        for c in range(ncol):
            for v in np.nonzero(colors == c)[0]:
                x_ref[v] = _toy_update(x_ref, v, nbr[v])
    # emulated ranks: local arrays indexed by slot
    loc = [sc.x[p["slot2vert"]].copy() for p in plans]
    for it in range(3):
        for c in range(ncol):
            for r, p in enumerate(plans):
                g2l = {int(g): s for s, g in enumerate(p["slot2vert"])}
                owned = p["slot2vert"][:p["n_owned"]]
                for s, v in enumerate(owned):
                    if colors[v] != c:
                        continue
                    xs = np.stack([loc[r][g2l[int(u)]] for u in nbr[v]])
                    loc[r][s] = 0.5 * loc[r][s] + 0.5 * np.mean(xs, axis=0) + 1e-3 * np.sin(loc[r][s] * 3.0)
            # exchange of colour c
            msgs = {}
            for a, pa in enumerate(plans):
                g2l = {int(g): s for s, g in enumerate(pa["slot2vert"])}
                for ia, b in enumerate(pa["peers"]):
                    q = c * len(pa["peers"]) + ia
                    sent = pa["send_verts"][pa["send_offsets"][q]:pa["send_offsets"][q + 1]]
                    msgs[(a, int(b))] = np.stack([loc[a][g2l[int(g)]] for g in sent]) if len(sent) else np.zeros((0, 3))
            for b, pb in enumerate(plans):
                for ib, a in enumerate(pb["peers"]):
                    qb = c * len(pb["peers"]) + ib
                    lo, cnt = pb["recv_slot"][qb], pb["recv_count"][qb]
                    loc[b][lo:lo + cnt] = msgs[(int(a), b)]
    for r, p in enumerate(plans):
        owned = p["slot2vert"][:p["n_owned"]]
        assert np.array_equal(loc[r][:p["n_owned"]], x_ref[owned])
