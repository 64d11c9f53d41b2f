"""spgx_color_vertices_noalloc (include/spgx.h): the device colouring for callers that may not allocate -- its scratch is a
static device array inside the library, inputs and output are the caller's device buffers.  No counterpart in the reference;
the check is against the existing entry points spgx_color_vertices_device and
spgx_color_vertices, bit for bit (integer work), on two calls in a row so that the second call finds the first call's
contents in the shared scratch.  Collected last on purpose."""
import numpy as np
import pytest

from spg_b200 import api, scenes as S

pytestmark = pytest.mark.gpu


def _device_copy(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.int32).reshape(-1)).cuda()


def _noalloc(ctx, torch, n_verts, sets):
    """colours through the C ABI out of caller-owned device buffers (torch is only the owner of that memory)"""
    d_sets = [_device_copy(torch, s) for s in sets]
    d_colors = torch.full((max(n_verts, 1),), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    rounds = api.color_vertices_noalloc(ctx, n_verts, [t.data_ptr() for t in d_sets], [s.size for s in sets], [s.shape[1] for s in sets],
                                        d_colors.data_ptr())
    return d_colors[:n_verts].cpu().numpy(), rounds


def test_noalloc_colouring_matches_existing_entry_point_twice_in_a_row(ctx):
    import torch

    big = S.cloth(96, 61, S.K_SPRING, S.K_QUADRATIC_BENDING)
    small = S.tet_beam(6, 5, 4, S.K_TET_SNH)
    # a large problem, then a smaller one over the SAME scratch (stale counts, frontiers and sizes of the first call), then
    # the large one again
    for sc in (big, small, big):
        sets = [np.ascontiguousarray(e.stencils, dtype=np.int32) for e in sc.energies]
        want, want_rounds = api.color_vertices_device(ctx, sc.n_vertices, sets)
        got, rounds = _noalloc(ctx, torch, sc.n_vertices, sets)
        assert np.array_equal(got, want)
        assert rounds == want_rounds
        assert np.array_equal(got, api.color_vertices(sc.n_vertices, sets))


def test_noalloc_colouring_edge_cases_and_limits(ctx):
    import torch

    # isolated vertices, a clique beyond one 64-colour window, an empty group
    got, _ = _noalloc(ctx, torch, 10, [np.array([[2, 7]], np.int32)])
    assert np.array_equal(got, [0, 0, 0, 0, 0, 0, 0, 1, 0, 0])
    n = 150
    i, j = np.triu_indices(n, 1)
    got, rounds = _noalloc(ctx, torch, n, [np.stack([i, j], 1).astype(np.int32)])
    assert np.array_equal(got, np.arange(n)) and rounds == n
    assert _noalloc(ctx, torch, 0, [np.zeros((0, 2), np.int32)])[0].size == 0
    # an entry outside the vertex range is reported, not followed
    with pytest.raises(api.SpgxError, match="invalid node indices"):
        _noalloc(ctx, torch, 5, [np.array([[1, 9]], np.int32)])
    # sizes beyond the static scratch are rejected before anything is launched (the pointers are never read)
    with pytest.raises(api.SpgxError, match="exceed the static scratch"):
        api.color_vertices_noalloc(ctx, api.NOALLOC_MAX_VERTS + 1, [0], [0], [2], 0)
    with pytest.raises(api.SpgxError, match="exceed the static scratch"):
        api.color_vertices_noalloc(ctx, 16, [0], [api.NOALLOC_MAX_ENTRIES + 2], [2], 0)
    # and the call after the failures is still right
    sc = S.cloth(17, 9)
    sets = [np.ascontiguousarray(e.stencils, dtype=np.int32) for e in sc.energies]
    got, _ = _noalloc(ctx, torch, sc.n_vertices, sets)
    assert np.array_equal(got, api.color_vertices(sc.n_vertices, sets))
