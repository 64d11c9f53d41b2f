"""ctypes binding of include/spgx.h plus small Python classes mirroring the reference's solver surface
(BaseSolver: addObject / setDt / setNumSubsteps / setGravity / step / reset, spg/include/spg/sim/solver/baseSolver.h:14-41;
VBD::setIterations is the knob the reference leaves as a TODO at vbd.h:33).

No arithmetic happens here. The library is loaded eagerly and a missing library is an ImportError: there is no
CPU fallback for the product path.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

from . import scenes as _scenes

import os

# SPGX_LIB: another build of the same sources (tests/test_extra_kind.py loads the one with an extra energy kind registered)
_LIB_PATH = Path(os.environ.get("SPGX_LIB") or Path(__file__).resolve().parent / "lib" / "libspgx.so")
if not _LIB_PATH.exists():
    raise ImportError(f"{_LIB_PATH} is missing: build it with `python -m spg_b200.build` "
                      "(__graft_entry__.build()). There is no CPU fallback.")
lib = C.CDLL(str(_LIB_PATH))

SPGX_VBD, SPGX_XPBD, SPGX_XPBD_SERIAL, SPGX_NEWTON = 0, 1, 2, 3
(SPGX_IE_BARAFF_WITKIN, SPGX_IE_NEWTON_DX, SPGX_IE_NEWTON_DV, SPGX_IE_BDF2, SPGX_IE_STATIC_NEWTON, SPGX_IE_QUASI_STATIC_NEWTON,
 SPGX_IE_NEWTON_ROBUST, SPGX_IE_QUASI_STATIC_NEWTON_ROBUST) = range(8)
SPGX_HALO_NONE, SPGX_HALO_NCCL, SPGX_HALO_P2P = 0, 1, 2
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_lp = C.POINTER(C.c_int64)


class Timings(C.Structure):
    _fields_ = [("inertial_ms", C.c_double), ("solve_ms", C.c_double), ("update_ms", C.c_double),
                ("total_ms", C.c_double), ("kernel_launches", C.c_int64)]


# every symbol include/spgx.h declares, with its signature (tests/test_abi.py checks the list against the header)
SIGNATURES = {
    "spgx_last_error": (C.c_char_p, []),
    "spgx_version": (C.c_char_p, []),
    "spgx_kind_stencil_size": (C.c_int, [C.c_int]),
    "spgx_kind_num_constraints": (C.c_int, [C.c_int]),
    "spgx_kind_num_model_params": (C.c_int, [C.c_int]),
    "spgx_kind_num_record_params": (C.c_int, [C.c_int]),
    "spgx_color_vertices": (C.c_int, [C.c_int, C.c_int, C.POINTER(_ip), _lp, _ip, _ip]),
    "spgx_color_stencils": (C.c_int, [_ip, C.c_int64, C.c_int, _ip]),
    "spgx_color_vertices_device": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(_ip), _lp, _ip, _ip, C.POINTER(C.c_int)]),
    "spgx_color_vertices_noalloc": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p), _lp, _ip, C.c_void_p, C.POINTER(C.c_int)]),
    "spgx_prepare_energy": (C.c_int, [C.c_int, C.c_int, _ip, _dp, C.c_int, _dp, _dp]),
    "spgx_ctx_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "spgx_ctx_destroy": (None, [C.c_void_p]),
    "spgx_ctx_sync": (C.c_int, [C.c_void_p]),
    "spgx_comm_unique_id": (C.c_int, [C.c_void_p]),
    "spgx_ctx_init_comm": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "spgx_group_set_partition": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _ip]),
    "spgx_group_partition_info": (C.c_int, [C.c_void_p, _lp]),
    "spgx_xpbd_partition_plan": (C.c_int, [C.c_int, C.c_int, C.POINTER(_ip), _lp, _ip, _ip, C.c_int, C.c_int, _lp, _ip, _ip, _lp, _ip,
                                           _lp, _ip, _lp, _ip]),
    "spgx_partition_plan": (C.c_int, [C.c_int, C.c_int, C.POINTER(_ip), _lp, _ip, _ip, C.c_int, C.c_int, _lp, _ip, _ip, _lp, _ip,
                                      _lp, _lp]),
    "spgx_group_p2p_export": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, _lp]),
    "spgx_group_p2p_connect": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), _lp]),
    "spgx_group_p2p_disconnect": (C.c_int, [C.c_void_p]),
    "spgx_group_set_halo_mode": (C.c_int, [C.c_void_p, C.c_int]),
    "spgx_group_get_halo_mode": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "spgx_group_kernel_info": (C.c_int, [C.c_void_p, _lp]),
    "spgx_host_register": (C.c_int, [C.c_void_p, C.c_uint64]),
    "spgx_host_alloc": (C.c_int, [C.c_uint64, C.POINTER(C.c_void_p)]),
    "spgx_host_free": (C.c_int, [C.c_void_p]),
    "spgx_upload_async": (C.c_int, [C.c_void_p, _dp, _dp]),
    "spgx_download_async": (C.c_int, [C.c_void_p, _dp, _dp]),
    "spgx_transfer_wait": (C.c_int, [C.c_void_p]),
    "spgx_host_unregister": (C.c_int, [C.c_void_p]),
    "spgx_group_create": (C.c_int, [C.c_void_p, C.c_int, _dp, _dp, _dp, C.POINTER(C.c_void_p)]),
    "spgx_group_destroy": (None, [C.c_void_p]),
    "spgx_group_add_energy": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _ip, _dp, C.POINTER(C.c_int)]),
    "spgx_group_remove_energy": (C.c_int, [C.c_void_p, C.c_int]),
    "spgx_energy_update_record": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, _dp, C.c_int]),
    "spgx_energy_get_records": (C.c_int, [C.c_void_p, C.c_int, _dp]),
    "spgx_group_finalize": (C.c_int, [C.c_void_p, C.c_int]),
    "spgx_group_num_colors": (C.c_int, [C.c_void_p, C.c_int]),
    "spgx_get_vertex_colors": (C.c_int, [C.c_void_p, _ip]),
    "spgx_get_vertex_incidence": (C.c_int, [C.c_void_p, _lp, _ip, _lp]),
    "spgx_get_stencil_colors": (C.c_int, [C.c_void_p, C.c_int, _ip]),
    "spgx_download": (C.c_int, [C.c_void_p, _dp, _dp]),
    "spgx_upload": (C.c_int, [C.c_void_p, _dp, _dp]),
    "spgx_reset": (C.c_int, [C.c_void_p]),
    "spgx_scale_masses": (C.c_int, [C.c_void_p, C.c_double]),
    "spgx_vbd_step": (C.c_int, [C.c_void_p, C.c_double, C.c_int, C.c_int, _dp, C.c_int, C.c_int]),
    "spgx_xpbd_step": (C.c_int, [C.c_void_p, C.c_double, C.c_int, C.c_int, _dp, C.c_int, C.c_int]),
    "spgx_last_timings": (C.c_int, [C.c_void_p, C.POINTER(Timings)]),
    "spgx_set_graph_mode": (C.c_int, [C.c_void_p, C.c_int]),
    "spgx_eval_vertex_force_hessian": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, _dp, _dp]),
    "spgx_eval_constraints": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _dp, _dp]),
    "spgx_model_bytes": (C.c_int, [C.c_void_p, _dp, _dp]),
    "spgx_check_finite": (C.c_int, [C.c_void_p, _lp]),
    "spgx_implicit_step": (C.c_int, [C.c_void_p, C.c_int, C.c_double, C.c_int, _dp, C.c_double, C.c_int, C.c_int, C.c_int]),
    "spgx_symplectic_step": (C.c_int, [C.c_void_p, C.c_double, C.c_int, _dp, C.c_int, C.c_int]),
    "spgx_last_cg_iterations": (C.c_int, [C.c_void_p, _ip, C.c_int, C.POINTER(C.c_int)]),
    "spgx_system_force": (C.c_int, [C.c_void_p, _dp, _dp]),
    "spgx_system_stiffness": (C.c_int, [C.c_void_p, _lp, _ip, _ip, _dp]),
    "spgx_element_hessian": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _dp]),
    "spgx_linear_solve": (C.c_int, [C.c_void_p, C.c_double, C.c_double, _dp, _dp, C.c_double, C.c_int, C.POINTER(C.c_int)]),
    # rigid bodies
    "spgx_rb_kind_stencil_size": (C.c_int, [C.c_int]),
    "spgx_rb_kind_num_model_params": (C.c_int, [C.c_int]),
    "spgx_rb_group_create": (C.c_int, [C.c_void_p, C.c_int, _dp, _dp, _dp, _dp, C.POINTER(C.c_void_p)]),
    "spgx_rb_group_destroy": (None, [C.c_void_p]),
    "spgx_rb_group_add_energy": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _ip, _dp, C.POINTER(C.c_int)]),
    "spgx_rb_group_remove_energy": (C.c_int, [C.c_void_p, C.c_int]),
    "spgx_rb_energy_update_anchor": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _dp]),
    "spgx_rb_group_finalize": (C.c_int, [C.c_void_p, C.c_int]),
    "spgx_rb_step": (C.c_int, [C.c_void_p, C.c_double, C.c_int, C.c_int, _dp, C.c_int]),
    "spgx_rb_download": (C.c_int, [C.c_void_p, _dp, _dp, _dp, _dp, _dp]),
    "spgx_rb_upload": (C.c_int, [C.c_void_p, _dp, _dp, _dp, _dp]),
    "spgx_rb_reset": (C.c_int, [C.c_void_p]),
    "spgx_rb_get_body_colors": (C.c_int, [C.c_void_p, _ip]),
    "spgx_rb_get_stencil_colors": (C.c_int, [C.c_void_p, C.c_int, _ip]),
    "spgx_rb_eval_constraint": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _dp, _dp]),
    "spgx_rb_eval_vertex": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, _dp, _dp]),
}
for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)
    _fn.restype = _res
    _fn.argtypes = _args


class SpgxError(RuntimeError):
    """What the C++ facade re-throws as std::runtime_error."""

    def __init__(self, status: int, message: str):
        super().__init__(message)
        self.status = status


def check(status: int) -> None:
    if status != 0:
        raise SpgxError(status, lib.spgx_last_error().decode())


def _d(a):
    return None if a is None else a.ctypes.data_as(_dp)


def _i(a):
    return a.ctypes.data_as(_ip)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


# ---- pure host functions ---------------------------------------------------------------------------
def color_vertices(n_verts: int, stencil_sets) -> np.ndarray:
    """coloring::colorVertices (graphColoring.cpp:72-84). stencil_sets: list of [n, S] int arrays."""
    sets = [_i32(s) for s in stencil_sets]
    ptrs = (_ip * len(sets))(*[_i(s) for s in sets])
    n_entries = np.array([s.size for s in sets], dtype=np.int64)
    sizes = np.array([s.shape[1] for s in sets], dtype=np.int32)
    colors = np.empty(n_verts, dtype=np.int32)
    check(lib.spgx_color_vertices(n_verts, len(sets), ptrs, n_entries.ctypes.data_as(_lp), _i(sizes), _i(colors)))
    return colors


def color_vertices_device(ctx, n_verts: int, stencil_sets):
    """color_vertices built on the device (spg_b200/csrc/colour.cu): (colours, dependency levels)"""
    sets = [_i32(s) for s in stencil_sets]
    ptrs = (_ip * len(sets))(*[_i(s) for s in sets])
    n_entries = np.array([s.size for s in sets], dtype=np.int64)
    sizes = np.array([s.shape[1] for s in sets], dtype=np.int32)
    colors = np.empty(n_verts, dtype=np.int32)
    rounds = C.c_int(0)
    check(lib.spgx_color_vertices_device(ctx.handle, n_verts, len(sets), ptrs, n_entries.ctypes.data_as(_lp), _i(sizes), _i(colors),
                                         C.byref(rounds)))
    return colors, rounds.value


NOALLOC_MAX_VERTS = 262144     # SPGX_NOALLOC_MAX_VERTS, include/spgx.h
NOALLOC_MAX_ENTRIES = 2097152  # SPGX_NOALLOC_MAX_ENTRIES


def color_vertices_noalloc(ctx, n_verts: int, d_entries, n_entries, stencil_sizes, d_colors: int) -> int:
    """spgx_color_vertices_noalloc: the device colouring out of the library's static scratch (no cudaMalloc / cudaFree).
    d_entries: DEVICE addresses (ints) of the flat int32 stencil arrays, n_entries / stencil_sizes: their lengths and
    stencil sizes, d_colors: DEVICE address of n_verts int32. Returns the number of dependency levels."""
    ptrs = (C.c_void_p * len(d_entries))(*[int(p) for p in d_entries])
    n_entries = np.ascontiguousarray(n_entries, dtype=np.int64)
    sizes = np.ascontiguousarray(stencil_sizes, dtype=np.int32)
    rounds = C.c_int(0)
    check(lib.spgx_color_vertices_noalloc(ctx.handle, n_verts, len(d_entries), ptrs, n_entries.ctypes.data_as(_lp), _i(sizes),
                                          C.c_void_p(int(d_colors)), C.byref(rounds)))
    return rounds.value


def color_stencils(stencils) -> np.ndarray:
    """coloring::colorStencils (graphColoring.cpp:51-70). stencils: [n, S] int array."""
    s = _i32(stencils)
    colors = np.empty(s.shape[0], dtype=np.int32)
    check(lib.spgx_color_stencils(_i(s), s.size, s.shape[1], _i(colors)))
    return colors


def prepare_energy(kind: int, stencils, params, x0) -> np.ndarray:
    s, p, x0 = _i32(stencils), _f64(params), _f64(x0)
    rec = np.empty((s.shape[0], lib.spgx_kind_num_record_params(kind)), dtype=np.float64)
    check(lib.spgx_prepare_energy(kind, s.shape[0], _i(s), _d(p), x0.shape[0], _d(x0), _d(rec)))
    return rec


def comm_unique_id() -> np.ndarray:
    """ncclGetUniqueId: 128 bytes to broadcast from rank 0 to all ranks."""
    buf = np.zeros(128, dtype=np.uint8)
    check(lib.spgx_comm_unique_id(buf.ctypes.data_as(C.c_void_p)))
    return buf


def partition_plan(n_verts: int, stencil_sets, rank: int, nranks: int, owner=None) -> dict:
    """Pure host: the local view (owned / halo slots, per-colour send / receive tables) rank `rank` of `nranks` builds
    for the VBD sweep (spg_b200/csrc/partition.h). owner=None: equal contiguous index ranges."""
    sets = [_i32(s) for s in stencil_sets]
    ptrs = (_ip * len(sets))(*[_i(s) for s in sets])
    n_entries = np.array([s.size for s in sets], dtype=np.int64)
    sizes_s = np.array([s.shape[1] for s in sets], dtype=np.int32)
    own = None if owner is None else _i32(owner)
    op = None if own is None else _i(own)
    sizes = np.zeros(5, dtype=np.int64)
    lp = lambda a: a.ctypes.data_as(_lp)
    check(lib.spgx_partition_plan(n_verts, len(sets), ptrs, lp(n_entries), _i(sizes_s), op, rank, nranks, lp(sizes),
                                  None, None, None, None, None, None))
    n_owned, n_local, n_peers, n_colors, n_send = (int(a) for a in sizes)
    slot2vert = np.empty(n_local, np.int32)
    peers = np.empty(n_peers, np.int32)
    send_off = np.empty(n_colors * n_peers + 1, np.int64)
    send_verts = np.empty(n_send, np.int32)
    recv_slot = np.empty(n_colors * n_peers, np.int64)
    recv_count = np.empty(n_colors * n_peers, np.int64)
    check(lib.spgx_partition_plan(n_verts, len(sets), ptrs, lp(n_entries), _i(sizes_s), op, rank, nranks, lp(sizes),
                                  _i(slot2vert), _i(peers), lp(send_off), _i(send_verts), lp(recv_slot), lp(recv_count)))
    return {"n_owned": n_owned, "n_local": n_local, "n_colors": n_colors, "peers": peers, "slot2vert": slot2vert,
            "send_offsets": send_off, "send_verts": send_verts, "recv_slot": recv_slot, "recv_count": recv_count}


def xpbd_partition_plan(n_verts: int, stencil_sets, rank: int, nranks: int, owner=None) -> dict:
    """Pure host: what rank `rank` of `nranks` holds and exchanges when the XPBD passes are split between ranks
    (spg_b200/csrc/xpbd_partition.h). A pass is one stencil colour of one energy, numbered energy-major."""
    sets = [_i32(s) for s in stencil_sets]
    ptrs = (_ip * len(sets))(*[_i(s) for s in sets])
    n_entries = np.array([s.size for s in sets], dtype=np.int64)
    sizes_s = np.array([s.shape[1] for s in sets], dtype=np.int32)
    own = None if owner is None else _i32(owner)
    op = None if own is None else _i(own)
    sizes = np.zeros(7, dtype=np.int64)
    lp = lambda a: a.ctypes.data_as(_lp)
    check(lib.spgx_xpbd_partition_plan(n_verts, len(sets), ptrs, lp(n_entries), _i(sizes_s), op, rank, nranks, lp(sizes),
                                       None, None, None, None, None, None, None, None))
    n_owned, n_halo, n_peers, n_passes, n_held, n_send, n_recv = (int(a) for a in sizes)
    halo, peers = np.empty(n_halo, np.int32), np.empty(n_peers, np.int32)
    held_off, held = np.empty(n_passes + 1, np.int64), np.empty(n_held, np.int32)
    send_off, send = np.empty(n_passes * n_peers + 1, np.int64), np.empty(n_send, np.int32)
    recv_off, recv = np.empty(n_passes * n_peers + 1, np.int64), np.empty(n_recv, np.int32)
    check(lib.spgx_xpbd_partition_plan(n_verts, len(sets), ptrs, lp(n_entries), _i(sizes_s), op, rank, nranks, lp(sizes),
                                       _i(halo), _i(peers), lp(held_off), _i(held), lp(send_off), _i(send), lp(recv_off), _i(recv)))
    return {"n_owned": n_owned, "halo": halo, "peers": peers, "n_passes": n_passes, "held_offsets": held_off, "held": held,
            "send_offsets": send_off, "send_verts": send, "recv_offsets": recv_off, "recv_verts": recv}


def pinned_empty(shape, dtype=np.float64) -> np.ndarray:
    """A page-locked numpy array (spgx_host_alloc: cudaHostAlloc) for the end-to-end upload / download path.  The memory is
    released when the array and every view of it are gone (registering numpy's own allocation instead would leave the
    pages registered after numpy frees them, and a later allocation straddling them fails in cudaMemcpyAsync)."""
    import weakref
    dtype = np.dtype(dtype)
    n = int(np.prod(shape))
    ptr = C.c_void_p()
    check(lib.spgx_host_alloc(max(n * dtype.itemsize, 1), C.byref(ptr)))
    buf = (C.c_char * max(n * dtype.itemsize, 1)).from_address(ptr.value)
    weakref.finalize(buf, lib.spgx_host_free, C.c_void_p(ptr.value))
    return np.frombuffer(buf, dtype=dtype, count=n).reshape(shape)


# ---- device objects ----------------------------------------------------------------------------------
class Context:
    def __init__(self, device: int = 0):
        self.handle = C.c_void_p()
        check(lib.spgx_ctx_create(device, C.byref(self.handle)))
        self.device = device

    def sync(self):
        check(lib.spgx_ctx_sync(self.handle))

    def init_comm(self, rank: int, nranks: int, unique_id: np.ndarray):
        """ncclCommInitRank (collective): unique_id = comm_unique_id() of rank 0, broadcast by the caller."""
        uid = np.ascontiguousarray(unique_id, dtype=np.uint8)
        check(lib.spgx_ctx_init_comm(self.handle, rank, nranks, uid.ctypes.data_as(C.c_void_p)))

    def close(self):
        if self.handle:
            lib.spgx_ctx_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Group:
    """A solver's device-resident copy of one ParticleGroup (BaseSolver::addObject, baseSolver.cpp:10-14)."""

    def __init__(self, ctx: Context, x, x0, mass):
        self.ctx = ctx
        x, x0, mass = _f64(x), _f64(x0), _f64(mass)
        self.n = int(mass.shape[0])
        self.handle = C.c_void_p()
        self.kinds = []
        self.n_stencils = []
        check(lib.spgx_group_create(ctx.handle, self.n, _d(x), _d(x0), _d(mass), C.byref(self.handle)))

    @classmethod
    def from_scene(cls, ctx: Context, scene: "_scenes.Scene") -> "Group":
        g = cls(ctx, scene.x, scene.x0, scene.mass)
        for e in scene.energies:
            g.add_energy(e.kind, e.stencils, e.params)
        return g

    def add_energy(self, kind: int, stencils, params) -> int:
        s, p = _i32(stencils), _f64(params)
        eid = C.c_int(-1)
        n = s.size // lib.spgx_kind_stencil_size(kind)
        check(lib.spgx_group_add_energy(self.handle, kind, n, _i(s), _d(p), C.byref(eid)))
        self.kinds.append(kind)
        self.n_stencils.append(n)
        return eid.value

    def remove_energy(self, eid: int):
        """ParticleGroup::removeEnergy (particleGroup.cpp:89-95); finalize again before the next step."""
        check(lib.spgx_group_remove_energy(self.handle, eid))
        del self.kinds[eid]
        del self.n_stencils[eid]

    def set_partition(self, rank: int, nranks: int, owner=None):
        """Domain decomposition (before finalize): this rank owns the vertices with owner[v] == rank (None: equal
        contiguous index ranges)."""
        own = None if owner is None else _i32(owner)
        check(lib.spgx_group_set_partition(self.handle, rank, nranks, None if own is None else _i(own)))

    def partition_info(self) -> dict:
        o = np.zeros(8, dtype=np.int64)
        check(lib.spgx_group_partition_info(self.handle, o.ctypes.data_as(_lp)))
        return {"n_owned": int(o[0]), "n_local": int(o[1]), "n_peers": int(o[2]), "local_range": (int(o[3]), int(o[4])),
                "owned_range": (int(o[5]), int(o[6])), "n_colors": int(o[7])}

    def finalize(self, solver_kind: int):
        check(lib.spgx_group_finalize(self.handle, solver_kind))

    # ---- peer-memory halo exchange (include/spgx.h: spgx_group_p2p_*) ----
    def p2p_export(self) -> bytes:
        n = C.c_int64(0)
        check(lib.spgx_group_p2p_export(self.handle, None, 0, C.byref(n)))
        buf = C.create_string_buffer(n.value)
        check(lib.spgx_group_p2p_export(self.handle, buf, n.value, C.byref(n)))
        return buf.raw[:n.value]

    def p2p_connect(self, blobs):
        """blobs[r] = p2p_export() of rank r (all ranks, e.g. from torch.distributed.all_gather_object)."""
        keep = [C.create_string_buffer(b, len(b)) for b in blobs]
        ptrs = (C.c_void_p * len(keep))(*[C.cast(k, C.c_void_p) for k in keep])
        sizes = (C.c_int64 * len(keep))(*[len(b) for b in blobs])
        check(lib.spgx_group_p2p_connect(self.handle, len(keep), ptrs, sizes))

    def connect_peers(self, dist):
        """Collective: all-gather the export blobs over `dist` (torch.distributed) and connect; returns the halo mode."""
        blobs = [None] * dist.get_world_size()
        dist.all_gather_object(blobs, self.p2p_export())
        err = None
        try:
            self.p2p_connect(blobs)
        except SpgxError as e:  # keep the collectives below matched on every rank
            err = str(e)
        errs = [None] * dist.get_world_size()
        dist.all_gather_object(errs, err)
        if any(e is not None for e in errs):
            # every rank must run the same exchange: one failure (e.g. CUDA IPC unavailable) sends all back to NCCL
            check(lib.spgx_group_p2p_disconnect(self.handle))
            dist.barrier()
            raise SpgxError("peer-memory exchange unavailable: " + "; ".join(f"rank {r}: {e}" for r, e in enumerate(errs) if e))
        dist.barrier()
        return self.halo_mode()

    def disconnect_peers(self, dist=None):
        """Unmap the neighbours' arrays; with `dist`, bracketed by barriers so no rank frees what another still maps."""
        if dist is not None:
            dist.barrier()
        check(lib.spgx_group_p2p_disconnect(self.handle))
        if dist is not None:
            dist.barrier()

    def set_halo_mode(self, mode: int):
        check(lib.spgx_group_set_halo_mode(self.handle, int(mode)))

    def halo_mode(self) -> int:
        m = C.c_int(0)
        check(lib.spgx_group_get_halo_mode(self.handle, C.byref(m)))
        return m.value

    def kernel_info(self) -> dict:
        o = np.zeros(8, dtype=np.int64)
        check(lib.spgx_group_kernel_info(self.handle, o.ctypes.data_as(_lp)))
        return {"streamed": bool(o[0] & 0xFF), "stream_mask": int(o[0] & 0xFF), "wedge_records": int(o[0] >> 8), "lanes": int(o[1]), "slices": int(o[2]), "max_neighbours": int(o[3]),
                "halo_mode": int(o[4]), "boundary_stores_per_sweep": int(o[5]), "colouring_rounds": int(o[6]),
                "variant": int(o[7])}

    def streamed(self) -> bool:
        return self.kernel_info()["streamed"]

    def records(self, eid: int) -> np.ndarray:
        rec = np.empty((self.n_stencils[eid], lib.spgx_kind_num_record_params(self.kinds[eid])), dtype=np.float64)
        check(lib.spgx_energy_get_records(self.handle, eid, _d(rec)))
        return rec

    def update_record(self, eid: int, stencil: int, offset: int, values):
        v = _f64(values)
        check(lib.spgx_energy_update_record(self.handle, eid, stencil, offset, _d(v), v.size))

    def num_colors(self, eid: int = 0) -> int:
        return lib.spgx_group_num_colors(self.handle, eid)

    def vertex_colors(self) -> np.ndarray:
        c = np.empty(self.n, dtype=np.int32)
        check(lib.spgx_get_vertex_colors(self.handle, _i(c)))
        return c

    def vertex_incidence(self):
        n = C.c_int64(0)
        check(lib.spgx_get_vertex_incidence(self.handle, None, None, C.byref(n)))
        rowptr = np.empty(self.n * len(self.kinds) + 1, dtype=np.int64)
        entries = np.empty((n.value, 2), dtype=np.int32)
        check(lib.spgx_get_vertex_incidence(self.handle, rowptr.ctypes.data_as(_lp), _i(entries), C.byref(n)))
        return rowptr, entries

    def stencil_colors(self, eid: int) -> np.ndarray:
        c = np.empty(self.n_stencils[eid], dtype=np.int32)
        check(lib.spgx_get_stencil_colors(self.handle, eid, _i(c)))
        return c

    def download(self, want_v: bool = True):
        x = np.empty((self.n, 3), dtype=np.float64)
        v = np.empty((self.n, 3), dtype=np.float64) if want_v else None
        check(lib.spgx_download(self.handle, _d(x), _d(v)))
        return x, v

    def download_into(self, x, v=None):
        check(lib.spgx_download(self.handle, _d(x), _d(v)))

    def upload_async(self, x=None, v=None):
        """x, v: pinned_empty arrays, global-sized; returns at once (see include/spgx.h)"""
        check(lib.spgx_upload_async(self.handle, None if x is None else _d(x), None if v is None else _d(v)))

    def download_async(self, x=None, v=None):
        check(lib.spgx_download_async(self.handle, None if x is None else _d(x), None if v is None else _d(v)))

    def transfer_wait(self):
        check(lib.spgx_transfer_wait(self.handle))

    def upload(self, x=None, v=None):
        x = None if x is None else _f64(x)
        v = None if v is None else _f64(v)
        check(lib.spgx_upload(self.handle, _d(x), _d(v)))

    def reset(self):
        check(lib.spgx_reset(self.handle))

    def scale_masses(self, factor: float):
        check(lib.spgx_scale_masses(self.handle, factor))

    def set_graph_mode(self, enabled: bool):
        check(lib.spgx_set_graph_mode(self.handle, int(enabled)))

    def vbd_step(self, dt, substeps, iterations, gravity=(0.0, -9.8, 0.0), frames=1, sync=True):
        g = _f64(gravity)
        check(lib.spgx_vbd_step(self.handle, dt, substeps, iterations, _d(g), frames, int(sync)))

    def xpbd_step(self, dt, substeps, iterations=1, gravity=(0.0, -9.8, 0.0), frames=1, sync=True):
        g = _f64(gravity)
        check(lib.spgx_xpbd_step(self.handle, dt, substeps, iterations, _d(g), frames, int(sync)))

    # ---- implicit-Euler family / symplectic Euler (group finalized with SPGX_NEWTON) ----
    def implicit_step(self, kind, dt, substeps=1, gravity=(0.0, -9.8, 0.0), cg_tolerance=1e-5, cg_max_iterations=300,
                      frames=1, sync=True):
        g = _f64(gravity)
        check(lib.spgx_implicit_step(self.handle, kind, dt, substeps, _d(g), cg_tolerance, cg_max_iterations, frames, int(sync)))

    def symplectic_step(self, dt, substeps=1, gravity=(0.0, -9.8, 0.0), frames=1, sync=True):
        g = _f64(gravity)
        check(lib.spgx_symplectic_step(self.handle, dt, substeps, _d(g), frames, int(sync)))

    def cg_iterations(self) -> np.ndarray:
        n = C.c_int(0)
        check(lib.spgx_last_cg_iterations(self.handle, None, 0, C.byref(n)))
        out = np.zeros(max(n.value, 1), dtype=np.int32)
        check(lib.spgx_last_cg_iterations(self.handle, _i(out), out.size, C.byref(n)))
        return out[: n.value].copy()

    def system_force(self, gravity=(0.0, -9.8, 0.0)) -> np.ndarray:
        g = _f64(gravity)
        f = np.empty((self.n, 3))
        check(lib.spgx_system_force(self.handle, _d(g), _d(f)))
        return f

    def system_stiffness(self):
        """K = -Hessian as block-sparse rows: (rowptr [V + 1], cols [nb], blocks [nb, 3, 3])."""
        nb = C.c_int64(0)
        check(lib.spgx_system_stiffness(self.handle, C.byref(nb), None, None, None))
        rowptr, cols = np.empty(self.n + 1, dtype=np.int32), np.empty(nb.value, dtype=np.int32)
        vals = np.empty((nb.value, 3, 3))
        check(lib.spgx_system_stiffness(self.handle, C.byref(nb), _i(rowptr), _i(cols), _d(vals)))
        return rowptr, cols, vals

    def element_hessian(self, eid: int, stencil: int) -> np.ndarray:
        k = 3 * lib.spgx_kind_stencil_size(self.kinds[eid])
        h = np.empty((k, k))
        check(lib.spgx_element_hessian(self.handle, eid, stencil, _d(h)))
        return h

    def linear_solve(self, c_mass, c_hessian, b, tolerance=1e-5, max_iterations=300):
        b = _f64(b)
        x = np.empty_like(b)
        it = C.c_int(0)
        check(lib.spgx_linear_solve(self.handle, c_mass, c_hessian, _d(b), _d(x), tolerance, max_iterations, C.byref(it)))
        return x, it.value

    def timings(self) -> Timings:
        t = Timings()
        check(lib.spgx_last_timings(self.handle, C.byref(t)))
        return t

    def eval_vertex_force_hessian(self, eid: int, stencil: int, local: int, hand_derived: int = 1):
        f = np.empty(3)
        h = np.empty((3, 3))
        check(lib.spgx_eval_vertex_force_hessian(self.handle, eid, stencil, local, int(hand_derived), _d(f), _d(h)))
        return f, h

    def eval_constraints(self, eid: int, stencil: int):
        kind = self.kinds[eid]
        nc = lib.spgx_kind_num_constraints(kind)
        k = 3 * lib.spgx_kind_stencil_size(kind)
        c = np.empty(nc)
        g = np.empty((k, nc))
        check(lib.spgx_eval_constraints(self.handle, eid, stencil, _d(c), _d(g)))
        return c, g

    def check_finite(self):
        """(non-finite position components, non-finite velocity components, first offending vertex or -1), counted on the device"""
        o = np.zeros(3, np.int64)
        check(lib.spgx_check_finite(self.handle, o.ctypes.data_as(_lp)))
        return int(o[0]), int(o[1]), int(o[2])

    def model_bytes(self):
        a, b = C.c_double(0), C.c_double(0)
        check(lib.spgx_model_bytes(self.handle, C.byref(a), C.byref(b)))
        return a.value, b.value

    def close(self):
        if self.handle:
            lib.spgx_group_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


SPGX_RB_SPRING, SPGX_RB_SPRING_SQUARED, SPGX_RB_SPRING_ANCHOR, SPGX_RB_ORIENTATION_ALIGNMENT = 0, 1, 2, 3


class RbGroup:
    """A solver's device-resident copy of one RigidBodyGroup (rigidBodyGroup.h:14-111) under VBD or XPBD."""

    def __init__(self, ctx: Context, x, theta, mass, local_inertia):
        self.ctx = ctx
        x, theta, mass, inertia = _f64(x), _f64(theta), _f64(mass), _f64(local_inertia)
        self.n = int(mass.shape[0])
        self.handle = C.c_void_p()
        self.kinds = []
        check(lib.spgx_rb_group_create(ctx.handle, self.n, _d(x), _d(theta), _d(mass), _d(inertia), C.byref(self.handle)))

    @classmethod
    def from_scene(cls, ctx: Context, scene: "_scenes.RbScene") -> "RbGroup":
        g = cls(ctx, scene.x, scene.theta, scene.mass, scene.local_inertia)
        for e in scene.energies:
            g.add_energy(e.kind, e.stencils, e.params)
        return g

    def add_energy(self, kind: int, stencils, params) -> int:
        s, p = _i32(stencils), _f64(params)
        eid = C.c_int(-1)
        n = s.size // max(lib.spgx_rb_kind_stencil_size(kind), 1)
        check(lib.spgx_rb_group_add_energy(self.handle, kind, n, _i(s), _d(p), C.byref(eid)))
        self.kinds.append(kind)
        return eid.value

    def remove_energy(self, eid: int):
        check(lib.spgx_rb_group_remove_energy(self.handle, eid))
        del self.kinds[eid]

    def update_anchor(self, eid: int, stencil: int, anchor):
        a = _f64(anchor)
        check(lib.spgx_rb_energy_update_anchor(self.handle, eid, stencil, _d(a)))

    def finalize(self, solver_kind: int):
        check(lib.spgx_rb_group_finalize(self.handle, solver_kind))

    def step(self, dt, substeps, iterations=1, gravity=(0.0, -9.8, 0.0), frames=1):
        g = _f64(gravity)
        check(lib.spgx_rb_step(self.handle, dt, substeps, iterations, _d(g), frames))

    def download(self):
        """x, theta, v, omega [n, 3] and the rotation matrices [n, 3, 3]"""
        x, th, v, om = (np.empty((self.n, 3)) for _ in range(4))
        R = np.empty((self.n, 3, 3))
        check(lib.spgx_rb_download(self.handle, _d(x), _d(th), _d(v), _d(om), _d(R)))
        return x, th, v, om, R

    def upload(self, x=None, theta=None, v=None, omega=None):
        a = [None if t is None else _f64(t) for t in (x, theta, v, omega)]
        check(lib.spgx_rb_upload(self.handle, *[None if t is None else _d(t) for t in a]))

    def reset(self):
        check(lib.spgx_rb_reset(self.handle))

    def body_colors(self) -> np.ndarray:
        c = np.empty(self.n, np.int32)
        check(lib.spgx_rb_get_body_colors(self.handle, _i(c)))
        return c

    def stencil_colors(self, eid: int, n: int) -> np.ndarray:
        c = np.empty(n, np.int32)
        check(lib.spgx_rb_get_stencil_colors(self.handle, eid, _i(c)))
        return c

    def eval_constraint(self, eid: int, stencil: int):
        k = 6 * lib.spgx_rb_kind_stencil_size(self.kinds[eid])
        c, g = C.c_double(0), np.empty(k)
        check(lib.spgx_rb_eval_constraint(self.handle, eid, stencil, C.byref(c), _d(g)))
        return c.value, g

    def eval_vertex(self, eid: int, stencil: int, local: int):
        f, h = np.empty(6), np.empty((6, 6))
        check(lib.spgx_rb_eval_vertex(self.handle, eid, stencil, local, _d(f), _d(h)))
        return f, h

    def close(self):
        if self.handle:
            lib.spgx_rb_group_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class _BaseSolver:
    """BaseSolver surface (baseSolver.h:14-41) for one or more particle groups."""
    kind = SPGX_VBD

    def __init__(self, ctx: Context | None = None, device: int = 0):
        self.ctx = ctx or Context(device)
        self.groups: list[Group] = []
        self._dt, self._substeps, self._gravity = 0.01, 1, (0.0, -9.8, 0.0)  # baseSolver.h:36-39
        self._iterations = 1

    def addObject(self, scene_or_group) -> Group:
        g = scene_or_group if isinstance(scene_or_group, Group) else Group.from_scene(self.ctx, scene_or_group)
        g.finalize(self.kind)
        self.groups.append(g)
        return g

    def setDt(self, dt: float):
        self._dt = float(dt)

    def setNumSubsteps(self, n: int):
        self._substeps = int(n)

    def setGravity(self, g):
        self._gravity = tuple(float(a) for a in g)

    def setIterations(self, n: int):
        self._iterations = int(n)

    def particleGroups(self):
        return self.groups

    def reset(self):
        for g in self.groups:
            g.reset()


class VBD(_BaseSolver):
    kind = SPGX_VBD

    def step(self, frames: int = 1, sync: bool = True):
        for g in self.groups:
            g.vbd_step(self._dt, self._substeps, self._iterations, self._gravity, frames, sync)


class XPBD(_BaseSolver):
    kind = SPGX_XPBD

    def __init__(self, ctx=None, device: int = 0, useParallelGaussSeidel: bool = True):
        super().__init__(ctx, device)
        self.kind = SPGX_XPBD if useParallelGaussSeidel else SPGX_XPBD_SERIAL

    def step(self, frames: int = 1, sync: bool = True):
        for g in self.groups:
            g.xpbd_step(self._dt, self._substeps, self._iterations, self._gravity, frames, sync)


class _ImplicitEuler(_BaseSolver):
    """ImplicitEulerBase (implicitEulerBase.h:11-25): m_cgThreshold 1e-5, m_cgMaxIterations 300."""
    kind = SPGX_NEWTON
    implicit_kind = 0
    cg_threshold, cg_max_iterations = 1e-5, 300

    def step(self, frames: int = 1, sync: bool = True):
        for g in self.groups:
            g.implicit_step(self.implicit_kind, self._dt, self._substeps, self._gravity, self.cg_threshold,
                            self.cg_max_iterations, frames, sync)


class ImplicitEulerBaraffWitkin(_ImplicitEuler):
    implicit_kind = SPGX_IE_BARAFF_WITKIN


class ImplicitEulerNewtonDx(_ImplicitEuler):
    implicit_kind = SPGX_IE_NEWTON_DX


class ImplicitEulerNewtonDv(_ImplicitEuler):
    implicit_kind = SPGX_IE_NEWTON_DV


class BDF2(_ImplicitEuler):
    implicit_kind = SPGX_IE_BDF2


class StaticNewton(_ImplicitEuler):
    implicit_kind = SPGX_IE_STATIC_NEWTON


class QuasiStaticNewton(_ImplicitEuler):
    implicit_kind = SPGX_IE_QUASI_STATIC_NEWTON


class ImplicitEulerNewtonRobust(_ImplicitEuler):
    implicit_kind = SPGX_IE_NEWTON_ROBUST


class QuasiStaticNewtonRobust(_ImplicitEuler):
    implicit_kind = SPGX_IE_QUASI_STATIC_NEWTON_ROBUST


class SimplecticEuler(_BaseSolver):
    """simplecticEuler.h:12 (the reference's spelling)."""
    kind = SPGX_NEWTON

    def step(self, frames: int = 1, sync: bool = True):
        for g in self.groups:
            g.symplectic_step(self._dt, self._substeps, self._gravity, frames, sync)
