"""TEST INFRASTRUCTURE: ctypes binding of the oracle libraries.

Two oracles expose the same C entry points (prefix `ref_`):
  oracle/_ref/libspgref.so  the UNMODIFIED reference (alexrodag/spg @ 75122ce) compiled from /root/reference
                            by oracle/Makefile + oracle/ref_driver.cpp  -> kind "reference"
  oracle/libspgport.so      the independent CPU restatement oracle/spg_port.cpp           -> kind "port"
`load()` prefers the reference build and falls back to the port. Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this module.
"""
from __future__ import annotations

import os
import ctypes as C
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
# SPGX_ORACLE_REF_LIB: the reference + a test's own energy (oracle/Makefile target ref_extra, tests/test_extra_kind.py)
REF_LIB = Path(os.environ.get("SPGX_ORACLE_REF_LIB") or ROOT / "oracle" / "_ref" / "libspgref.so")
PORT_LIB = ROOT / "oracle" / "libspgport.so"

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_lp = C.POINTER(C.c_int64)

STENCIL_SIZE = [2, 2, 4, 4, 3, 3, 3, 4, 4, 1, 4, 2, 4, 1]
NUM_CONSTRAINTS = [1, 1, 1, 3, 3, 4, 3, 2, 2, 1, 1, 1, 2, 1]


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _d(a):
    return None if a is None else a.ctypes.data_as(_dp)


def _i(a):
    return a.ctypes.data_as(_ip)


class OracleError(RuntimeError):
    pass


class Oracle:
    def __init__(self, path: Path, kind: str):
        self.kind = kind
        self.path = path
        self.lib = C.CDLL(str(path))
        L = self.lib
        L.ref_last_error.restype = C.c_char_p
        L.ref_group_create.restype = C.c_void_p
        L.ref_group_create.argtypes = [C.c_int, _dp, _dp, _dp]
        L.ref_group_destroy.argtypes = [C.c_void_p]
        L.ref_group_add_energy.argtypes = [C.c_void_p, C.c_int, C.c_int, _ip, _dp]
        L.ref_group_get_state.argtypes = [C.c_void_p, _dp, _dp]
        L.ref_update_anchor.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp]
        L.ref_group_set_state.argtypes = [C.c_void_p, _dp, _dp]
        L.ref_group_get_masses.argtypes = [C.c_void_p, _dp, _dp]
        L.ref_energy_value.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp]
        L.ref_energy_grad_hess.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, _dp]
        L.ref_vertex_force_hessian.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, _dp, _dp]
        L.ref_constraints_and_gradient.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, _dp]
        L.ref_project_position.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double]
        L.ref_effective_stiffness.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, _dp]
        L.ref_rest_data.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, C.POINTER(C.c_int)]
        L.ref_color_vertices.argtypes = [C.c_void_p, _ip]
        L.ref_color_stencils.argtypes = [C.c_void_p, C.c_int, _ip]
        L.ref_color_vertices_flat.argtypes = [C.c_int, C.c_int, C.POINTER(_ip), _ip, _ip, _ip]
        L.ref_color_stencils_flat.argtypes = [_ip, C.c_int, C.c_int, _ip]
        L.ref_solver_create.restype = C.c_void_p
        L.ref_solver_create.argtypes = [C.c_int, C.c_void_p, C.c_double, C.c_int, C.c_int, _dp]
        L.ref_solver_destroy.argtypes = [C.c_void_p]
        L.ref_solver_step.argtypes = [C.c_void_p, C.c_int]
        L.ref_solver_reset.argtypes = [C.c_void_p]
        L.ref_solver_get_state.argtypes = [C.c_void_p, _dp, _dp]
        L.ref_solver_set_state.argtypes = [C.c_void_p, _dp, _dp]
        L.ref_vbd_vertex_colors.argtypes = [C.c_void_p, _ip, C.POINTER(C.c_int)]
        L.ref_vbd_incidence.argtypes = [C.c_void_p, _lp, _ip, _lp]
        L.ref_vbd_inject_colors.argtypes = [C.c_void_p, _ip, C.c_int]
        L.ref_xpbd_stencil_colors.argtypes = [C.c_void_p, C.c_int, _ip, C.POINTER(C.c_int)]
        L.ref_xpbd_inject_colors.argtypes = [C.c_void_p, C.c_int, _ip, C.c_int]
        if hasattr(L, "ref_rb_group_create"):  # the compiled reference only: rigid bodies
            L.ref_rb_group_create.restype = C.c_void_p
            L.ref_rb_group_create.argtypes = [C.c_int, _dp, _dp, _dp, _dp]
            L.ref_rb_group_destroy.argtypes = [C.c_void_p]
            L.ref_rb_group_add_energy.argtypes = [C.c_void_p, C.c_int, C.c_int, _ip, _dp]
            L.ref_rb_group_get_state.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp, _dp]
            L.ref_rb_group_set_state.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp]
            L.ref_rb_constraint.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, _dp]
            L.ref_rb_vertex_force_hessian.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, _dp, _dp]
            L.ref_rb_solver_create.restype = C.c_void_p
            L.ref_rb_solver_create.argtypes = [C.c_int, C.c_void_p, C.c_double, C.c_int, C.c_int, _dp]
            L.ref_rb_solver_destroy.argtypes = [C.c_void_p]
            L.ref_rb_solver_step.argtypes = [C.c_void_p, C.c_int]
            L.ref_rb_solver_get_state.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp, _dp]
            L.ref_rb_solver_set_state.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp]
            L.ref_rb_solver_update_anchor.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, _dp]
        L.ref_omp_max_threads.restype = C.c_int
        # implicit-Euler family probes (both oracles export them)
        L.ref_solver_cg_iterations.argtypes = [C.c_void_p, _ip, C.c_int, C.POINTER(C.c_int)]
        L.ref_system_force.argtypes = [C.c_void_p, _dp]
        L.ref_system_matrix.argtypes = [C.c_void_p, C.c_int, _lp, _ip, _ip, _dp]
        L.ref_linear_solve.argtypes = [C.c_void_p, C.c_int, C.c_int64, _ip, _ip, _dp, _dp, _dp, C.POINTER(C.c_int)]
        L.ref_omp_set_threads.argtypes = [C.c_int]
        # live topology edits on the solver's copy of the group (picking)
        L.ref_solver_add_energy.argtypes = [C.c_void_p, C.c_int, C.c_int, _ip, _dp]
        L.ref_solver_remove_energy.argtypes = [C.c_void_p, C.c_int]
        L.ref_solver_update_anchor.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp]
        if kind == "reference":
            L.ref_scene_cloth.restype = C.c_void_p
            L.ref_scene_cloth.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int]
            L.ref_scene_beam.restype = C.c_void_p
            L.ref_scene_beam.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int]
            L.ref_scene_destroy.argtypes = [C.c_void_p]
            for f in ("ref_scene_num_vertices", "ref_scene_num_energies"):
                getattr(L, f).argtypes = [C.c_void_p]
            for f in ("ref_scene_energy_kind", "ref_scene_energy_num_stencils"):
                getattr(L, f).argtypes = [C.c_void_p, C.c_int]
            L.ref_scene_get_vertices.argtypes = [C.c_void_p, _dp, _dp, _dp]
            L.ref_scene_get_energy.argtypes = [C.c_void_p, C.c_int, _ip, _dp]

    def check(self, status):
        if status != 0:
            raise OracleError(self.lib.ref_last_error().decode())

    def max_threads(self) -> int:
        return self.lib.ref_omp_max_threads()

    def set_threads(self, n: int):
        self.lib.ref_omp_set_threads(n)

    def color_vertices(self, n_verts, stencil_sets):
        sets = [_i32(s) for s in stencil_sets]
        ptrs = (_ip * len(sets))(*[_i(s) for s in sets])
        n_entries = np.array([s.size for s in sets], dtype=np.int32)
        sizes = np.array([s.shape[1] for s in sets], dtype=np.int32)
        colors = np.empty(n_verts, dtype=np.int32)
        self.check(self.lib.ref_color_vertices_flat(n_verts, len(sets), ptrs, _i(n_entries), _i(sizes), _i(colors)))
        return colors

    def color_stencils(self, stencils):
        s = _i32(stencils)
        colors = np.empty(s.shape[0], dtype=np.int32)
        self.check(self.lib.ref_color_stencils_flat(_i(s), s.size, s.shape[1], _i(colors)))
        return colors

    def scene(self, builder: str, *args):
        """Arrays of the reference-side scene builders (reference oracle only)."""
        from spg_b200.scenes import Scene, SceneEnergy, NUM_MODEL_PARAMS
        L = self.lib
        h = getattr(L, f"ref_scene_{builder}")(*args)
        try:
            nv, ne = L.ref_scene_num_vertices(h), L.ref_scene_num_energies(h)
            x, x0, m = np.empty((nv, 3)), np.empty((nv, 3)), np.empty(nv)
            L.ref_scene_get_vertices(h, _d(x), _d(x0), _d(m))
            sc = Scene(x, x0, m, name=builder)
            for e in range(ne):
                kind, n = L.ref_scene_energy_kind(h, e), L.ref_scene_energy_num_stencils(h, e)
                st = np.empty((n, STENCIL_SIZE[kind]), dtype=np.int32)
                p = np.empty((n, NUM_MODEL_PARAMS[kind]))
                L.ref_scene_get_energy(h, e, _i(st), _d(p))
                sc.energies.append(SceneEnergy(kind, st, p))
            return sc
        finally:
            L.ref_scene_destroy(h)


class OracleGroup:
    def __init__(self, oracle: Oracle, scene):
        self.o = oracle
        self.n = scene.n_vertices
        x, x0, m = _f64(scene.x), _f64(scene.x0), _f64(scene.mass)
        self.h = C.c_void_p(oracle.lib.ref_group_create(self.n, _d(x), _d(x0), _d(m)))
        self.kinds = []
        self.n_stencils = []
        for e in scene.energies:
            st, p = _i32(e.stencils), _f64(e.params)
            oracle.check(oracle.lib.ref_group_add_energy(self.h, e.kind, st.shape[0], _i(st), _d(p)))
            self.kinds.append(e.kind)
            self.n_stencils.append(st.shape[0])

    def close(self):
        if self.h:
            self.o.lib.ref_group_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def state(self):
        x, v = np.empty((self.n, 3)), np.empty((self.n, 3))
        self.o.lib.ref_group_get_state(self.h, _d(x), _d(v))
        return x, v

    def set_state(self, x=None, v=None):
        x = None if x is None else _f64(x)
        v = None if v is None else _f64(v)
        self.o.lib.ref_group_set_state(self.h, _d(x), _d(v))

    def update_anchor(self, e, stencil, anchor):
        a = _f64(anchor)
        self.o.check(self.o.lib.ref_update_anchor(self.h, e, stencil, _d(a)))

    def masses(self):
        m, w = np.empty(self.n), np.empty(self.n)
        self.o.lib.ref_group_get_masses(self.h, _d(m), _d(w))
        return m, w

    def energy(self, e, stencil):
        out = C.c_double(0)
        self.o.check(self.o.lib.ref_energy_value(self.h, e, stencil, C.byref(out)))
        return out.value

    def grad_hess(self, e, stencil):
        k = 3 * STENCIL_SIZE[self.kinds[e]]
        g, h = np.empty(k), np.empty((k, k))
        self.o.check(self.o.lib.ref_energy_grad_hess(self.h, e, stencil, _d(g), _d(h)))
        return g, h

    def vertex_force_hessian(self, e, stencil, local):
        f, h = np.empty(3), np.empty((3, 3))
        self.o.check(self.o.lib.ref_vertex_force_hessian(self.h, e, stencil, local, _d(f), _d(h)))
        return f, h

    def constraints(self, e, stencil):
        kind = self.kinds[e]
        k, nc = 3 * STENCIL_SIZE[kind], NUM_CONSTRAINTS[kind]
        c, g = np.empty(nc), np.empty((k, nc))
        self.o.check(self.o.lib.ref_constraints_and_gradient(self.h, e, stencil, _d(c), _d(g)))
        return c, g

    def project(self, e, stencil, dt):
        self.o.check(self.o.lib.ref_project_position(self.h, e, stencil, dt))

    def stiffness(self, e, stencil):
        nc = NUM_CONSTRAINTS[self.kinds[e]]
        k, a = np.empty((nc, nc)), np.empty((nc, nc))
        self.o.check(self.o.lib.ref_effective_stiffness(self.h, e, stencil, _d(k), _d(a)))
        return k, a

    def rest_data(self, e, stencil):
        out = np.empty(64)
        n = C.c_int(0)
        self.o.check(self.o.lib.ref_rest_data(self.h, e, stencil, _d(out), C.byref(n)))
        return out[: n.value].copy()

    def color_vertices(self):
        c = np.empty(self.n, dtype=np.int32)
        self.o.check(self.o.lib.ref_color_vertices(self.h, _i(c)))
        return c

    def color_stencils(self, e):
        c = np.empty(self.n_stencils[e], dtype=np.int32)
        self.o.check(self.o.lib.ref_color_stencils(self.h, e, _i(c)))
        return c


class OracleSolver:
    """kind: 0 = VBD, 1 = XPBD parallel Gauss-Seidel, 2 = XPBD serial, 3 = ImplicitEulerBaraffWitkin,
    4 = ImplicitEulerNewtonDx, 5 = ImplicitEulerNewtonDv, 6 = SimplecticEuler."""

    def __init__(self, group: OracleGroup, kind: int, dt=0.01, substeps=1, iterations=1, gravity=(0.0, -9.8, 0.0)):
        self.g = group
        self.o = group.o
        grav = _f64(gravity)
        self.h = C.c_void_p(self.o.lib.ref_solver_create(kind, group.h, dt, substeps, iterations, _d(grav)))

    def close(self):
        if self.h:
            self.o.lib.ref_solver_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def step(self, frames=1):
        self.o.check(self.o.lib.ref_solver_step(self.h, frames))

    def reset(self):
        self.o.check(self.o.lib.ref_solver_reset(self.h))

    def state(self):
        x, v = np.empty((self.g.n, 3)), np.empty((self.g.n, 3))
        self.o.lib.ref_solver_get_state(self.h, _d(x), _d(v))
        return x, v

    def set_state(self, x=None, v=None):
        x = None if x is None else _f64(x)
        v = None if v is None else _f64(v)
        self.o.lib.ref_solver_set_state(self.h, _d(x), _d(v))

    # ---- live topology edits (apps/solverCompare.cpp:717-726, :1446-1455): the energy joins / leaves the solver's own
    # copy of the group and the solver is flagged with requirePrecomputationUpdate ----
    def add_energy(self, kind, stencils, params):
        st, pr = _i32(stencils), _f64(params)
        self.o.check(self.o.lib.ref_solver_add_energy(self.h, kind, st.shape[0], _i(st), _d(pr)))

    def remove_energy(self, e):
        self.o.check(self.o.lib.ref_solver_remove_energy(self.h, e))

    def update_anchor(self, e, stencil, anchor):
        a = _f64(anchor)
        self.o.check(self.o.lib.ref_solver_update_anchor(self.h, e, stencil, _d(a)))

    # ---- implicit-Euler family (kinds 3-5) ----
    def cg_iterations(self):
        out = np.zeros(4096, dtype=np.int32)
        n = C.c_int(0)
        self.o.check(self.o.lib.ref_solver_cg_iterations(self.h, _i(out), out.size, C.byref(n)))
        return out[: n.value].copy()

    def system_force(self):
        f = np.empty((self.g.n, 3))
        self.o.check(self.o.lib.ref_system_force(self.h, _d(f)))
        return f

    def system_matrix(self, which=0):
        """which = 0: stiffness K = -Hessian (getSystemStiffnessMatrix), 1: mass. Returns a scipy CSR matrix."""
        import scipy.sparse as sp
        nnz = C.c_int64(0)
        self.o.check(self.o.lib.ref_system_matrix(self.h, which, C.byref(nnz), None, None, None))
        r, c, v = np.empty(nnz.value, dtype=np.int32), np.empty(nnz.value, dtype=np.int32), np.empty(nnz.value)
        self.o.check(self.o.lib.ref_system_matrix(self.h, which, C.byref(nnz), _i(r), _i(c), _d(v)))
        return sp.csr_matrix((v, (r, c)), shape=(3 * self.g.n, 3 * self.g.n))

    def linear_solve(self, A, b):
        """solveLinearSystem on a scipy sparse matrix: returns (x, CG iterations)."""
        A = A.tocoo()
        r, c, v, b = _i32(A.row), _i32(A.col), _f64(A.data), _f64(b)
        x = np.empty(b.size)
        it = C.c_int(0)
        self.o.check(self.o.lib.ref_linear_solve(self.h, b.size, v.size, _i(r), _i(c), _d(v), _d(b), _d(x), C.byref(it)))
        return x, it.value

    def vbd_colors(self):
        c = np.empty(self.g.n, dtype=np.int32)
        n = C.c_int(0)
        self.o.check(self.o.lib.ref_vbd_vertex_colors(self.h, _i(c), C.byref(n)))
        return c, n.value

    def vbd_incidence(self):
        n = C.c_int64(0)
        self.o.check(self.o.lib.ref_vbd_incidence(self.h, None, None, C.byref(n)))
        rowptr = np.empty(self.g.n * len(self.g.kinds) + 1, dtype=np.int64)
        entries = np.empty((n.value, 2), dtype=np.int32)
        self.o.check(self.o.lib.ref_vbd_incidence(self.h, rowptr.ctypes.data_as(_lp), _i(entries), C.byref(n)))
        return rowptr, entries

    def inject_vbd_colors(self, colors):
        c = _i32(colors)
        self.o.check(self.o.lib.ref_vbd_inject_colors(self.h, _i(c), c.size))

    def xpbd_colors(self, e):
        c = np.empty(self.g.n_stencils[e], dtype=np.int32)
        n = C.c_int(0)
        self.o.check(self.o.lib.ref_xpbd_stencil_colors(self.h, e, _i(c), C.byref(n)))
        return c, n.value

    def inject_xpbd_colors(self, e, colors):
        c = _i32(colors)
        self.o.check(self.o.lib.ref_xpbd_inject_colors(self.h, e, _i(c), c.size))


_cache: dict = {}


class OracleRbGroup:
    """the reference's RigidBodyGroup with its energies (compiled reference only)"""
    STENCIL = [2, 2, 1, 2]

    def __init__(self, oracle: Oracle, scene):
        if not hasattr(oracle.lib, "ref_rb_group_create"):
            raise OracleError("this oracle has no rigid bodies (the CPU restatement covers particle groups only)")
        self.o = oracle
        self.n = scene.n_bodies
        x, th, m, I = _f64(scene.x), _f64(scene.theta), _f64(scene.mass), _f64(scene.local_inertia)
        self.h = C.c_void_p(oracle.lib.ref_rb_group_create(self.n, _d(x), _d(th), _d(m), _d(I)))
        self.kinds = []
        for e in scene.energies:
            st, p = _i32(e.stencils), _f64(e.params)
            oracle.check(oracle.lib.ref_rb_group_add_energy(self.h, e.kind, st.shape[0], _i(st), _d(p)))
            self.kinds.append(e.kind)

    def state(self):
        x, th, v, om = (np.empty((self.n, 3)) for _ in range(4))
        R = np.empty((self.n, 3, 3))
        self.o.lib.ref_rb_group_get_state(self.h, _d(x), _d(th), _d(v), _d(om), _d(R))
        return x, th, v, om, R

    def set_state(self, x=None, theta=None, v=None, omega=None):
        a = [None if t is None else _f64(t) for t in (x, theta, v, omega)]
        self.o.lib.ref_rb_group_set_state(self.h, *[None if t is None else _d(t) for t in a])

    def constraint(self, e, stencil):
        c, g = np.empty(1), np.empty(6 * self.STENCIL[self.kinds[e]])
        self.o.check(self.o.lib.ref_rb_constraint(self.h, e, stencil, _d(c), _d(g)))
        return float(c[0]), g

    def vertex_force_hessian(self, e, stencil, local):
        f, h = np.empty(6), np.empty((6, 6))
        self.o.check(self.o.lib.ref_rb_vertex_force_hessian(self.h, e, stencil, local, _d(f), _d(h)))
        return f, h

    def close(self):
        if self.h:
            self.o.lib.ref_rb_group_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class OracleRbSolver:
    """kind: 0 = VBD, 1 = XPBD parallel Gauss-Seidel, 2 = XPBD serial, on a copy of the rigid-body group"""

    def __init__(self, group: OracleRbGroup, kind: int, dt=0.01, substeps=1, iterations=1, gravity=(0.0, -9.8, 0.0)):
        self.g, self.o = group, group.o
        grav = _f64(gravity)
        self.h = C.c_void_p(self.o.lib.ref_rb_solver_create(kind, group.h, dt, substeps, iterations, _d(grav)))

    def step(self, frames=1):
        self.o.check(self.o.lib.ref_rb_solver_step(self.h, frames))

    def state(self):
        x, th, v, om = (np.empty((self.g.n, 3)) for _ in range(4))
        R = np.empty((self.g.n, 3, 3))
        self.o.lib.ref_rb_solver_get_state(self.h, _d(x), _d(th), _d(v), _d(om), _d(R))
        return x, th, v, om, R

    def set_state(self, x=None, theta=None, v=None, omega=None):
        a = [None if t is None else _f64(t) for t in (x, theta, v, omega)]
        self.o.lib.ref_rb_solver_set_state(self.h, *[None if t is None else _d(t) for t in a])

    def update_anchor(self, e, stencil, anchor):
        a = _f64(anchor)
        self.o.check(self.o.lib.ref_rb_solver_update_anchor(self.h, self.g.h, e, stencil, _d(a)))

    def close(self):
        if self.h:
            self.o.lib.ref_rb_solver_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def load(prefer: str = "reference") -> Oracle:
    """The strongest oracle available: the compiled reference if present, else the CPU port."""
    order = [("reference", REF_LIB), ("port", PORT_LIB)]
    if prefer == "port":
        order.reverse()
    for kind, path in order:
        if path.exists():
            if kind not in _cache:
                _cache[kind] = Oracle(path, kind)
            return _cache[kind]
    raise OracleError("no oracle library built: run __graft_entry__.build() (make -C oracle)")


def available(kind: str) -> bool:
    return (REF_LIB if kind == "reference" else PORT_LIB).exists()
