"""Headless, vectorised restatements of the reference's scene builders (host side, build time only).

They emit flat arrays (positions, rest positions, lumped masses and, per energy in insertion order, the
stencils plus the model parameters the reference passes to addStencil) for the C ABI in include/spgx.h.

Reference: apps/solverCompare.cpp:118-189 (buildTetrahedralBeamMesh), :260-330 (createTetrahedralBeam),
:332-365 (buildTriangleSquareMesh), :367-498 (createCloth); spg/include/spg/geom/triangleMesh.h:36-50
(edge table in std::map<Int2> key order, opposing vertices in face order).  The reference computes grid
coordinates in `float` and widens them to double; that arithmetic is kept (np.float32) so x0 matches bit for
bit, and the sequential accumulation order of areas / volumes is kept (np.cumsum, np.add.at) so the lumped
masses do as well.  tests/test_scenes.py pins these against the reference driver.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

K_SPRING, K_SPRING_CONTINUUM, K_DISCRETE_BENDING, K_QUADRATIC_BENDING = 0, 1, 2, 3
K_MEMBRANE_STVK, K_MEMBRANE_CHOI, K_MEMBRANE_BW, K_TET_STVK, K_TET_SNH, K_ANCHOR = 4, 5, 6, 7, 8, 9
K_BENDING_BW, K_SPRING_SQUARED_CONTINUUM, K_TET_SSNH, K_STATIC_LOAD = 10, 11, 12, 13
KIND_NAMES = ["spring", "spring_continuum", "discrete_bending", "quadratic_bending", "membrane_stvk",
              "membrane_choi", "membrane_bw", "tet_stvk", "tet_snh", "anchor", "bending_bw", "spring_squared_continuum",
              "tet_ssnh", "static_load"]
STENCIL_SIZE = [2, 2, 4, 4, 3, 3, 3, 4, 4, 1, 4, 2, 4, 1]
NUM_MODEL_PARAMS = [2, 2, 2, 2, 2, 3, 3, 2, 2, 4, 2, 2, 2, 4]


@dataclass
class SceneEnergy:
    kind: int
    stencils: np.ndarray  # [n, S] int32
    params: np.ndarray    # [n, NM] float64


@dataclass
class Scene:
    x: np.ndarray      # [V, 3] float64 current positions
    x0: np.ndarray     # [V, 3] float64 rest positions
    mass: np.ndarray   # [V] float64
    energies: list = field(default_factory=list)
    name: str = ""

    @property
    def n_vertices(self) -> int:
        return int(self.mass.shape[0])


def _sequential_sum(a: np.ndarray) -> float:
    """Left-to-right sum (what a C++ `total += a[i]` loop computes); np.sum is pairwise."""
    return float(np.cumsum(a)[-1]) if a.size else 0.0


def triangle_square_mesh(n_rows: int, n_cols: int, width: float = 1.0, height: float = 1.0):
    """buildTriangleSquareMesh (apps/solverCompare.cpp:332-365): vertices, vertices0, faces."""
    f32 = np.float32
    row_step = f32(height) / f32(n_rows - 1)
    col_step = f32(width) / f32(n_cols - 1)
    rows, cols = np.meshgrid(np.arange(n_rows), np.arange(n_cols), indexing="ij")
    cx = (col_step * cols.astype(f32)).astype(np.float64).ravel()
    ry = (row_step * rows.astype(f32)).astype(np.float64).ravel()
    zero = np.zeros_like(cx)
    verts0 = np.stack([cx, ry, zero], axis=1)
    verts = np.stack([cx, zero, -ry], axis=1)
    r, c = np.meshgrid(np.arange(1, n_rows), np.arange(1, n_cols), indexing="ij")
    r, c = r.ravel(), c.ravel()
    a = (r - 1) * n_cols + c - 1  # (row-1, col-1)
    b = (r - 1) * n_cols + c      # (row-1, col)
    d = r * n_cols + c - 1        # (row,   col-1)
    e = r * n_cols + c            # (row,   col)
    even = ((r + c) % 2 == 0)
    f0 = np.where(even[:, None], np.stack([d, a, b], 1), np.stack([a, e, d], 1))
    f1 = np.where(even[:, None], np.stack([d, b, e], 1), np.stack([a, b, e], 1))
    faces = np.empty((2 * r.size, 3), dtype=np.int32)
    faces[0::2] = f0
    faces[1::2] = f1
    return verts, verts0, faces


def triangle_edges(faces: np.ndarray):
    """TriangleMesh edge table (triangleMesh.h:36-50): unique edges in lexicographic (min, max) order and,
    per edge, the opposing vertices in face order. Returns edges [E, 2], count [E], opp [E, 2] (-1 padded)."""
    nf = faces.shape[0]
    i0 = faces[:, [0, 1, 2]].ravel().astype(np.int64)
    i1 = faces[:, [1, 2, 0]].ravel().astype(np.int64)
    i2 = faces[:, [2, 0, 1]].ravel().astype(np.int64)
    lo, hi = np.minimum(i0, i1), np.maximum(i0, i1)
    order = np.lexsort((np.arange(3 * nf), hi, lo))  # stable: insertion order breaks ties
    lo, hi, opp = lo[order], hi[order], i2[order]
    first = np.ones(lo.size, dtype=bool)
    first[1:] = (lo[1:] != lo[:-1]) | (hi[1:] != hi[:-1])
    start = np.flatnonzero(first)
    count = np.diff(np.append(start, lo.size))
    edges = np.stack([lo[start], hi[start]], 1).astype(np.int32)
    opps = np.full((start.size, 2), -1, dtype=np.int32)
    opps[:, 0] = opp[start]
    two = count >= 2
    opps[two, 1] = opp[start[two] + 1]
    return edges, count, opps


def cloth(n_rows: int, n_cols: int, membrane: int = K_MEMBRANE_STVK, bending: int = K_DISCRETE_BENDING) -> Scene:
    """createCloth(mass=1, 1x1, n_rows x n_cols, membrane, bending) (apps/solverCompare.cpp:367-498).

    `membrane` may also be K_SPRING / K_SPRING_CONTINUUM / K_SPRING_SQUARED_CONTINUUM: a spring of stiffness 1e2 on every mesh edge with its
    rest length from the rest pose (the BASELINE config-2 scene, built through the public API as SURVEY 8d says).
    Energies in insertion order: bending, membrane, anchors (2 corner anchors, k = 1e5, at current positions)."""
    verts, verts0, faces = triangle_square_mesh(n_rows, n_cols)
    nv = verts.shape[0]
    p0, p1, p2 = verts0[faces[:, 0]], verts0[faces[:, 1]], verts0[faces[:, 2]]
    area = 0.5 * np.sqrt(np.sum(np.cross(p1 - p0, p2 - p1) ** 2, axis=1))
    # the reference's norm() is sqrt(x*x + y*y + z*z) in that order; the rest pose is planar so only z^2 is non-zero
    total_area = _sequential_sum(area)
    particle_area = np.zeros(nv)
    np.add.at(particle_area, faces.ravel(), np.repeat(area / 3.0, 3))
    density = float(np.float32(1.0)) / total_area
    mass = particle_area * density

    edges, count, opps = triangle_edges(faces)
    interior = count == 2
    bend_st = np.concatenate([edges[interior], opps[interior]], axis=1).astype(np.int32)
    bend_p = np.tile(np.array([0.0, 1e-2]), (bend_st.shape[0], 1))
    scene = Scene(verts, verts0, mass, name=f"cloth{n_rows}x{n_cols}")
    scene.energies.append(SceneEnergy(bending, bend_st, bend_p))
    if membrane in (K_SPRING, K_SPRING_CONTINUUM, K_SPRING_SQUARED_CONTINUUM):
        d = verts0[edges[:, 0]] - verts0[edges[:, 1]]
        rest = np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2])
        mem_p = np.stack([rest, np.full(rest.shape, 1e2)], 1)
        scene.energies.append(SceneEnergy(membrane, edges.copy(), mem_p))
    else:
        if membrane == K_MEMBRANE_STVK:
            mem_p = np.tile(np.array([1e2, 0.0]), (faces.shape[0], 1))
        else:
            mem_p = np.tile(np.array([1e2, 1e2, 1e2]), (faces.shape[0], 1))
        scene.energies.append(SceneEnergy(membrane, faces.copy(), mem_p))
    ids = np.array([0, n_rows - 1], dtype=np.int32)
    anchor_p = np.concatenate([verts[ids], np.full((2, 1), 1e5)], axis=1)
    scene.energies.append(SceneEnergy(K_ANCHOR, ids[:, None].copy(), anchor_p))
    return scene


def _det3_cols(u, v, w):
    """Eigen's 3x3 determinant (LU/Determinant.h:21-24) of the matrix with columns u, v, w, vectorised."""
    def h(a, b, c):
        m0 = (u[:, 0], v[:, 0], w[:, 0])
        m1 = (u[:, 1], v[:, 1], w[:, 1])
        m2 = (u[:, 2], v[:, 2], w[:, 2])
        return m0[a] * (m1[b] * m2[c] - m1[c] * m2[b])
    return h(0, 1, 2) + h(1, 2, 0) + h(2, 0, 1)


def tet_beam_mesh(nx: int, ny: int, nz: int, side=(4.0, 1.0, 1.0)):
    """buildTetrahedralBeamMesh (apps/solverCompare.cpp:118-189): vertices (== vertices0), tets."""
    f32 = np.float32
    sx, sy, sz = (f32(s) for s in side)
    i, j, k = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    half = f32(0.5)

    def coord(s, idx, n):
        # -side * 0.5f + i * side / (n - 1): float arithmetic, left to right
        return (-s * half + idx.astype(f32) * s / f32(n - 1)).astype(np.float64).ravel()

    verts = np.stack([coord(sx, i, nx), coord(sy, j, ny), coord(sz, k, nz)], axis=1)
    ci, cj, ck = np.meshgrid(np.arange(nx - 1), np.arange(ny - 1), np.arange(nz - 1), indexing="ij")
    ci, cj, ck = ci.ravel(), cj.ravel(), ck.ravel()
    ids = np.empty((ci.size, 8), dtype=np.int64)  # id1..id8 -> columns 0..7: index = di + 2*dk + 4*dj
    for dj in range(2):
        for dk in range(2):
            for di in range(2):
                ids[:, di + 2 * dk + 4 * dj] = (ci + di) * ny * nz + (cj + dj) * nz + ck + dk

    def swap(mask, pairs):
        for a, b in pairs:
            tmp = ids[mask, a].copy()
            ids[mask, a] = ids[mask, b]
            ids[mask, b] = tmp

    swap(ci % 2 == 1, [(0, 1), (2, 3), (4, 5), (6, 7)])
    swap(cj % 2 == 1, [(0, 4), (1, 5), (2, 6), (3, 7)])
    swap(ck % 2 == 1, [(0, 2), (1, 3), (4, 6), (5, 7)])
    # (id1,id3,id4,id7) (id7,id8,id4,id6) (id4,id2,id1,id6) (id1,id5,id7,id6) (id7,id4,id6,id1), 0-based columns
    combos = [(0, 2, 3, 6), (6, 7, 3, 5), (3, 1, 0, 5), (0, 4, 6, 5), (6, 3, 5, 0)]
    tets = np.empty((ci.size, 5, 4), dtype=np.int64)
    for t, cmb in enumerate(combos):
        tets[:, t, :] = ids[:, cmb]
    tets = tets.reshape(-1, 4)
    vol = tet_volumes(verts, tets)
    flip = ~(vol > 0)
    tets[flip, 0], tets[flip, 1] = tets[flip, 1].copy(), tets[flip, 0].copy()
    return verts, tets.astype(np.int32)


def tet_volumes(verts, tets):
    a = verts[tets[:, 0]]
    return _det3_cols(verts[tets[:, 1]] - a, verts[tets[:, 2]] - a, verts[tets[:, 3]] - a) / 6.0


def tet_beam(nx: int, ny: int, nz: int, kind: int = K_TET_SNH) -> Scene:
    """createTetrahedralBeam(1.f, 4, 1, 1, nx, ny, nz, 1e3f, 0.2f, kind) as createSceneObjects calls it
    (apps/solverCompare.cpp:260-330, :679-691). Energies in insertion order: anchors (x = 0 face, k = 1e4), tets."""
    verts, tets = tet_beam_mesh(nx, ny, nz)
    nv = verts.shape[0]
    vol = tet_volumes(verts, tets)
    total = _sequential_sum(vol)
    particle_vol = np.zeros(nv)
    np.add.at(particle_vol, tets.ravel(), np.repeat(vol / 4.0, 4))
    mass = particle_vol * (float(np.float32(1.0)) / total)
    scene = Scene(verts.copy(), verts.copy(), mass, name=f"beam{nx}x{ny}x{nz}")
    jj, kk = np.meshgrid(np.arange(ny), np.arange(nz), indexing="ij")
    ids = (jj * nz + kk).ravel().astype(np.int32)
    anchor_p = np.concatenate([verts[ids], np.full((ids.size, 1), 1e4)], axis=1)
    scene.energies.append(SceneEnergy(K_ANCHOR, ids[:, None].copy(), anchor_p))
    young, poisson = float(np.float32(1e3)), float(np.float32(0.2))
    scene.energies.append(SceneEnergy(kind, tets, np.tile(np.array([young, poisson]), (tets.shape[0], 1))))
    return scene


# ---- rigid-body scenes (apps/solverCompare.cpp:524-660): width = height = depth = mass = 1 (floats, :699) -------------------
RB_SPRING, RB_SPRING_SQUARED, RB_SPRING_ANCHOR, RB_ORIENTATION_ALIGNMENT = 0, 1, 2, 3


@dataclass
class RbScene:
    x: np.ndarray              # [n, 3] centres of mass
    theta: np.ndarray          # [n, 3] axis * angle
    mass: np.ndarray           # [n]
    local_inertia: np.ndarray  # [n, 3, 3] body frame
    energies: list = field(default_factory=list)
    name: str = ""

    @property
    def n_bodies(self) -> int:
        return int(self.mass.shape[0])


def _box_inertia(mass, width, height, depth):
    f32 = np.float32
    w, h, d = f32(width), f32(height), f32(depth)
    # 1.0 / 12.0 * mass * (height * height + depth * depth): the sums of squares are float arithmetic (solverCompare.cpp:541-543)
    I = np.zeros((3, 3))
    I[0, 0] = 1.0 / 12.0 * mass * float(h * h + d * d)
    I[1, 1] = 1.0 / 12.0 * mass * float(w * w + d * d)
    I[2, 2] = 1.0 / 12.0 * mass * float(w * w + h * h)
    return I


def rb_anchored(mass=1.0, width=1.0, height=1.0, depth=1.0) -> RbScene:
    """createAnchoredRigidBody (solverCompare.cpp:524-555): one box held by three anchor springs"""
    w, h = float(np.float32(width)), float(np.float32(height))
    sc = RbScene(np.zeros((1, 3)), np.zeros((1, 3)), np.array([mass]), _box_inertia(mass, width, height, depth)[None], name="anchored rigid body")
    params = np.array([[0, h * 0.5, 0, 0, h * 0.5, -1, 100], [w * 0.5, -h * 0.5, 0, w, -h * 0.5, 1, 100],
                       [-w * 0.5, -h * 0.5, 0, -w, -h * 0.5, 1, 100]], dtype=np.float64)
    sc.energies.append(SceneEnergy(RB_SPRING_ANCHOR, np.zeros((3, 1), np.int32), params))
    return sc


def rb_articulated(n_bodies: int, mass=1.0, width=1.0, height=1.0, depth=1.0) -> RbScene:
    """createArticulatedRigidBody (solverCompare.cpp:557-603): a row of boxes joined by squared springs of rest length 0
    and orientation alignment, the first one pinned by three stiff anchors"""
    w, h = float(np.float32(width)), float(np.float32(height))
    x = np.zeros((n_bodies, 3))
    x[:, 0] = w * np.arange(n_bodies)
    I = np.repeat(_box_inertia(mass, width, height, depth)[None], n_bodies, 0)
    sc = RbScene(x, np.zeros((n_bodies, 3)), np.full(n_bodies, mass), I, name="articulated rigid body")
    anchors = np.array([[0, h * 0.5, 0, 0, h * 0.5, 0, 1e10], [0, -h * 0.5, 0, 0, -h * 0.5, 0, 1e10],
                        [-w * 0.5, 0, 0, -w * 0.5, 0, 0, 1e10]], dtype=np.float64)
    sc.energies.append(SceneEnergy(RB_SPRING_ANCHOR, np.zeros((3, 1), np.int32), anchors))
    if n_bodies > 1:
        st = np.stack([np.arange(n_bodies - 1), np.arange(1, n_bodies)], 1).astype(np.int32)
        sp = np.tile(np.array([w * 0.5, 0, 0, -w * 0.5, 0, 0, 0, 1e10]), (n_bodies - 1, 1))
        sc.energies.append(SceneEnergy(RB_SPRING_SQUARED, st, sp))
        sc.energies.append(SceneEnergy(RB_ORIENTATION_ALIGNMENT, st.copy(), np.full((n_bodies - 1, 1), 1e5)))
    return sc


def rb_chain(n_bodies: int, mass=1.0, width=1.0, height=1.0, depth=1.0) -> RbScene:
    """createRigidBodyChain (solverCompare.cpp:605-660): boxes along a diagonal joined corner to corner by springs of rest
    length sqrt(2 width) (a FLOAT square root, :654), the first one hung from an anchor"""
    w, h = float(np.float32(width)), float(np.float32(height))
    x = np.zeros((n_bodies, 3))
    step = float(np.float32(height) + np.float32(1))  # -(height + 1) * i: float arithmetic
    x[:, 0] = -step * np.arange(n_bodies)
    x[:, 2] = -step * np.arange(n_bodies)
    I = np.repeat(_box_inertia(mass, width, height, depth)[None], n_bodies, 0)
    sc = RbScene(x, np.zeros((n_bodies, 3)), np.full(n_bodies, mass), I, name="rigid body chain")
    sc.energies.append(SceneEnergy(RB_SPRING_ANCHOR, np.zeros((1, 1), np.int32), np.array([[w * 0.5, h * 0.5, 0, w * 0.5, h * 0.5, 0, 100.0]])))
    if n_bodies > 1:
        st = np.stack([np.arange(n_bodies - 1), np.arange(1, n_bodies)], 1).astype(np.int32)
        rest = float(np.sqrt(np.float32(width) * np.float32(2)))
        sp = np.tile(np.array([-w * 0.5, -h * 0.5, 0, w * 0.5, h * 0.5, 0, rest, 100.0]), (n_bodies - 1, 1))
        sc.energies.append(SceneEnergy(RB_SPRING, st, sp))
    return sc
