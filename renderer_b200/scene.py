"""Synthetic scenes for the harness (tests, smoke, bench): host-side input
builders only, no compute on the product path.

The entity generator is counter-based -- entity i depends only on (seed, i) --
so a GPU shard can regenerate its slice on the device (rcull_synth_entities)
and the CPU oracle can regenerate the same slice (oracle_synth_entities); all
three are bit-identical because every float step is a single IEEE-rounded
+, -, *, / or sqrt on values derived from 24-bit integers.

Scene recipes follow SURVEY.md section 8(d): seed 0xC0FFEE, M = 16 synthetic
meshes, default camera (camera_controller.rs:21-24), the two reference lights
(src/main.rs:365-384) plus two mirrored ones for the 5-view configuration.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import MESH_DTYPE, View

SEED = 0xC0FFEE
N_MESHES = 16

_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)
_GOLD = np.uint64(0x9E3779B97F4A7C15)
_F = np.float32


def _mix64(z: np.ndarray) -> np.ndarray:
    z = (z ^ (z >> np.uint64(30))) * _M1
    z = (z ^ (z >> np.uint64(27))) * _M2
    return z ^ (z >> np.uint64(31))


def _chan(seed: int, i: np.ndarray, c: int) -> np.ndarray:
    with np.errstate(over="ignore"):
        return _mix64(np.uint64(seed) + (i * np.uint64(16) + np.uint64(c + 1)) * _GOLD)


def _unit24(h: np.ndarray) -> np.ndarray:
    return (h >> np.uint64(40)).astype(np.float32) * _F(2.0**-24)


def synth_entities(seed: int, first: int, n: int, half_box: float, n_meshes: int = N_MESHES):
    """Returns SoA columns (pos_scale [n,4], rot [n,4] as i,j,k,w, mesh [n] u32)."""
    i = np.arange(first, first + n, dtype=np.uint64)
    hb = _F(half_box)
    ps = np.empty((n, 4), np.float32)
    for a in range(3):
        ps[:, a] = (_unit24(_chan(seed, i, a)) * _F(2) - _F(1)) * hb
    ps[:, 3] = _unit24(_chan(seed, i, 3)) * _F(3.75) + _F(0.25)
    c = np.stack([_unit24(_chan(seed, i, 4 + a)) * _F(2) - _F(1) for a in range(4)], axis=1)
    n2 = ((c[:, 0] * c[:, 0] + c[:, 1] * c[:, 1]) + c[:, 2] * c[:, 2]) + c[:, 3] * c[:, 3]
    small = n2 < _F(1e-4)
    nn = np.sqrt(np.where(small, _F(1), n2)).astype(np.float32)
    c = (c / nn[:, None]).astype(np.float32)
    c[small] = np.array([0, 0, 0, 1], np.float32)
    # UnitQuaternion::new_normalize (scene_loader.rs:590)
    nn = np.sqrt((c[:, 0] * c[:, 0] + c[:, 2] * c[:, 2]) + (c[:, 1] * c[:, 1] + c[:, 3] * c[:, 3])).astype(np.float32)
    rot = (c / nn[:, None]).astype(np.float32)
    mesh = (((_chan(seed, i, 8) >> np.uint64(32)) * np.uint64(n_meshes)) >> np.uint64(32)).astype(np.uint32)
    return ps, rot, mesh


def synth_mesh_table(n_meshes: int = N_MESHES, seed: int = SEED) -> np.ndarray:
    """M synthetic meshes: AABB half-sizes in [0.2, 3] around small offsets, 1-3 LODs,
    index_len a multiple of 3 in [36, 30000]; offsets are running sums like
    consolidate_mesh_buffers.rs:100-159."""
    rng = np.random.Generator(np.random.PCG64(seed))
    t = np.zeros(n_meshes, MESH_DTYPE)
    next_index = 0
    next_vertex = 0
    for m in range(n_meshes):
        half = rng.uniform(0.2, 3.0, 3).astype(np.float32)
        off = rng.uniform(-0.5, 0.5, 3).astype(np.float32)
        t[m]["aabb_min"] = off - half
        t[m]["aabb_max"] = off + half
        n_lods = int(rng.integers(1, 4))
        t[m]["n_lods"] = n_lods
        length = int(rng.integers(12, 10001)) * 3
        for l in range(n_lods):
            t[m]["lod"][l] = (next_index, length)
            next_index += length
            length = max(36, (length // 2) // 3 * 3)
        t[m]["vertex_offset"] = next_vertex
        next_vertex += int(rng.integers(24, 20000))
    return t


def synth_geometry(meshes: np.ndarray, seed: int = SEED, last_vertices: int = 4096, coherent: bool = False):
    """Consolidated position / index buffers for a synthetic mesh table: mesh m owns vertices
    [vertex_offset[m], vertex_offset[m+1]) inside its AABB, and every LOD's index range
    [first_index, first_index + index_len) holds triangles over them (indices relative to
    vertex_offset, as generate_work.comp:122 adds it back).
    coherent=False: uniformly random positions and random index triples -- no vertex reuse between
    neighbouring triangles, the worst case for the position gathers.
    coherent=True: a jittered W x H grid sheet folded through the AABB, triangles in row-major cell order
    (two per cell, alternating winding by row, wrapped to fill index_len) -- consecutive triangles share
    vertices the way an indexed, cache-optimised glTF mesh does."""
    rng = np.random.Generator(np.random.PCG64(seed ^ 0x9E3779B9))
    voff = meshes["vertex_offset"].astype(np.int64)
    n_vert = np.append(voff[1:], voff[-1] + last_vertices) - voff
    positions = np.empty((int(voff[-1] + last_vertices), 3), np.float32)
    n_idx = int(max(int(m["lod"][l][0]) + int(m["lod"][l][1]) for m in meshes for l in range(int(m["n_lods"]))))
    indices = np.zeros(n_idx, np.uint32)
    for m, mesh in enumerate(meshes):
        lo, hi = mesh["aabb_min"].astype(np.float64), mesh["aabb_max"].astype(np.float64)
        nv = int(n_vert[m])
        if not coherent:
            positions[voff[m] : voff[m] + nv] = (lo + rng.random((nv, 3)) * (hi - lo)).astype(np.float32)
            for l in range(int(mesh["n_lods"])):
                first, length = int(mesh["lod"][l][0]), int(mesh["lod"][l][1])
                indices[first : first + length] = rng.integers(0, nv, length, dtype=np.uint32)
            continue
        w = max(2, int(np.sqrt(nv)))
        h = max(2, nv // w)
        u, v = np.meshgrid(np.linspace(0.0, 1.0, w), np.linspace(0.0, 1.0, h))
        u = np.clip(u + (rng.random(u.shape) - 0.5) * 0.3 / w, 0.0, 1.0)
        v = np.clip(v + (rng.random(v.shape) - 0.5) * 0.3 / h, 0.0, 1.0)
        z = 0.5 + 0.5 * np.sin(6.0 * u + m) * np.cos(5.0 * v)
        grid = np.stack([u.ravel(), v.ravel(), z.ravel()], axis=1)
        pts = np.empty((nv, 3))
        pts[: w * h] = grid
        pts[w * h :] = grid[: nv - w * h]  # vertices the grid does not use repeat its first ones
        positions[voff[m] : voff[m] + nv] = (lo + pts * (hi - lo)).astype(np.float32)
        a = (np.arange(h - 1)[:, None] * w + np.arange(w - 1)[None, :])
        flip = (np.arange(h - 1)[:, None] % 2 == 1)  # alternate the winding by row: both facings occur
        t0 = np.where(flip[..., None], np.stack([a, a + w, a + 1], -1), np.stack([a, a + 1, a + w], -1))
        t1 = np.where(flip[..., None], np.stack([a + 1, a + w, a + w + 1], -1), np.stack([a + 1, a + w + 1, a + w], -1))
        cells = np.stack([t0, t1], axis=2).reshape(-1, 3).astype(np.uint32)  # row-major cells, two triangles each
        for l in range(int(mesh["n_lods"])):
            first, length = int(mesh["lod"][l][0]), int(mesh["lod"][l][1])
            stride = 1 << l  # coarser LODs skip cells
            tri = cells[(np.arange(length // 3) * stride) % len(cells)]
            indices[first : first + length] = tri.reshape(-1)[:length]
    return positions, indices


def to_components(pos_scale: np.ndarray, rot: np.ndarray):
    """SoA columns -> the reference's AoS components (Position, Rotation, Scale)."""
    return (
        np.ascontiguousarray(pos_scale[:, :3]),
        np.ascontiguousarray(rot),
        np.ascontiguousarray(pos_scale[:, 3]),
    )


# ---- views (host side of row A5, computed by the library's host helpers) -----------
CAMERA_POS = np.array([0.0, 1.0, 2.0], np.float32)  # camera_controller.rs:21
CAMERA_ROT = np.array([0.0, 0.0, 0.0, 1.0], np.float32)  # identity
LIGHT_POS = np.array(
    [
        [30.0, 20.0, -40.1],  # src/main.rs:368
        [0.1, 17.0, 0.1],  # src/main.rs:377
        [-30.0, 20.0, 40.1],  # mirrored (config 3 needs 4 shadow views)
        [-0.1, 17.0, -0.1],
    ],
    np.float32,
)


def _fp(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def camera_view(aspect: float = 16.0 / 9.0, pos=CAMERA_POS, rot=CAMERA_ROT) -> np.ndarray:
    lib = _lib.load()
    v = View()
    pos = np.ascontiguousarray(pos, np.float32)
    rot = np.ascontiguousarray(rot, np.float32)
    lib.rcull_project_camera(aspect, _fp(pos), _fp(rot), None, None, None, C.byref(v))
    return np.ctypeslib.as_array(v.planes).copy().astype(np.float32)


def camera_pv(aspect: float = 16.0 / 9.0, pos=CAMERA_POS, rot=CAMERA_ROT) -> np.ndarray:
    """projection * view of the camera, column-major (CameraMatrices.pv, generate_work.comp:33)."""
    lib = _lib.load()
    v = View()
    pos = np.ascontiguousarray(pos, np.float32)
    rot = np.ascontiguousarray(rot, np.float32)
    pv = np.zeros(16, np.float32)
    lib.rcull_project_camera(aspect, _fp(pos), _fp(rot), None, None, _fp(pv), C.byref(v))
    return pv


def quat_look_at_lh(direction, up=(0.0, 1.0, 0.0)) -> np.ndarray:
    """UnitQuaternion::look_at_lh(dir, up) (main.rs:369-372), via the basis rows."""
    z = np.asarray(direction, np.float64)
    z = z / np.linalg.norm(z)
    x = np.cross(np.asarray(up, np.float64), z)
    x = x / np.linalg.norm(x)
    y = np.cross(z, x)
    m = np.stack([x, y, z])  # world -> view rotation
    tr = m[0, 0] + m[1, 1] + m[2, 2]
    if tr > 0:
        s = np.sqrt(tr + 1.0) * 2
        q = [(m[2, 1] - m[1, 2]) / s, (m[0, 2] - m[2, 0]) / s, (m[1, 0] - m[0, 1]) / s, 0.25 * s]
    elif m[0, 0] > m[1, 1] and m[0, 0] > m[2, 2]:
        s = np.sqrt(1.0 + m[0, 0] - m[1, 1] - m[2, 2]) * 2
        q = [0.25 * s, (m[0, 1] + m[1, 0]) / s, (m[0, 2] + m[2, 0]) / s, (m[2, 1] - m[1, 2]) / s]
    elif m[1, 1] > m[2, 2]:
        s = np.sqrt(1.0 + m[1, 1] - m[0, 0] - m[2, 2]) * 2
        q = [(m[0, 1] + m[1, 0]) / s, 0.25 * s, (m[1, 2] + m[2, 1]) / s, (m[0, 2] - m[2, 0]) / s]
    else:
        s = np.sqrt(1.0 + m[2, 2] - m[0, 0] - m[1, 1]) * 2
        q = [(m[0, 2] + m[2, 0]) / s, (m[1, 2] + m[2, 1]) / s, 0.25 * s, (m[1, 0] - m[0, 1]) / s]
    q = np.asarray(q, np.float64)
    return (q / np.linalg.norm(q)).astype(np.float32)


def light_view(pos) -> np.ndarray:
    lib = _lib.load()
    pos = np.ascontiguousarray(pos, np.float32)
    rot = quat_look_at_lh(-pos.astype(np.float64))  # look at the origin (main.rs:369-371)
    v = View()
    lib.rcull_light_view(_fp(pos), _fp(rot), None, None, None, C.byref(v))
    return np.ctypeslib.as_array(v.planes).copy().astype(np.float32)


def scene_views(n_views: int = 1) -> np.ndarray:
    """views[0] = default camera; views[1..] = shadow views of LIGHT_POS."""
    vs = [camera_view()]
    for k in range(n_views - 1):
        vs.append(light_view(LIGHT_POS[k % len(LIGHT_POS)] * (1.0 + 0.25 * (k // len(LIGHT_POS)))))
    return np.stack(vs).astype(np.float32)


# ---- BASELINE config 4: Sponza instanced, per-mesh compaction -------------------------
def sponza_mesh_table(path: str | None = None):
    """Mesh table from the committed extract of the reference's Sponza.gltf
    (tests/golden/sponza_meshes.json, made by tests/golden/make_sponza_table.py).
    Returns (table, node_scale)."""
    import json
    import os

    if path is None:
        path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "sponza_meshes.json")
    d = json.load(open(path))
    t = np.zeros(len(d["meshes"]), MESH_DTYPE)
    for i, m in enumerate(d["meshes"]):
        t[i]["aabb_min"], t[i]["aabb_max"] = m["aabb_min"], m["aabb_max"]
        t[i]["n_lods"] = len(m["lods"])
        for l, (first, length) in enumerate(m["lods"]):
            t[i]["lod"][l] = (first, length)
        t[i]["vertex_offset"] = m["vertex_offset"]
    return t, float(d["node_scale"])


def sponza_instances(n_entities: int, n_meshes: int, node_scale: float, spacing: float = 40.0, seed: int = SEED):
    """n_entities = copies x n_meshes primitives (the last copy may be cut short): every
    primitive of a copy is its own entity, as scene_loader.rs:588-599 spawns them, sharing the
    copy's transform.  Copies sit on a jittered 3-D grid centred on the origin, random yaw,
    uniform scale = the glTF node scale (scene_loader.rs:760-765 takes scale[0])."""
    copies = -(-n_entities // n_meshes)
    side = int(np.ceil(copies ** (1.0 / 3.0)))
    c = np.arange(copies, dtype=np.uint64)
    grid = np.stack([c % side, (c // side) % side, c // (side * side)], axis=1).astype(np.float32)
    jitter = np.stack([_unit24(_chan(seed, c, a)) for a in range(3)], axis=1) - _F(0.5)
    pos = ((grid - _F((side - 1) / 2.0)) + jitter * _F(0.5)) * _F(spacing)
    yaw = _unit24(_chan(seed, c, 3)).astype(np.float64) * 2.0 * np.pi
    q = np.zeros((copies, 4), np.float64)
    q[:, 1], q[:, 3] = np.sin(yaw / 2), np.cos(yaw / 2)
    q = (q / np.linalg.norm(q, axis=1, keepdims=True)).astype(np.float32)
    nn = np.sqrt((q[:, 0] * q[:, 0] + q[:, 2] * q[:, 2]) + (q[:, 1] * q[:, 1] + q[:, 3] * q[:, 3])).astype(np.float32)
    q = (q / nn[:, None]).astype(np.float32)  # UnitQuaternion::new_normalize (scene_loader.rs:590)
    e = np.arange(n_entities, dtype=np.int64)
    copy_of, mesh = e // n_meshes, (e % n_meshes).astype(np.uint32)
    ps = np.empty((n_entities, 4), np.float32)
    ps[:, :3] = pos[copy_of]
    ps[:, 3] = _F(node_scale)
    return ps, np.ascontiguousarray(q[copy_of]), mesh
