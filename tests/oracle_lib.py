"""ctypes wrapper of oracle/librcull_oracle.so -- TEST INFRASTRUCTURE.
Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATH = os.path.join(ROOT, "oracle", "librcull_oracle.so")

MESH_DTYPE = np.dtype(
    [("aabb_min", "<f4", (3,)), ("aabb_max", "<f4", (3,)), ("n_lods", "<u4"), ("lod", "<u4", (6, 2)), ("vertex_offset", "<i4")]
)
CMD_DTYPE = np.dtype(
    [("indexCount", "<u4"), ("instanceCount", "<u4"), ("firstIndex", "<u4"), ("vertexOffset", "<i4"), ("firstInstance", "<u4")]
)
_P, _u32, _u64, _f = C.c_void_p, C.c_uint32, C.c_uint64, C.c_float
_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(PATH)
        L.oracle_model_matrix.argtypes = [_P, _P, _f, _P]
        L.oracle_aabb.argtypes = [_P] * 5
        L.oracle_coarse_culled.argtypes = [_P, _P, _P]
        L.oracle_coarse_culled.restype = C.c_int
        L.oracle_pick_lod.argtypes = [_P, _P, _u32, _f]
        L.oracle_pick_lod.restype = _u32
        L.oracle_frustum_planes.argtypes = [_P, _P]
        L.oracle_project_camera.argtypes = [_f, _P, _P, _P, _P, _P, _P]
        L.oracle_light_view.argtypes = [_P] * 6
        L.oracle_quat_look_at_lh.argtypes = [_P, _P, _P]
        L.oracle_mat4_mul.argtypes = [_P, _P, _P]
        L.oracle_frame.argtypes = [_P, _P, _P, _P, _u64, _P, _u32, _P, _u32, _P, _f, _u32, _P, _P, _P, _P, _P, _P, _P, C.c_int]
        L.oracle_frame.restype = C.c_int
        L.oracle_frame_fused.argtypes = L.oracle_frame.argtypes
        L.oracle_frame_fused.restype = C.c_int
        L.oracle_emit_per_mesh.argtypes = [_P, _P, _P, _u64, _P, _u32, _P, _f, _u32, _P, _P, _P, _P]
        L.oracle_emit_per_mesh.restype = C.c_int
        L.oracle_tri_cull.argtypes = [_P, _P, _u32, _P, _u32, _P, _P, _u64, _P, _u64, _P, _u64, _P, _P, _P, C.c_int]
        L.oracle_tri_cull.restype = C.c_int
        L.oracle_as_instances.argtypes = [_P, _P, _P, _u64, _u32, _P]
        L.oracle_synth_entities.argtypes = [_u64, _u64, _u64, _f, _u32, _P, _P, _P, _P]
        _lib = L
    return _lib


def p(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


# NOTE: p() takes a raw address; every temporary is bound to a name first so that it
# stays alive across the call.
def model_matrix(pos, quat, scale):
    m = np.empty(16, np.float32)
    a, b = f32(pos), f32(quat)
    lib().oracle_model_matrix(p(a), p(b), float(scale), p(m))
    return m


def aabb(m, mesh_min, mesh_max):
    lo, hi = np.empty(3, np.float32), np.empty(3, np.float32)
    a, b, c = f32(m), f32(mesh_min), f32(mesh_max)
    lib().oracle_aabb(p(a), p(b), p(c), p(lo), p(hi))
    return lo, hi


def coarse_culled(mins, maxs, planes):
    a, b, c = f32(mins), f32(maxs), f32(planes)
    return bool(lib().oracle_coarse_culled(p(a), p(b), p(c)))


def pick_lod(cam, pos, n_lods, lod_distance=10.0):
    a, b = f32(cam), f32(pos)
    return int(lib().oracle_pick_lod(p(a), p(b), n_lods, lod_distance))


def frustum_planes(pv):
    out = np.empty((6, 4), np.float32)
    a = f32(pv)
    lib().oracle_frustum_planes(p(a), p(out))
    return out


def project_camera(aspect, pos, rot):
    proj, view, pv = (np.empty(16, np.float32) for _ in range(3))
    planes = np.empty((6, 4), np.float32)
    a, b = f32(pos), f32(rot)
    lib().oracle_project_camera(aspect, p(a), p(b), p(proj), p(view), p(pv), p(planes))
    return proj, view, pv, planes


def light_view(pos, rot):
    proj, view, pv = (np.empty(16, np.float32) for _ in range(3))
    planes = np.empty((6, 4), np.float32)
    a, b = f32(pos), f32(rot)
    lib().oracle_light_view(p(a), p(b), p(proj), p(view), p(pv), p(planes))
    return proj, view, pv, planes


def quat_look_at_lh(direction, up=(0, 1, 0)):
    q = np.empty(4, np.float32)
    a, b = f32(direction), f32(up)
    lib().oracle_quat_look_at_lh(p(a), p(b), p(q))
    return q


def synth_entities(seed, first, n, half_box, n_meshes):
    pos, quat = np.empty((n, 3), np.float32), np.empty((n, 4), np.float32)
    scale, mesh = np.empty(n, np.float32), np.empty(n, np.uint32)
    lib().oracle_synth_entities(seed, first, n, half_box, n_meshes, p(pos), p(quat), p(scale), p(mesh))
    return pos, quat, scale, mesh


def frame(pos, quat, scale, mesh_id, meshes, planes, cam_pos, lod_distance=10.0, draw_index_base=0,
          n_threads=1, want=("model", "aabb", "flags", "cmds", "visible"), fused=False):
    """Runs oracle_frame (or oracle_frame_fused: same results, one pass per entity + parallel emit);
    returns a dict of the requested outputs (lists trimmed to counts)."""
    pos, quat, scale = f32(pos).reshape(-1, 3), f32(quat).reshape(-1, 4), f32(scale).reshape(-1)
    mesh_id = np.ascontiguousarray(mesh_id, np.uint32)
    meshes = np.ascontiguousarray(meshes, MESH_DTYPE)
    planes = f32(planes).reshape(-1, 6, 4)
    n, v = len(scale), len(planes)
    out = {
        "model": np.empty((n, 16), np.float32) if "model" in want else None,
        "aabb": np.empty((n, 6), np.float32) if "aabb" in want else None,
        "flags": np.empty((v, n), np.uint8) if "flags" in want else None,
        "cmds": np.empty((v, max(n, 1)), CMD_DTYPE) if "cmds" in want else None,
        "visible": np.empty((v, max(n, 1)), np.uint32) if "visible" in want else None,
    }
    counts, totals = np.zeros(v, np.uint32), np.zeros(v, np.uint64)
    cam = f32(cam_pos)
    rc = (lib().oracle_frame_fused if fused else lib().oracle_frame)(p(pos), p(quat), p(scale), p(mesh_id), n, p(meshes), len(meshes), p(planes), v,
                            p(cam), lod_distance, draw_index_base, p(out["model"]), p(out["aabb"]),
                            p(out["flags"]), p(out["cmds"]), p(out["visible"]), p(counts), p(totals), n_threads)
    if rc != 0:
        raise ValueError("oracle_frame: mesh id out of range")
    out["counts"], out["totals"] = counts, totals
    if out["cmds"] is not None:
        out["cmds"] = [out["cmds"][k, : counts[k]] for k in range(v)]
    if out["visible"] is not None:
        out["visible"] = [out["visible"][k, : counts[k]] for k in range(v)]
    return out


def emit_per_mesh(pos, mesh_id, flags, meshes, cam_pos, lod_distance=10.0, draw_index_base=0):
    pos = f32(pos).reshape(-1, 3)
    n = len(pos)
    meshes = np.ascontiguousarray(meshes, MESH_DTYPE)
    cmds = np.empty(len(meshes) * 6, CMD_DTYPE)
    visible = np.empty(max(n, 1), np.uint32)
    count, inst = _u32(0), _u64(0)
    mesh_id, flags, cam = np.ascontiguousarray(mesh_id, np.uint32), np.ascontiguousarray(flags, np.uint8), f32(cam_pos)
    rc = lib().oracle_emit_per_mesh(p(pos), p(mesh_id), p(flags),
                                    n, p(meshes), len(meshes), p(cam), lod_distance, draw_index_base,
                                    p(cmds), p(visible), C.byref(count), C.byref(inst))
    if rc != 0:
        raise ValueError("oracle_emit_per_mesh failed")
    return cmds[: count.value], visible[: inst.value]


def src_first_index(pos, mesh_id, cmds, meshes, cam_pos, lod_distance=10.0, draw_index_base=0):
    """indexOffset push constant of every command (cull_pipeline.rs:544-548): first_index of the
    LOD pick_lod chooses for the command's entity."""
    pos = f32(pos)
    cam = f32(cam_pos)
    out = np.empty(len(cmds), np.uint32)
    for c, cmd in enumerate(cmds):
        e = int(cmd["firstInstance"]) - draw_index_base
        m = meshes[int(mesh_id[e])]
        lod = pick_lod(cam, pos[e], int(m["n_lods"]), lod_distance)
        out[c] = m["lod"][lod][0]
    return out


def tri_cull(cmds, src_first, model, pv, positions, indices, out_capacity, draw_index_base=0, n_threads=1):
    """A8 + A7 for one view's command list.  Returns dict(indices, cmds, count, kept_indices)."""
    cmds = np.ascontiguousarray(cmds, dtype=CMD_DTYPE)
    src_first = np.ascontiguousarray(src_first, dtype=np.uint32)
    model, pvv = f32(model), f32(pv).reshape(16)
    positions = f32(positions)
    indices = np.ascontiguousarray(indices, dtype=np.uint32)
    out_idx = np.zeros(int(out_capacity), np.uint32)
    out_cmds = np.zeros(len(cmds), CMD_DTYPE)
    count = np.zeros(1, np.uint32)
    kept = np.zeros(1, np.uint64)
    rc = lib().oracle_tri_cull(p(cmds), p(src_first), len(cmds), p(model), draw_index_base, p(pvv), p(positions),
                               len(positions) // 3 if positions.ndim == 1 else len(positions), p(indices), len(indices),
                               p(out_idx), int(out_capacity), p(out_cmds), p(count), p(kept), int(n_threads))
    if rc != 0:
        raise ValueError("oracle_tri_cull: index or vertex out of range")
    return dict(indices=out_idx, cmds=out_cmds[: int(count[0])], count=int(count[0]), kept_indices=int(kept[0]))


AS_INSTANCE_DTYPE = np.dtype(
    [("transform", "<f4", (3, 4)), ("custom_index_and_mask", "<u4"), ("sbt_offset_and_flags", "<u4"), ("blas", "<u8")]
)


def as_instances(model, mesh_id, blas, draw_index_base=0):
    model = f32(model)
    mesh_id = np.ascontiguousarray(mesh_id, dtype=np.uint32)
    blas = np.ascontiguousarray(blas, dtype=np.uint64)
    out = np.zeros(len(mesh_id), AS_INSTANCE_DTYPE)
    lib().oracle_as_instances(p(model), p(mesh_id), p(blas), len(mesh_id), draw_index_base, p(out))
    return out
