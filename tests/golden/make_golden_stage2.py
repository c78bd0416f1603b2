#!/usr/bin/env python
"""Generates tests/golden/stage2_per_mesh.npz from the CPU oracle: committed targets for the two
output paths config1_10k.npz does not cover --

  * stage 2 (SURVEY 8a rows A8 + A7): 1500 entities in a +-15 box, camera view; the compacted triangle-stage
    commands, their count, the kept-index total and the culled index ranges the commands own;
  * per-mesh instanced mode (BASELINE configs[3] shape): 900 entities in a +-30 box, camera + 2 shadow views,
    draw_index_base 17; one command per non-empty (mesh, lod) slot and the bucketed visible list per view.

Like config1_10k.npz these come from oracle/ (PARITY UNPINNED): they pin the oracle against
silent drift and give the GPU tests a target that does not depend on rebuilding the oracle.

Run from the repo root:  python tests/golden/make_golden_stage2.py
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle_lib as O  # noqa: E402
from renderer_b200 import scene  # noqa: E402

TRI_N, TRI_HALF_BOX = 1500, 15.0
HEAD = 4096  # culled indices kept verbatim; the rest is pinned by a digest


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def owned_indices(cmds, index_buffer) -> np.ndarray:
    """The culled-index ranges the commands own, concatenated in command order (the rest of the buffer is undefined)."""
    owned = [index_buffer[int(c["firstIndex"]): int(c["firstIndex"]) + int(c["indexCount"])] for c in cmds]
    return np.concatenate(owned) if owned else np.zeros(0, np.uint32)
PM_N, PM_HALF_BOX, PM_VIEWS, PM_BASE = 900, 30.0, 3, 17


def build():
    meshes = scene.synth_mesh_table()
    positions, indices = scene.synth_geometry(meshes)
    out = {"meshes": meshes, "cam_pos": scene.CAMERA_POS, "seed": np.uint64(scene.SEED)}

    # ---- stage 2 ----
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, TRI_N, TRI_HALF_BOX, len(meshes))
    pos, quat, scale = scene.to_components(ps, rot)
    views, pv = scene.camera_view()[None], scene.camera_pv()
    ref = O.frame(pos, quat, scale, mesh, meshes, views, scene.CAMERA_POS)
    total = int(ref["totals"][0])
    src = O.src_first_index(pos, mesh, ref["cmds"][0], meshes, scene.CAMERA_POS)
    tri = O.tri_cull(ref["cmds"][0], src, ref["model"], pv, positions, indices, max(total, 3))
    owned = owned_indices(tri["cmds"], tri["indices"])
    out.update({
        "tri_n": np.int64(TRI_N), "tri_half_box": np.float32(TRI_HALF_BOX), "tri_view": views, "tri_pv": pv,
        "tri_entity_count": ref["counts"][:1], "tri_entity_total": ref["totals"][:1],
        "tri_count": np.uint32(tri["count"]), "tri_kept": np.uint64(tri["kept_indices"]), "tri_cmds": tri["cmds"],
        "tri_owned_head": owned[:HEAD].copy(), "tri_owned_len": np.uint64(len(owned)), "tri_owned_sha256": np.array(sha(owned)),
    })

    # ---- per-mesh instanced mode ----
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, PM_N, PM_HALF_BOX, len(meshes))
    pos, quat, scale = scene.to_components(ps, rot)
    views = scene.scene_views(PM_VIEWS)
    ref = O.frame(pos, quat, scale, mesh, meshes, views, scene.CAMERA_POS, want=("flags",))
    out.update({"pm_n": np.int64(PM_N), "pm_half_box": np.float32(PM_HALF_BOX), "pm_views": views, "pm_base": np.uint32(PM_BASE)})
    for k in range(PM_VIEWS):
        cmds, vis = O.emit_per_mesh(pos, mesh, ref["flags"][k], meshes, scene.CAMERA_POS, 10.0, PM_BASE)
        out[f"pm_cmds_{k}"], out[f"pm_visible_{k}"] = cmds, vis
    return out


if __name__ == "__main__":
    path = os.path.join(HERE, "stage2_per_mesh.npz")
    np.savez_compressed(path, **build())
    print("wrote", path, os.path.getsize(path), "bytes")
