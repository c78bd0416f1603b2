#!/usr/bin/env python
"""Generates tests/golden/config1_10k.npz -- BASELINE configs[0]: 10 k random-TRS
entities, default camera + the 4 shadow views -- from the CPU oracle.

The reference is Rust (bevy_ecs 0.6 + nalgebra 0.29) and holds no golden vector
for this path; it cannot be built or imported here, so these vectors come from
oracle/ (PARITY UNPINNED).  They pin the oracle against silent
drift and give the GPU tests a committed target that does not depend on
recompiling the oracle on the GPU box.

Run from the repo root:  python tests/golden/make_golden.py
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

N, N_VIEWS, HALF_BOX = 10_000, 5, 60.0


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def canon(a: np.ndarray) -> np.ndarray:
    """-0.0 -> +0.0 so the digest is insensitive to the sign of zero."""
    return a + np.float32(0.0)


def build():
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, N, HALF_BOX)
    meshes = scene.synth_mesh_table()
    views = scene.scene_views(N_VIEWS)
    pos, quat, scale = scene.to_components(ps, rot)
    ref = O.frame(pos, quat, scale, mesh, meshes, views, scene.CAMERA_POS)
    out = {
        "n": np.int64(N), "half_box": np.float32(HALF_BOX), "seed": np.uint64(scene.SEED),
        "views": views, "meshes": meshes, "cam_pos": scene.CAMERA_POS,
        "inputs_sha256": np.array([sha(ps), sha(rot), sha(mesh)]),
        "counts": ref["counts"], "totals": ref["totals"],
        "flags_packed": np.packbits(ref["flags"], axis=1),
        "model_head": ref["model"][:64].copy(), "aabb_head": ref["aabb"][:64].copy(),
        "model_sha256": np.array(sha(canon(ref["model"]))), "aabb_sha256": np.array(sha(canon(ref["aabb"]))),
    }
    for k in range(N_VIEWS):
        out[f"visible_{k}"] = ref["visible"][k]
        out[f"cmds_{k}"] = ref["cmds"][k]
    return out


if __name__ == "__main__":
    path = os.path.join(HERE, "config1_10k.npz")
    np.savez_compressed(path, **build())
    print("wrote", path, os.path.getsize(path), "bytes")
