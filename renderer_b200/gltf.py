"""glTF 2.0 -> mesh table (+ consolidated geometry) ingestion for the cull path.

Host-side data preparation only -- nothing here computes a frame.  Mirrors what the reference
does before the per-frame stage can run:

* `load_gltf` (src/renderer/systems/scene_loader.rs:642-789): every primitive of every mesh,
  skipping primitives without a base colour texture (:656-667) and those with fewer than 100
  positions (:677); AABB = the POSITION accessor's min/max (`primitive.bounding_box()`,
  :694-698); LOD 0 = the primitive's index accessor.
* `consolidate_mesh_buffers` (src/renderer/systems/consolidate_mesh_buffers.rs:99-162):
  `first_index` and `vertex_offset` are running sums in load order; positions and indices of
  all meshes end up in one vec3 and one u32 buffer.

The simplified LODs 1..5 come from meshopt::simplify(_sloppy) in the reference
(scene_loader.rs:739-756) -- a third-party mesh decimator that is not restated here: tables
hold LOD 0 only unless `synthetic_lod1=True` asks for a placeholder second LOD (the first half
of the triangle list) so that pick_lod's second branch can be exercised.

Geometry (needed by the per-triangle stage only) is read when the buffers are reachable:
`.bin` files next to the `.gltf`, or `data:` URIs.  A missing `.bin` (the reference tree ships
Sponza without it) leaves `positions`/`indices` as None; the entity-level table needs only the JSON.
"""
from __future__ import annotations

import base64
import json
import os
from dataclasses import dataclass

import numpy as np

from ._lib import MESH_DTYPE

_COMPONENT = {5120: np.int8, 5121: np.uint8, 5122: np.int16, 5123: np.uint16, 5125: np.uint32, 5126: np.float32}
_WIDTH = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4, "MAT4": 16}


@dataclass
class GltfScene:
    meshes: np.ndarray  # MESH_DTYPE [M]
    vertex_counts: np.ndarray  # [M]
    positions: np.ndarray | None  # float32 [sum(vertex_counts)][3] or None (buffer not reachable)
    indices: np.ndarray | None  # uint32 [sum(index_len)] or None
    node_scale: float  # uniform scale of scene node 0 (main.rs applies it to every instance), 1.0 if absent
    skipped: int  # primitives dropped by the reference's filters


def _buffers(g: dict, base_dir: str):
    out = []
    for b in g.get("buffers", []):
        uri = b.get("uri")
        data = None
        if uri is None:
            pass  # GLB-embedded: not used by the reference's assets
        elif uri.startswith("data:"):
            data = base64.b64decode(uri.split(",", 1)[1])
        else:
            path = os.path.join(base_dir, uri)
            if os.path.exists(path):
                data = open(path, "rb").read()
        out.append(data)
    return out


def _read_accessor(g: dict, bufs, index: int) -> np.ndarray | None:
    a = g["accessors"][index]
    if "bufferView" not in a or "sparse" in a:
        return None
    v = g["bufferViews"][a["bufferView"]]
    raw = bufs[v["buffer"]]
    if raw is None:
        return None
    dt = np.dtype(_COMPONENT[a["componentType"]])
    width = _WIDTH[a["type"]]
    start = v.get("byteOffset", 0) + a.get("byteOffset", 0)
    stride = v.get("byteStride") or dt.itemsize * width
    count = a["count"]
    if stride == dt.itemsize * width:
        arr = np.frombuffer(raw, dtype=dt, count=count * width, offset=start).reshape(count, width)
    else:
        arr = np.lib.stride_tricks.as_strided(
            np.frombuffer(raw, dtype=np.uint8, offset=start), shape=(count, width * dt.itemsize), strides=(stride, 1)
        ).copy().view(dt).reshape(count, width)
    return arr


def load_gltf(path: str, *, min_positions: int = 100, require_base_color_texture: bool = True,
              synthetic_lod1: bool = False) -> GltfScene:
    """Mesh table in load order; raises ValueError on a primitive without POSITION min/max or indices."""
    g = json.load(open(path))
    bufs = _buffers(g, os.path.dirname(os.path.abspath(path)))
    acc = g["accessors"]
    rows, vcounts, pos_parts, idx_parts = [], [], [], []
    next_index = next_vertex = skipped = 0
    have_geometry = True
    for mesh in g.get("meshes", []):
        for prim in mesh["primitives"]:
            if require_base_color_texture:  # scene_loader.rs:656-667
                mat = g["materials"][prim["material"]] if "material" in prim else {}
                if "baseColorTexture" not in mat.get("pbrMetallicRoughness", {}):
                    skipped += 1
                    continue
            if "POSITION" not in prim.get("attributes", {}) or "indices" not in prim:
                raise ValueError(f"{path}: primitive without POSITION or indices (reader.read_positions().expect, scene_loader.rs:672)")
            pos = acc[prim["attributes"]["POSITION"]]
            if pos["count"] < min_positions:  # scene_loader.rs:677
                skipped += 1
                continue
            if "min" not in pos or "max" not in pos:
                raise ValueError(f"{path}: POSITION accessor without min/max (primitive.bounding_box())")
            n_idx = acc[prim["indices"]]["count"] // 3 * 3
            lods = [(next_index, n_idx)]
            next_index += n_idx
            if synthetic_lod1:
                lod1 = max(3, (n_idx // 2) // 3 * 3)
                lods.append((next_index, lod1))
                next_index += lod1
            rows.append((pos["min"], pos["max"], lods, next_vertex))
            vcounts.append(pos["count"])
            next_vertex += pos["count"]
            p = _read_accessor(g, bufs, prim["attributes"]["POSITION"])
            i = _read_accessor(g, bufs, prim["indices"])
            if p is None or i is None:
                have_geometry = False
            elif have_geometry:
                pos_parts.append(np.ascontiguousarray(p, np.float32))
                ix = np.ascontiguousarray(i.reshape(-1)[:n_idx], np.uint32)
                idx_parts.append(ix)
                if synthetic_lod1:
                    idx_parts.append(ix[:lods[1][1]])
    table = np.zeros(len(rows), MESH_DTYPE)
    for m, (lo, hi, lods, voff) in enumerate(rows):
        table[m]["aabb_min"], table[m]["aabb_max"] = lo, hi
        table[m]["n_lods"] = len(lods)
        for l, fl in enumerate(lods):
            table[m]["lod"][l] = fl
        table[m]["vertex_offset"] = voff
    scale = 1.0
    nodes = g.get("nodes", [])
    if nodes and "scale" in nodes[0]:
        scale = float(nodes[0]["scale"][0])
    geo = have_geometry and len(rows) > 0
    return GltfScene(
        meshes=table, vertex_counts=np.asarray(vcounts, np.int64),
        positions=np.concatenate(pos_parts) if geo else None,
        indices=np.concatenate(idx_parts) if geo else None,
        node_scale=scale, skipped=skipped,
    )
