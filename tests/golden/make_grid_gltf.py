#!/usr/bin/env python
"""Writes tests/golden/grid.gltf: a self-authored glTF 2.0 file (embedded base64 buffer) for the
ingestion tests -- primitive 0: an 11 x 11 vertex grid (121 positions, 200 triangles, textured
material) that the reference's loader would keep; primitive 1: a quad of 4 positions (dropped:
fewer than 100 positions, scene_loader.rs:677); primitive 2: a second grid without a base colour
texture (dropped, scene_loader.rs:656-667); mesh 1 / primitive 0: a 12 x 10 grid with u16 indices
and an interleaved (byteStride 24) position view.  Node 0 carries scale 0.5."""
import base64, json, os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def grid(nx, ny, z, dtype):
    xs, ys = np.meshgrid(np.linspace(-1, 1, nx, dtype=np.float32), np.linspace(0, 3, ny, dtype=np.float32))
    pos = np.stack([xs.ravel(), ys.ravel(), np.full(nx * ny, z, np.float32) + 0.1 * xs.ravel() ** 2], axis=1).astype(np.float32)
    idx = []
    for j in range(ny - 1):
        for i in range(nx - 1):
            a = j * nx + i
            idx += [a, a + 1, a + nx, a + 1, a + nx + 1, a + nx]
    return pos, np.asarray(idx, dtype)


blob = bytearray()
views, accessors = [], []


def add_view(data: bytes, stride=None):
    while len(blob) % 4:
        blob.append(0)
    v = {"buffer": 0, "byteOffset": len(blob), "byteLength": len(data)}
    if stride:
        v["byteStride"] = stride
    blob.extend(data)
    views.append(v)
    return len(views) - 1


def add_positions(pos, interleave=False):
    if interleave:  # position + 12 bytes of padding per vertex, byteStride 24
        data = np.zeros((len(pos), 6), np.float32)
        data[:, :3] = pos
        view = add_view(data.tobytes(), stride=24)
    else:
        view = add_view(pos.tobytes())
    accessors.append({"bufferView": view, "componentType": 5126, "count": len(pos), "type": "VEC3",
                      "min": pos.min(0).tolist(), "max": pos.max(0).tolist()})
    return len(accessors) - 1


def add_indices(idx):
    view = add_view(idx.tobytes())
    accessors.append({"bufferView": view, "componentType": 5125 if idx.dtype == np.uint32 else 5123, "count": len(idx), "type": "SCALAR"})
    return len(accessors) - 1


p0, i0 = grid(11, 11, 0.0, np.uint32)
quad = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 1, 0]], np.float32)
p2, i2 = grid(10, 10, 2.0, np.uint32)
p3, i3 = grid(12, 10, -1.0, np.uint16)
prims0 = [
    {"attributes": {"POSITION": add_positions(p0)}, "indices": add_indices(i0), "material": 0},
    {"attributes": {"POSITION": add_positions(quad)}, "indices": add_indices(np.array([0, 1, 2, 1, 3, 2], np.uint32)), "material": 0},
    {"attributes": {"POSITION": add_positions(p2)}, "indices": add_indices(i2), "material": 1},
]
prims1 = [{"attributes": {"POSITION": add_positions(p3, interleave=True)}, "indices": add_indices(i3), "material": 0}]
doc = {
    "asset": {"version": "2.0", "generator": "renderer_b200 tests/golden/make_grid_gltf.py"},
    "scene": 0, "scenes": [{"nodes": [0]}],
    "nodes": [{"mesh": 0, "scale": [0.5, 0.5, 0.5], "children": [1]}, {"mesh": 1}],
    "meshes": [{"primitives": prims0}, {"primitives": prims1}],
    "materials": [{"pbrMetallicRoughness": {"baseColorTexture": {"index": 0}}}, {"pbrMetallicRoughness": {"baseColorFactor": [1, 0, 0, 1]}}],
    "textures": [{"source": 0}], "images": [{"uri": "none.png"}],
    "accessors": accessors, "bufferViews": views,
    "buffers": [{"byteLength": len(blob), "uri": "data:application/octet-stream;base64," + base64.b64encode(bytes(blob)).decode()}],
}
path = os.path.join(HERE, "grid.gltf")
json.dump(doc, open(path, "w"))
print("wrote", path, os.path.getsize(path), "bytes")
