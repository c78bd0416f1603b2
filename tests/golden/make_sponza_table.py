#!/usr/bin/env python
"""Extracts the mesh table of BASELINE config 4 from the reference's Sponza.gltf JSON
(vendor/glTF-Sample-Models/2.0/Sponza/glTF/Sponza.gltf; the .bin is missing from the
tree and not needed at entity level) into tests/golden/sponza_meshes.json.

Mirrors what scene_loader.rs does per primitive: skip primitives with < 100 positions
(scene_loader.rs:677), AABB = POSITION accessor min/max (scene_loader.rs:694-698), LOD 0
index count = indices accessor count; first_index / vertex_offset are running sums in
primitive order, as consolidate_mesh_buffers.rs:100-159 assigns them; node scale
0.008 (glTF node 0) is recorded for the instance generator.  LODs 1..5 come from
meshopt::simplify_sloppy at run time in the reference (scene_loader.rs:739-756); their
sizes cannot be reproduced here, so one synthetic LOD 1 of ~half the indices (rounded to
whole triangles) is recorded to exercise pick_lod.

Run HERE (needs /root/reference):  python tests/golden/make_sponza_table.py
"""
import json
import os

SRC = "/root/reference/vendor/glTF-Sample-Models/2.0/Sponza/glTF/Sponza.gltf"
HERE = os.path.dirname(os.path.abspath(__file__))

g = json.load(open(SRC))
acc = g["accessors"]
meshes, next_index, next_vertex = [], 0, 0
for prim in g["meshes"][0]["primitives"]:
    pos = acc[prim["attributes"]["POSITION"]]
    if pos["count"] < 100:
        continue
    n_idx = acc[prim["indices"]]["count"]
    lod1 = max(3, (n_idx // 2) // 3 * 3)
    meshes.append({
        "aabb_min": pos["min"], "aabb_max": pos["max"], "vertex_count": pos["count"],
        "lods": [[next_index, n_idx], [next_index + n_idx, lod1]], "vertex_offset": next_vertex,
    })
    next_index += n_idx + lod1
    next_vertex += pos["count"]
out = {"source": "glTF-Sample-Models/2.0/Sponza/glTF/Sponza.gltf", "node_scale": g["nodes"][0]["scale"][0],
       "primitives_total": len(g["meshes"][0]["primitives"]), "meshes": meshes}
path = os.path.join(HERE, "sponza_meshes.json")
json.dump(out, open(path, "w"))
print("wrote", path, len(meshes), "meshes of", out["primitives_total"], "primitives,", os.path.getsize(path), "bytes")
