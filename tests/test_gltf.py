"""glTF ingestion: mesh table and consolidated geometry in the reference's load
order (scene_loader.rs:642-789, consolidate_mesh_buffers.rs:99-162).  CPU only."""
import json
import os

import numpy as np
import pytest

from renderer_b200 import gltf, scene

HERE = os.path.dirname(os.path.abspath(__file__))
GRID = os.path.join(HERE, "golden", "grid.gltf")
REF_SPONZA = "/root/reference/vendor/glTF-Sample-Models/2.0/Sponza/glTF/Sponza.gltf"


def test_grid_fixture_table_and_geometry():
    s = gltf.load_gltf(GRID)
    assert len(s.meshes) == 2 and s.skipped == 2  # the quad (< 100 positions) and the untextured grid are dropped
    assert s.node_scale == 0.5
    assert s.vertex_counts.tolist() == [121, 120]
    m0, m1 = s.meshes
    assert m0["n_lods"] == 1 and tuple(m0["lod"][0]) == (0, 600) and m0["vertex_offset"] == 0
    assert tuple(m1["lod"][0]) == (600, 11 * 9 * 6) and m1["vertex_offset"] == 121  # running sums
    assert s.positions.shape == (241, 3) and s.indices.shape == (600 + 594,) and s.indices.dtype == np.uint32
    # AABB = accessor min/max = the actual extent of the positions
    for m, lo, n in ((m0, 0, 121), (m1, 121, 120)):
        assert np.allclose(m["aabb_min"], s.positions[lo : lo + n].min(0))
        assert np.allclose(m["aabb_max"], s.positions[lo : lo + n].max(0))
    # indices are relative to the mesh's vertex_offset and in range (u16 source widened to u32)
    assert s.indices[:600].max() == 120 and s.indices[600:].max() == 119
    # the interleaved (byteStride 24) view was de-interleaved: x spans [-1, 1] over 12 columns
    assert np.allclose(np.unique(s.positions[121:, 0]), np.linspace(-1, 1, 12, dtype=np.float32))


def test_grid_fixture_synthetic_lod1_layout():
    s = gltf.load_gltf(GRID, synthetic_lod1=True)
    m0, m1 = s.meshes
    assert m0["n_lods"] == 2 and tuple(m0["lod"][0]) == (0, 600) and tuple(m0["lod"][1]) == (600, 300)
    assert tuple(m1["lod"][0]) == (900, 594) and tuple(m1["lod"][1]) == (1494, 297)
    assert len(s.indices) == 1494 + 297
    assert np.array_equal(s.indices[600:900], s.indices[:300])  # placeholder LOD 1 = first half of LOD 0


def test_filters_can_be_disabled():
    s = gltf.load_gltf(GRID, min_positions=0, require_base_color_texture=False)
    assert len(s.meshes) == 4 and s.skipped == 0
    assert s.vertex_counts.tolist() == [121, 4, 100, 120]


def test_matches_the_committed_sponza_extract():
    """The committed mesh table of BASELINE config 4 (tests/golden/sponza_meshes.json) is what this
    loader produces from the reference's Sponza.gltf (JSON only: the .bin is not in the tree)."""
    if not os.path.exists(REF_SPONZA):
        pytest.skip("reference tree not present on this box")
    s = gltf.load_gltf(REF_SPONZA, synthetic_lod1=True, require_base_color_texture=False)
    want, scale = scene.sponza_mesh_table()
    assert s.positions is None and s.indices is None  # Sponza.bin is missing (.MISSING_LARGE_BLOBS)
    assert abs(s.node_scale - scale) < 1e-9
    assert len(s.meshes) == len(want)
    assert s.meshes.tobytes() == want.tobytes()
