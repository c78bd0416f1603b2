"""GPU parity of stage 2 (SURVEY 8a rows A8 + A7): per-triangle cull of the camera view's
survivors and compaction of the commands that kept a triangle, against the CPU oracle's
restatement of generate_work.comp / compact_draw_stream.comp.  Commands, counts and the
culled index buffer must be bit-exact (ascending-triangle order is the canonical order of
both; the reference's own order inside a mesh is atomic-arrival order)."""
import numpy as np
import pytest

import oracle_lib as O
from renderer_b200 import CullContext, RcullError, scene

pytestmark = pytest.mark.gpu

MESHES = scene.synth_mesh_table()
POSITIONS, INDICES = scene.synth_geometry(MESHES)


def run_tri(n, half_box, *, base=0, cam_pos=scene.CAMERA_POS, capacity=None, threads=8, seed=scene.SEED, geometry=None):
    MESHES, POSITIONS, INDICES = geometry if geometry is not None else (globals()['MESHES'], globals()['POSITIONS'], globals()['INDICES'])
    ps, rot, mesh = scene.synth_entities(seed, 0, n, half_box, len(MESHES))
    views = scene.camera_view(pos=cam_pos)[None]
    pv = scene.camera_pv(pos=cam_pos)
    pos, quat, scale = scene.to_components(ps, rot)
    ref = O.frame(pos, quat, scale, mesh, MESHES, views, cam_pos, draw_index_base=base, n_threads=threads)
    total = int(ref["totals"][0])
    cap = capacity if capacity is not None else max(total, 3)
    with CullContext(max(n, 1), max_views=1, max_meshes=len(MESHES), capacity_tri_indices=cap) as ctx:
        ctx.set_mesh_table(MESHES)
        ctx.set_geometry(POSITIONS, INDICES)
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(views, cam_pos, n=n, draw_index_base=base)
        ctx.tri_cull(pv, draw_index_base=base)
        count, kept = ctx.tri_counts()
        src = O.src_first_index(pos, mesh, ref["cmds"][0], MESHES, cam_pos, draw_index_base=base)
        want = O.tri_cull(ref["cmds"][0], src, ref["model"], pv, POSITIONS, INDICES, cap, draw_index_base=base, n_threads=threads)
        assert count == want["count"]
        assert kept == want["kept_indices"]
        got_cmds = ctx.tri_cmds(count)
        assert np.array_equal(got_cmds, want["cmds"])
        got_idx = ctx.tri_indices(0, cap)
        for c in want["cmds"]:  # only the ranges the commands own are defined
            a, b = int(c["firstIndex"]), int(c["firstIndex"]) + int(c["indexCount"])
            assert np.array_equal(got_idx[a:b], want["indices"][a:b])
        assert ctx.launch_count in (3, 4)  # frame + tri_cull (one kernel per kind of LOD present) + compact_draw_stream
        return count, kept, total, int(ref["counts"][0])


def test_tri_cull_small_scene():
    count, kept, total, visible = run_tri(20_000, 40.0)
    assert 0 < count <= visible and 0 < kept < total


@pytest.mark.parametrize("n", [1, 33, 1500])
def test_tri_cull_ragged(n):
    run_tri(n, 15.0)


def test_tri_cull_draw_index_base():
    run_tri(5_000, 30.0, base=1_000_000)


def test_tri_cull_nothing_visible():
    count, kept, total, visible = run_tri(2_000, 40.0, cam_pos=np.array([5000.0, 0.0, 0.0], np.float32))
    assert (count, kept, visible) == (0, 0, 0)


def test_tri_cull_config2_1m_camera():
    count, kept, total, visible = run_tri(1_000_000, 250.0, threads=16)
    assert visible > 10_000 and count > 0 and kept % 3 == 0


def test_tri_cull_errors_are_loud():
    with CullContext(100, max_views=1, max_meshes=len(MESHES)) as ctx:
        with pytest.raises(RcullError):  # context created without the stage
            ctx.set_geometry(POSITIONS, INDICES)
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, 2000, 15.0, len(MESHES))
    views = scene.camera_view()[None]
    with CullContext(2000, max_views=1, max_meshes=len(MESHES), capacity_tri_indices=300) as ctx:
        ctx.set_mesh_table(MESHES)
        with pytest.raises(RcullError):  # no geometry yet
            ctx.tri_cull(scene.camera_pv())
        ctx.set_geometry(POSITIONS, INDICES)
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(views, scene.CAMERA_POS)
        ctx.tri_cull(scene.camera_pv())
        with pytest.raises(RcullError):  # culled index buffer too small
            ctx.tri_counts()
    with CullContext(2000, max_views=1, max_meshes=len(MESHES), capacity_tri_indices=1 << 22) as ctx:
        ctx.set_mesh_table(MESHES)
        ctx.set_geometry(POSITIONS[:100], INDICES)  # vertices missing
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(views, scene.CAMERA_POS)
        ctx.tri_cull(scene.camera_pv())
        with pytest.raises(RcullError):
            ctx.tri_counts()


def test_tri_cull_on_gltf_geometry():
    """Meshes and consolidated buffers from the glTF loader (tests/golden/grid.gltf: real connected
    triangle grids, two LODs, u16 indices widened) through frame -> tri_cull -> compact."""
    import os
    from renderer_b200 import gltf
    g = gltf.load_gltf(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "grid.gltf"), synthetic_lod1=True)
    count, kept, total, visible = run_tri(3_000, 14.0, geometry=(g.meshes, g.positions, g.indices))
    assert visible > 100 and 0 < kept < total


def test_tri_cull_coherent_grid_meshes():
    """The bench's mesh flavour: grid sheets with shared vertices and both windings."""
    pos, idx = scene.synth_geometry(MESHES, coherent=True)
    count, kept, total, visible = run_tri(20_000, 40.0, geometry=(MESHES, pos, idx))
    assert 0 < kept < total


def test_tri_cull_ndc_boundaries_and_nonpositive_w():
    """Clip coordinates placed exactly on and one ulp around the NDC limits, with w > 0, tiny, zero and
    negative: the GPU skips the six divisions when a vertex is provably inside (w > 0, |x| <= w, |y| <= w)
    and must still agree with the shader arithmetic everywhere.  Identity model, pv = (x, y, z, w = z)."""
    rng = np.random.default_rng(77)
    f = np.float32
    xy = np.array([-2, np.nextafter(f(-1), f(-2)), -1, np.nextafter(f(-1), f(0)), -0.5, 0, 0.5,
                   np.nextafter(f(1), f(0)), 1, np.nextafter(f(1), f(2)), 2, 1e-30, -1e-30], f)
    zs = np.array([-1, -1e-30, 0, 1e-30, 0.5, 1, 1, 1, 2], f)
    n_tri = 6000
    positions = np.stack([rng.choice(xy, 3 * n_tri), rng.choice(xy, 3 * n_tri), rng.choice(zs, 3 * n_tri)], axis=1).astype(f)
    indices = np.arange(3 * n_tri, dtype=np.uint32)
    meshes = np.zeros(1, scene.MESH_DTYPE)
    meshes[0]["aabb_min"], meshes[0]["aabb_max"], meshes[0]["n_lods"] = (-2, -2, -1), (2, 2, 2), 1
    meshes[0]["lod"][0] = (0, 3 * n_tri)
    ps = np.array([[0, 0, 0, 1]], f)
    rot = np.array([[0, 0, 0, 1]], f)
    mesh = np.zeros(1, np.uint32)
    views = np.zeros((1, 6, 4), f)
    views[..., 3] = -1.0  # s = -1, e = 0: inside every plane
    pv = np.zeros(16, f)  # column-major: x' = x, y' = y, z' = z, w' = z
    pv[0] = pv[5] = pv[10] = pv[11] = 1.0
    cam = scene.CAMERA_POS
    pos, quat, scale = scene.to_components(ps, rot)
    ref = O.frame(pos, quat, scale, mesh, meshes, views, cam)
    assert ref["counts"][0] == 1 and np.array_equal(ref["model"].reshape(4, 4), np.eye(4, dtype=f))
    src = O.src_first_index(pos, mesh, ref["cmds"][0], meshes, cam)
    want = O.tri_cull(ref["cmds"][0], src, ref["model"], pv, positions, indices, 3 * n_tri)
    with CullContext(1, max_views=1, max_meshes=1, capacity_tri_indices=3 * n_tri) as ctx:
        ctx.set_mesh_table(meshes)
        ctx.set_geometry(positions, indices)
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(views, cam)
        ctx.tri_cull(pv)
        count, kept = ctx.tri_counts()
        assert (count, kept) == (want["count"], want["kept_indices"])
        assert 0 < kept < 3 * n_tri
        assert np.array_equal(ctx.tri_cmds(count), want["cmds"])
        assert np.array_equal(ctx.tri_indices(0, kept), want["indices"][:kept])
