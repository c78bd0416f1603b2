"""CPU tests of the oracle (oracle/rcull_oracle.c): hand-checkable known answers,
an independent numpy restatement of the same operation order, and the committed
golden vectors.  The reference holds no test or fixture for this path,
so these are what pins the oracle; see DESIGN.md "parity unpinned"."""
import os

import numpy as np
import pytest

import oracle_lib as O
from renderer_b200 import scene

F = np.float32
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "config1_10k.npz")


# ---- A1: model matrix (src/ecs.rs:61-62) -------------------------------------------
def test_model_matrix_identity_rotation_is_scale_and_translation():
    m = O.model_matrix([1.5, -2.0, 3.25], [0, 0, 0, 1], 2.0).reshape(4, 4).T  # row, col
    want = np.array([[2, 0, 0, 1.5], [0, 2, 0, -2.0], [0, 0, 2, 3.25], [0, 0, 0, 1]], F)
    assert np.array_equal(m, want)


def test_model_matrix_quarter_turn_about_y():
    h = F(np.sqrt(0.5))
    m = O.model_matrix([0, 0, 0], [0, h, 0, h], 1.0).reshape(4, 4).T
    # +90 degrees about Y maps x -> -z and z -> x
    assert np.allclose(m @ np.array([1, 0, 0, 1], F), [0, 0, -1, 1], atol=1e-6)
    assert np.allclose(m @ np.array([0, 0, 1, 1], F), [1, 0, 0, 1], atol=1e-6)
    assert np.array_equal(m[3], [0, 0, 0, 1])


def _np_model(ps, q):
    """Independent numpy float32 restatement of row A1 (vectorised over entities)."""
    i, j, k, w, s = q[:, 0], q[:, 1], q[:, 2], q[:, 3], ps[:, 3]
    ww, ii, jj, kk = w * w, i * i, j * j, k * k
    ij, wk, wj = i * j * F(2), w * k * F(2), w * j * F(2)
    ik, jk, wi = i * k * F(2), j * k * F(2), w * i * F(2)
    m = np.zeros((len(s), 16), F)
    m[:, 0] = (ww + ii - jj - kk) * s; m[:, 4] = (ij - wk) * s; m[:, 8] = (wj + ik) * s
    m[:, 1] = (wk + ij) * s; m[:, 5] = (ww - ii + jj - kk) * s; m[:, 9] = (jk - wi) * s
    m[:, 2] = (ik - wj) * s; m[:, 6] = (wi + jk) * s; m[:, 10] = (ww - ii - jj + kk) * s
    m[:, 12:15] = ps[:, :3]
    m[:, 15] = 1
    return m


def _np_aabb(m, mn, mx):
    """Row A2 with all 8 corners, corner order of src/ecs.rs:151-159."""
    lo = np.full((len(m), 3), np.finfo(F).max, F)
    hi = np.full((len(m), 3), np.finfo(F).min, F)
    for cy in (mn[:, 1], mx[:, 1]):
        for cz in (mn[:, 2], mx[:, 2]):
            for cx in (mn[:, 0], mx[:, 0]):
                for a in range(3):
                    v = m[:, a] * cx
                    v = m[:, 4 + a] * cy + v
                    v = m[:, 8 + a] * cz + v
                    v = m[:, 12 + a] * F(1) + v
                    v = v / F(1)
                    lo[:, a] = np.minimum(lo[:, a], v)
                    hi[:, a] = np.maximum(hi[:, a], v)
    c, h = (hi + lo) / F(2), (hi - lo) / F(2)
    return c - h, c + h


def _np_culled(mins, maxs, planes):
    h, c = (maxs - mins) * F(0.5), (mins + maxs) * F(0.5)
    out = np.zeros(len(h), bool)
    for p in planes:
        a = np.abs(p[:3])
        e = (h[:, 0] * a[0] + h[:, 1] * a[1]) + h[:, 2] * a[2]
        s = (p[0] * c[:, 0] + p[2] * c[:, 2]) + (p[1] * c[:, 1] + p[3] * F(1))
        out |= (s - e) > 0
    return out


def test_numpy_restatement_matches_c_oracle_on_config1():
    n = 10_000
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, n, 60.0)
    meshes = scene.synth_mesh_table()
    views = scene.scene_views(5)
    pos, quat, scale = scene.to_components(ps, rot)
    ref = O.frame(pos, quat, scale, mesh, meshes, views, scene.CAMERA_POS)
    m = _np_model(ps, rot)
    assert np.array_equal(m, ref["model"])
    mins, maxs = _np_aabb(m, meshes["aabb_min"][mesh], meshes["aabb_max"][mesh])
    assert np.array_equal(np.concatenate([mins, maxs], axis=1), ref["aabb"])
    for k in range(5):
        culled = _np_culled(mins, maxs, views[k])
        assert np.array_equal(culled.astype(np.uint8), ref["flags"][k])
        vis = np.nonzero(~culled)[0].astype(np.uint32)
        assert np.array_equal(vis, ref["visible"][k])
        d = scene.CAMERA_POS[None, :] - ps[vis, :3]
        dist = np.sqrt((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2])
        lod = ((dist > F(10.0)) & (meshes["n_lods"][mesh[vis]] > 1)).astype(np.int64)
        length = meshes["lod"][mesh[vis], lod, 1].astype(np.uint64)
        cmds = ref["cmds"][k]
        assert np.array_equal(cmds["indexCount"], length.astype(np.uint32))
        first = np.concatenate([[0], np.cumsum(length)[:-1]]).astype(np.uint64) & np.uint64(0xFFFFFFFF)
        assert np.array_equal(cmds["firstIndex"], first.astype(np.uint32))
        assert np.array_equal(cmds["firstInstance"], vis)
        assert np.all(cmds["instanceCount"] == 1)
        assert np.array_equal(cmds["vertexOffset"], meshes["vertex_offset"][mesh[vis]])
        assert ref["totals"][k] == length.sum()


# ---- A3: plane test boundary (cull_pipeline.rs:113, strict >) ------------------------
def test_box_touching_a_plane_is_visible():
    planes = np.zeros((6, 4), F)
    planes[:, 3] = -1.0  # five planes that never cull
    planes[0] = [1, 0, 0, -2]  # outside iff x - 2 > e
    assert not O.coarse_culled([0, -1, -1], [2, 1, 1], planes)  # s - e = -1 - 1
    assert not O.coarse_culled([2, -1, -1], [4, 1, 1], planes)  # s - e = 1 - 1 = 0: NOT culled
    assert O.coarse_culled(np.nextafter(F(2), F(3)) + np.array([0, -1, -1], F), [4, 1, 1], planes)


def test_pick_lod_threshold_is_strict():
    cam = [0, 0, 0]
    assert O.pick_lod(cam, [10, 0, 0], 3) == 0  # distance == 10: LOD 0 (helpers.rs:6 uses >)
    assert O.pick_lod(cam, [np.nextafter(F(10), F(11)), 0, 0], 3) == 1
    assert O.pick_lod(cam, [500, 0, 0], 1) == 0  # single LOD


# ---- A5: planes ---------------------------------------------------------------------
def test_default_camera_sees_forward_only():
    planes = scene.camera_view()
    inside = lambda p: not O.coarse_culled(np.array(p, F) - F(0.1), np.array(p, F) + F(0.1), planes)
    assert inside([0, 1, 12])  # in front of (0,1,2) looking down +z
    assert not inside([0, 1, -10])  # behind
    assert not inside([0, 1, 150])  # past far = 100
    assert not inside([200, 1, 12])  # far off to the side


def test_library_host_helpers_match_oracle_planes():
    _, _, pv, planes = O.project_camera(16 / 9, scene.CAMERA_POS, scene.CAMERA_ROT)
    assert np.array_equal(scene.camera_view(), planes)
    assert np.array_equal(O.frustum_planes(pv), planes)
    for lp in scene.LIGHT_POS:
        rot = scene.quat_look_at_lh(-lp.astype(np.float64))
        assert np.array_equal(scene.light_view(lp), O.light_view(lp, rot)[3])


# ---- generators and golden vectors ----------------------------------------------------
def test_scene_generator_twins_agree():
    ps, rot, mesh = scene.synth_entities(scene.SEED, 777, 5000, 250.0)
    pos, quat, scale, mesh2 = O.synth_entities(scene.SEED, 777, 5000, 250.0, scene.N_MESHES)
    assert np.array_equal(ps[:, :3], pos) and np.array_equal(ps[:, 3], scale)
    assert np.array_equal(rot, quat) and np.array_equal(mesh, mesh2)


def test_golden_vectors_config1():
    import sys
    sys.path.insert(0, os.path.dirname(GOLDEN))
    import make_golden

    g = np.load(GOLDEN)
    now = make_golden.build()
    for key in g.files:
        assert np.array_equal(g[key], now[key]), key


def test_empty_and_bad_inputs():
    meshes = scene.synth_mesh_table()
    views = scene.scene_views(2)
    e = np.zeros((0, 3), F)
    r = O.frame(e, np.zeros((0, 4), F), np.zeros(0, F), np.zeros(0, np.uint32), meshes, views, scene.CAMERA_POS)
    assert r["counts"].tolist() == [0, 0] and r["totals"].tolist() == [0, 0]
    with pytest.raises(ValueError):
        O.frame(np.zeros((1, 3), F), np.array([[0, 0, 0, 1]], F), np.ones(1, F), np.array([99], np.uint32), meshes, views, scene.CAMERA_POS)


def test_threads_do_not_change_results():
    ps, rot, mesh = scene.synth_entities(3, 0, 20_000, 80.0)
    pos, quat, scale = scene.to_components(ps, rot)
    meshes, views = scene.synth_mesh_table(), scene.scene_views(3)
    a = O.frame(pos, quat, scale, mesh, meshes, views, scene.CAMERA_POS, n_threads=1)
    b = O.frame(pos, quat, scale, mesh, meshes, views, scene.CAMERA_POS, n_threads=5)
    assert np.array_equal(a["model"], b["model"]) and np.array_equal(a["flags"], b["flags"])
    for k in range(3):
        assert np.array_equal(a["cmds"][k], b["cmds"][k])


# ---- stage 2: per-triangle cull (generate_work.comp) + compact_draw_stream ------------------
@pytest.mark.parametrize("n,n_views,threads,want", [
    (0, 2, 4, ("model", "aabb", "flags", "cmds", "visible")),
    (1, 1, 1, ("model", "aabb", "flags", "cmds", "visible")),
    (1023, 3, 8, ("model", "aabb", "flags", "cmds", "visible")),
    (20_000, 5, 1, ("model", "aabb", "flags", "cmds", "visible")),
    (20_000, 5, 7, ("model", "aabb", "flags", "cmds", "visible")),
    (50_001, 8, 16, ("cmds", "visible")),
    (50_001, 2, 3, ("flags",)),
])
def test_fused_schedule_is_bit_identical_to_the_reference_shaped_passes(n, n_views, threads, want):
    """oracle_frame_fused (the stronger CPU baseline of bench.py) only reschedules the same per-entity functions:
    every output equals oracle_frame's, for any thread count and any subset of outputs, draw_index_base included."""
    ps, rot, mesh = scene.synth_entities(scene.SEED + 3, 0, n, 45.0)
    meshes = scene.synth_mesh_table()
    views = scene.scene_views(n_views)
    pos, quat, scale = scene.to_components(ps, rot)
    a = O.frame(pos, quat, scale, mesh, meshes, views, scene.CAMERA_POS, draw_index_base=11, want=want)
    b = O.frame(pos, quat, scale, mesh, meshes, views, scene.CAMERA_POS, draw_index_base=11, want=want, n_threads=threads, fused=True)
    assert np.array_equal(a["counts"], b["counts"]) and np.array_equal(a["totals"], b["totals"])
    for key in ("model", "aabb", "flags"):
        if key in want:
            assert a[key].tobytes() == b[key].tobytes(), key
    for key in ("cmds", "visible"):
        if key in want:
            for k in range(n_views):
                assert np.array_equal(a[key][k], b[key][k]), (key, k)


def test_fused_schedule_wraps_first_index_like_the_serial_emit():
    """index_len sums beyond 2^32: firstIndex is a u32 that wraps, the index totals stay exact -- across thread ranges too."""
    n = 6000
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, n, 6.0)
    meshes = scene.synth_mesh_table().copy()
    meshes["lod"][:, :, 1] = 3 * 11_200_000
    views = scene.scene_views(2)
    pos, quat, scale = scene.to_components(ps, rot)
    a = O.frame(pos, quat, scale, mesh, meshes, views, scene.CAMERA_POS, want=("cmds",))
    b = O.frame(pos, quat, scale, mesh, meshes, views, scene.CAMERA_POS, want=("cmds",), n_threads=5, fused=True)
    assert int(a["totals"][1]) > 2**32 and np.array_equal(a["totals"], b["totals"])
    for k in range(2):
        assert np.array_equal(a["cmds"][k], b["cmds"][k])


def _one_cmd(index_len, first_instance=0, first_index=0):
    c = np.zeros(1, O.CMD_DTYPE)
    c[0] = (index_len, 1, first_index, 0, first_instance)
    return c


def test_tri_cull_kats():
    """Identity model and pv: clip == vertex (w = 1).  The determinant of (v0.xyw, v1.xyw, v2.xyw)
    is twice the signed area in xy: > 0 is culled (generate_work.comp:143), a triangle with all
    three x > 1 is culled (:152-155), one straddling the edge is kept."""
    eye = np.eye(4, dtype=np.float32).reshape(-1)
    pos = np.array([
        [0, 0, 0.5], [0, 0.5, 0.5], [0.5, 0, 0.5],     # det < 0 -> kept
        [2, 0, 0.5], [2, 0.5, 0.5], [2.5, 0, 0.5],     # same winding, all x > 1 -> culled
        [0.9, 0, 0.5], [0.9, 0.5, 0.5], [1.5, 0, 0.5],  # straddles x = 1 -> kept
    ], np.float32)
    idx = np.array([0, 1, 2,  0, 2, 1,  3, 4, 5,  6, 7, 8], np.uint32)  # second triangle: flipped winding -> det > 0 -> culled
    r = O.tri_cull(_one_cmd(12), np.zeros(1, np.uint32), eye, eye, pos, idx, 12)
    assert r["count"] == 1 and r["kept_indices"] == 6
    assert r["cmds"][0]["indexCount"] == 6 and r["cmds"][0]["firstIndex"] == 0
    assert r["indices"][:6].tolist() == [0, 1, 2, 6, 7, 8]


def test_tri_cull_compacts_empty_commands_and_keeps_order():
    eye = np.eye(4, dtype=np.float32).reshape(-1)
    pos = np.array([[0, 0, 0.5], [0, 0.5, 0.5], [0.5, 0, 0.5]], np.float32)
    idx = np.array([0, 1, 2,  0, 2, 1], np.uint32)  # kept, culled
    cmds = np.zeros(3, O.CMD_DTYPE)
    cmds[0] = (3, 1, 0, 0, 10)  # kept triangle
    cmds[1] = (3, 1, 3, 0, 11)  # culled triangle -> indexCount 0 -> dropped (compact_draw_stream.comp:41)
    cmds[2] = (6, 1, 6, 0, 12)  # both -> one kept
    src = np.array([0, 3, 0], np.uint32)
    model = np.tile(eye, 3)
    r = O.tri_cull(cmds, src, model, eye, pos, idx, 12, draw_index_base=10)
    assert r["count"] == 2 and r["kept_indices"] == 6
    assert r["cmds"]["firstInstance"].tolist() == [10, 12]
    assert r["cmds"]["indexCount"].tolist() == [3, 3]
    assert r["indices"][0:3].tolist() == [0, 1, 2] and r["indices"][6:9].tolist() == [0, 1, 2]


def test_tri_cull_matches_numpy_restatement():
    """Independent numpy restatement of the shader arithmetic (float32, same operation order)
    over a random scene through the whole frame -> tri pipeline."""
    from renderer_b200 import scene as S
    meshes = S.synth_mesh_table()
    positions, indices = S.synth_geometry(meshes)
    n = 300
    ps, rot, mesh = S.synth_entities(S.SEED, 0, n, 12.0, len(meshes))
    pos, quat, scale = S.to_components(ps, rot)
    _, _, pv, planes = O.project_camera(16.0 / 9.0, S.CAMERA_POS, S.CAMERA_ROT)
    ref = O.frame(pos, quat, scale, mesh, meshes, planes[None], S.CAMERA_POS)
    cmds = ref["cmds"][0]
    src = O.src_first_index(pos, mesh, cmds, meshes, S.CAMERA_POS)
    total = int(ref["totals"][0])
    r = O.tri_cull(cmds, src, ref["model"], pv, positions, indices, max(total, 3))
    f = np.float32

    def mul(m, v):  # GLSL mat4 * vec4, columns left to right; m column-major [16], v [k][4]
        out = np.empty_like(v)
        for row in range(4):
            a = f(m[0 + row]) * v[:, 0]
            a = a + f(m[4 + row]) * v[:, 1]
            a = a + f(m[8 + row]) * v[:, 2]
            a = a + f(m[12 + row]) * v[:, 3]
            out[:, row] = a
        return out

    kept_total, n_out = 0, 0
    with np.errstate(all="ignore"):
        for c, cmd in enumerate(cmds):
            nt = int(cmd["indexCount"]) // 3
            tri = indices[int(src[c]) : int(src[c]) + 3 * nt].reshape(nt, 3)
            model = ref["model"].reshape(-1, 16)[int(cmd["firstInstance"])]
            clip = []
            for k in range(3):
                v = positions[int(cmd["vertexOffset"]) + tri[:, k].astype(np.int64)]
                v4 = np.concatenate([v, np.ones((nt, 1), np.float32)], axis=1)
                clip.append(mul(pv, mul(model, v4)))
            a, b, d = clip
            t0 = b[:, 1] * d[:, 3] - d[:, 1] * b[:, 3]
            t1 = a[:, 1] * d[:, 3] - d[:, 1] * a[:, 3]
            t2 = a[:, 1] * b[:, 3] - b[:, 1] * a[:, 3]
            det = (a[:, 0] * t0 - b[:, 0] * t1) + d[:, 0] * t2
            x = [q[:, 0] / q[:, 3] for q in clip]
            y = [q[:, 1] / q[:, 3] for q in clip]
            culled = det > 0
            culled |= (x[0] < -1) & (x[1] < -1) & (x[2] < -1)
            culled |= (x[0] > 1) & (x[1] > 1) & (x[2] > 1)
            culled |= (y[0] < -1) & (y[1] < -1) & (y[2] < -1)
            culled |= (y[0] > 1) & (y[1] > 1) & (y[2] > 1)
            keep = tri[~culled].reshape(-1)
            if len(keep):
                assert r["cmds"][n_out]["firstInstance"] == cmd["firstInstance"]
                assert r["cmds"][n_out]["indexCount"] == len(keep)
                fi = int(cmd["firstIndex"])
                assert np.array_equal(r["indices"][fi : fi + len(keep)], keep)
                n_out += 1
                kept_total += len(keep)
    assert n_out == r["count"] and kept_total == r["kept_indices"] and n_out > 0


def test_as_instance_records():
    """acceleration_strucures.rs:419-449: 3x4 row-major transform, custom index | 0xFF mask, flags, BLAS address."""
    m = O.model_matrix([1.0, 2.0, 3.0], [0.0, 0.0, 0.0, 1.0], 2.0)
    r = O.as_instances(np.tile(m, 2), [1, 0], np.array([0x1111, 0x2222_0000_0000], np.uint64), draw_index_base=0x1000005)
    assert r["transform"][0].tolist() == [[2, 0, 0, 1], [0, 2, 0, 2], [0, 0, 2, 3]]
    assert r["custom_index_and_mask"].tolist() == [0xFF000005, 0xFF000006]  # 24-bit draw index, mask 0xFF
    assert r["sbt_offset_and_flags"].tolist() == [1 << 24, 1 << 24]
    assert r["blas"].tolist() == [0x2222_0000_0000, 0x1111]
