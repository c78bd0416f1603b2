"""Committed golden vectors for stage 2 and the per-mesh instanced mode
(tests/golden/stage2_per_mesh.npz, made by make_golden_stage2.py from the oracle).
CPU: the oracle still reproduces them.  GPU: librcull produces them through the C ABI."""
import os
import sys

import numpy as np
import pytest

from renderer_b200 import scene

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden", "stage2_per_mesh.npz")
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden_stage2 as G  # noqa: E402


def check_stage2(g, count, kept, cmds, index_buffer):
    assert (int(count), int(kept)) == (int(g["tri_count"]), int(g["tri_kept"]))
    assert np.array_equal(cmds, g["tri_cmds"])
    owned = G.owned_indices(g["tri_cmds"], index_buffer)
    assert len(owned) == int(g["tri_owned_len"])
    assert np.array_equal(owned[: G.HEAD], g["tri_owned_head"])
    assert G.sha(owned) == str(g["tri_owned_sha256"])


def check_per_mesh(g, k, cmds, visible):
    assert np.array_equal(cmds, g[f"pm_cmds_{k}"]), f"view {k} per-mesh commands"
    assert np.array_equal(visible, g[f"pm_visible_{k}"]), f"view {k} bucketed visible list"


def test_oracle_reproduces_the_fixture():
    g = np.load(GOLDEN)
    now = G.build()
    assert sorted(g.files) == sorted(now.keys())
    for key in g.files:
        assert np.array_equal(g[key], now[key]), key


def test_fixture_is_self_consistent():
    g = np.load(GOLDEN)
    cmds = g["tri_cmds"]
    assert len(cmds) == int(g["tri_count"]) <= int(g["tri_entity_count"][0])
    assert int(cmds["indexCount"].sum()) == int(g["tri_kept"]) == int(g["tri_owned_len"])
    assert np.all(cmds["indexCount"] % 3 == 0) and np.all(cmds["indexCount"] > 0)  # compact_draw_stream.comp:41
    assert np.all(np.diff(cmds["firstInstance"].astype(np.int64)) > 0)  # stable, ascending draw_index
    for k in range(len(g["pm_views"])):
        c, v = g[f"pm_cmds_{k}"], g[f"pm_visible_{k}"]
        assert int(c["instanceCount"].sum()) == len(v) and np.all(c["instanceCount"] > 0)
        assert np.array_equal(c["firstInstance"], np.concatenate([[0], np.cumsum(c["instanceCount"])[:-1]]))
        assert v.min() >= int(g["pm_base"]) and v.max() < int(g["pm_base"]) + int(g["pm_n"])


@pytest.mark.gpu
def test_gpu_stage2_matches_the_fixture():
    from renderer_b200 import CullContext

    g = np.load(GOLDEN)
    meshes = g["meshes"]
    positions, indices = scene.synth_geometry(meshes)
    n = int(g["tri_n"])
    ps, rot, mesh = scene.synth_entities(int(g["seed"]), 0, n, float(g["tri_half_box"]), len(meshes))
    cap = max(int(g["tri_entity_total"][0]), 3)
    with CullContext(n, max_views=1, max_meshes=len(meshes), capacity_tri_indices=cap) as ctx:
        ctx.set_mesh_table(meshes)
        ctx.set_geometry(positions, indices)
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(g["tri_view"], g["cam_pos"], n=n)
        counts, totals = ctx.counts()
        assert counts[0] == g["tri_entity_count"][0] and totals[0] == g["tri_entity_total"][0]
        ctx.tri_cull(g["tri_pv"])
        count, kept = ctx.tri_counts()
        check_stage2(g, count, kept, ctx.tri_cmds(count), ctx.tri_indices(0, cap))


@pytest.mark.gpu
def test_gpu_per_mesh_mode_matches_the_fixture():
    from renderer_b200 import PER_MESH_INSTANCED, CullContext

    g = np.load(GOLDEN)
    meshes, views = g["meshes"], g["pm_views"]
    n = int(g["pm_n"])
    ps, rot, mesh = scene.synth_entities(int(g["seed"]), 0, n, float(g["pm_half_box"]), len(meshes))
    with CullContext(n, max_views=len(views), max_meshes=len(meshes), output_mode=PER_MESH_INSTANCED) as ctx:
        ctx.set_mesh_table(meshes)
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(views, g["cam_pos"], draw_index_base=int(g["pm_base"]))
        counts, totals = ctx.counts()
        for k in range(len(views)):
            assert counts[k] == len(g[f"pm_cmds_{k}"]) and totals[k] == len(g[f"pm_visible_{k}"])
            check_per_mesh(g, k, ctx.cmds(k, int(counts[k])), ctx.visible(k, int(totals[k])))


@pytest.mark.gpu
def test_gpu_slot_counts_are_the_instance_counts_of_every_slot():
    """rcull_get_slot_counts / rcull_download_slot_counts: what a range-shard merge of the per-mesh lists
    all_gathers (shard.plan_merge_per_mesh) -- every slot's survivors, empty slots included."""
    from renderer_b200 import PER_MESH_INSTANCED, CullContext, RcullError

    g = np.load(GOLDEN)
    meshes, views = g["meshes"], g["pm_views"]
    n, n_slots = int(g["pm_n"]), len(meshes) * 6
    ps, rot, mesh = scene.synth_entities(int(g["seed"]), 0, n, float(g["pm_half_box"]), len(meshes))
    with CullContext(n, max_views=len(views), max_meshes=len(meshes), output_mode=PER_MESH_INSTANCED) as ctx:
        ctx.set_mesh_table(meshes)
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(views, g["cam_pos"], draw_index_base=int(g["pm_base"]))
        for k in range(len(views)):
            got = ctx.slot_counts(k, n_slots)
            assert np.array_equal(got[got > 0], g[f"pm_cmds_{k}"]["instanceCount"])
            assert int(got.sum()) == len(g[f"pm_visible_{k}"])
        with pytest.raises(RcullError):
            ctx.slot_counts(0, n_slots + 1)
    with CullContext(n, max_views=1, max_meshes=len(meshes)) as ctx:  # per-entity mode has no slots
        ctx.set_mesh_table(meshes)
        with pytest.raises(RcullError):
            ctx.slot_counts(0, n_slots)


def merge_per_mesh_parts(parts, meshes):
    """parts[r] = (bucketed visible list, slot counts) of rank r for one view -> (merged commands, merged visible list)."""
    from renderer_b200.shard import per_mesh_commands, per_mesh_permutation, plan_merge_per_mesh

    plan = plan_merge_per_mesh(np.stack([p[1] for p in parts]))
    staged = np.concatenate([p[0] for p in parts])
    return per_mesh_commands(meshes, plan), staged[per_mesh_permutation(plan)], plan.slot_counts.astype(np.uint32)


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 3])
def test_gpu_per_mesh_range_shards_merge_to_the_single_range_output(world):
    """SURVEY 8e, per-mesh mode: `world` contiguous ranges run one after the other on this GPU (global draw
    indices, rcull_get_slot_counts), merged on the host with shard.plan_merge_per_mesh -- commands and bucketed
    visible lists equal what one context produces for the whole range."""
    from renderer_b200 import PER_MESH_INSTANCED, CullContext
    from renderer_b200.shard import shard_range

    meshes, views, n = scene.synth_mesh_table(), scene.scene_views(2), 40_001
    n_slots = len(meshes) * 6
    ps, rot, mesh = scene.synth_entities(scene.SEED + 9, 0, n, 35.0, len(meshes))

    def run(lo, hi):
        with CullContext(hi - lo, max_views=2, max_meshes=len(meshes), output_mode=PER_MESH_INSTANCED) as ctx:
            ctx.set_mesh_table(meshes)
            ctx.upload_entities(ps[lo:hi], rot[lo:hi], mesh[lo:hi])
            ctx.frame(views, scene.CAMERA_POS, n=hi - lo, draw_index_base=lo)
            counts, totals = ctx.counts()
            return [(ctx.cmds(k, int(counts[k])), ctx.visible(k, int(totals[k])), ctx.slot_counts(k, n_slots)) for k in range(2)]

    whole = run(0, n)
    parts = [run(*shard_range(n, world, r)) for r in range(world)]
    for k in range(2):
        cmds, visible, slot_counts = merge_per_mesh_parts([(parts[r][k][1], parts[r][k][2]) for r in range(world)], meshes)
        assert len(whole[k][1]) > 0
        assert np.array_equal(visible, whole[k][1]), f"view {k} merged visible list"
        assert np.array_equal(cmds, whole[k][0]), f"view {k} merged commands"
        assert np.array_equal(slot_counts, whole[k][2])
