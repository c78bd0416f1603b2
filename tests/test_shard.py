"""CPU tests of the range-shard merge: planning arithmetic, and the
exchange itself on 2 (and 3, ragged) gloo ranks.  Each rank's per-shard draw lists
come from the oracle with LOCAL firstIndex and GLOBAL draw indices -- what one
GPU's rcull_frame produces for its range -- and the merged lists must equal the
single-range lists byte for byte."""
import os
import socket

import numpy as np
import pytest

import oracle_lib as O
from renderer_b200 import scene
from renderer_b200.shard import (PerMeshShardMerger, ShardMerger, per_mesh_commands, per_mesh_permutation, plan_merge,
                                  plan_merge_per_mesh, shard_range)

N_TOTAL, N_VIEWS, HALF_BOX = 30_001, 3, 40.0


def test_shard_range_covers_everything_once():
    for n, w in [(0, 1), (1, 4), (10, 3), (30_001, 8), (128 * 1024 * 1024, 8)]:
        spans = [shard_range(n, w, r) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        sizes = [hi - lo for lo, hi in spans]
        assert max(sizes) - min(sizes) <= 1


def test_plan_merge_prefixes():
    counts = np.array([[3, 0], [0, 5], [7, 1]])
    totals = np.array([[30, 0], [0, 50], [2**33, 1]], dtype=np.uint64)
    p = plan_merge(counts, totals)
    assert p.count_prefix.tolist() == [[0, 0], [3, 0], [3, 5]]
    assert p.index_prefix.tolist() == [[0, 0], [30, 0], [30, 50]]
    assert p.counts.tolist() == [10, 6] and p.totals.tolist() == [30 + 2**33, 51]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _scene():
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, N_TOTAL, HALF_BOX)
    pos, quat, scale = scene.to_components(ps, rot)
    return pos, quat, scale, mesh, scene.synth_mesh_table(), scene.scene_views(N_VIEWS)


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pos, quat, scale, mesh, meshes, views = _scene()
        lo, hi = shard_range(N_TOTAL, world, rank)
        mine = O.frame(pos[lo:hi], quat[lo:hi], scale[lo:hi], mesh[lo:hi], meshes, views, scene.CAMERA_POS, draw_index_base=lo)
        merger = ShardMerger(N_VIEWS, N_TOTAL, torch.device("cpu"))
        plan = merger.gather_meta(mine["counts"].astype(np.int64), mine["totals"].astype(np.int64))
        counts_all = merger._meta_all.numpy()[:, :N_VIEWS].astype(np.uint64)
        for k in range(N_VIEWS):  # what rcull_rebase_into does on the device
            off, n = int(plan.count_prefix[rank, k]), int(mine["counts"][k])
            cmds = mine["cmds"][k].copy()
            cmds["firstIndex"] = (cmds["firstIndex"].astype(np.uint64) + plan.index_prefix[rank, k]).astype(np.uint32)
            merger.merged_cmds[k][off * 5:(off + n) * 5] = torch.from_numpy(cmds.view(np.int32).reshape(-1).copy())
            merger.merged_visible[k][off:off + n] = torch.from_numpy(mine["visible"][k].view(np.int32).copy())
        merger.exchange(counts_all)
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), counts=plan.counts, totals=plan.totals, recv=merger.bytes_received,
                 **{f"cmds{k}": merger.merged_cmds[k][: int(plan.counts[k]) * 5].numpy() for k in range(N_VIEWS)},
                 **{f"vis{k}": merger.merged_visible[k][: int(plan.counts[k])].numpy() for k in range(N_VIEWS)})
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_merged_lists_equal_single_range_lists(world, tmp_path):
    import torch.multiprocessing as mp

    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    pos, quat, scale, mesh, meshes, views = _scene()
    ref = O.frame(pos, quat, scale, mesh, meshes, views, scene.CAMERA_POS)
    own_total = 0
    for rank in range(world):  # every rank holds the complete merged lists
        g = np.load(os.path.join(tmp_path, f"rank{rank}.npz"))
        assert np.array_equal(g["counts"], ref["counts"]) and np.array_equal(g["totals"], ref["totals"])
        for k in range(N_VIEWS):
            assert np.array_equal(g[f"cmds{k}"].view(np.uint32), ref["cmds"][k].view(np.uint32).reshape(-1)), (rank, k)
            assert np.array_equal(g[f"vis{k}"].view(np.uint32), ref["visible"][k]), (rank, k)
        own_total += int(ref["counts"].sum()) * 24 - int(g["recv"])
    assert own_total == int(ref["counts"].sum()) * 24  # received bytes = 24 * (all - own), summed over ranks


# ---- per-mesh instanced mode over range shards (last paragraph) ----------------------
def test_plan_merge_per_mesh_places_every_bucket():
    c = np.array([[2, 0, 1], [0, 3, 1]])  # [world, slots]
    p = plan_merge_per_mesh(c)
    assert p.slot_counts.tolist() == [2, 3, 2] and p.slot_start.tolist() == [0, 2, 5] and p.total == 7
    assert p.dest.tolist() == [[0, 2, 5], [2, 2, 6]] and p.src.tolist() == [[0, 2, 2], [0, 0, 3]]
    assert per_mesh_permutation(p).tolist() == [0, 1, 3, 4, 5, 2, 6]
    rng = np.random.default_rng(3)
    c = rng.integers(0, 4, size=(5, 40)) * (rng.random((5, 40)) < 0.5)
    p = plan_merge_per_mesh(c)
    perm = per_mesh_permutation(p)
    assert sorted(perm.tolist()) == list(range(int(c.sum())))  # a permutation: every staged entry lands once
    slot_of_staged = np.concatenate([np.repeat(np.arange(40), row) for row in c])
    rank_of_staged = np.repeat(np.arange(5), c.sum(axis=1))
    merged_slot, merged_rank = slot_of_staged[perm], rank_of_staged[perm]
    assert np.all(np.diff(merged_slot) >= 0)  # slot-major
    same = np.diff(merged_slot) == 0
    assert np.all(np.diff(merged_rank)[same] >= 0)  # ranks in order inside a slot
    assert np.all(np.diff(perm)[same & (np.diff(merged_rank) == 0)] == 1)  # a rank's bucket stays contiguous and in order


def _slot_counts(pos, mesh, flags, meshes, n_slots):
    vis = np.nonzero(flags == 0)[0]
    lod = np.array([O.pick_lod(scene.CAMERA_POS, pos[i], int(meshes["n_lods"][mesh[i]])) for i in vis], np.int64)
    return np.bincount(mesh[vis].astype(np.int64) * 6 + lod, minlength=n_slots)


def _per_mesh_worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pos, quat, scale, mesh, meshes, views = _scene()
        n_slots = len(meshes) * 6
        lo, hi = shard_range(N_TOTAL, world, rank)
        mine = O.frame(pos[lo:hi], quat[lo:hi], scale[lo:hi], mesh[lo:hi], meshes, views, scene.CAMERA_POS, want=("flags",))
        out = {}
        for k in range(N_VIEWS):  # what one GPU's per-mesh frame produces for its range: bucketed list, global draw indices
            _, vis = O.emit_per_mesh(pos[lo:hi], mesh[lo:hi], mine["flags"][k], meshes, scene.CAMERA_POS, 10.0, lo)
            merger = PerMeshShardMerger(n_slots, N_TOTAL, torch.device("cpu"))
            plan = merger.gather_counts(_slot_counts(pos[lo:hi], mesh[lo:hi], mine["flags"][k], meshes, n_slots))
            merged = merger.merge_visible(torch.from_numpy(vis.view(np.int32).copy()))
            out[f"vis{k}"] = merged.numpy().view(np.uint32)
            out[f"cmds{k}"] = per_mesh_commands(meshes, plan)
        np.savez(os.path.join(out_dir, f"pm_rank{rank}.npz"), **out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_per_mesh_merged_lists_equal_single_range_lists(world, tmp_path):
    import torch.multiprocessing as mp

    mp.spawn(_per_mesh_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    pos, quat, scale, mesh, meshes, views = _scene()
    ref = O.frame(pos, quat, scale, mesh, meshes, views, scene.CAMERA_POS, want=("flags",))
    for k in range(N_VIEWS):
        want_cmds, want_vis = O.emit_per_mesh(pos, mesh, ref["flags"][k], meshes, scene.CAMERA_POS, 10.0, 0)
        for rank in range(world):  # every rank holds the complete merged output
            g = np.load(os.path.join(tmp_path, f"pm_rank{rank}.npz"))
            assert np.array_equal(g[f"vis{k}"], want_vis), (rank, k)
            assert np.array_equal(g[f"cmds{k}"], want_cmds), (rank, k)
