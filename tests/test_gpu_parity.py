"""GPU parity: librcull (through the C ABI) against the CPU oracle on the same
seeded inputs.  Integer/index outputs must be bit-exact; model matrices are
compared with == on floats (bit-exact up to the sign of zero, well inside the
1e-6 relative tolerance north_star allows)."""
import numpy as np
import pytest

import oracle_lib as O
from renderer_b200 import CullContext, RcullError, scene

pytestmark = pytest.mark.gpu

MESHES = scene.synth_mesh_table()


def run_and_compare(n, n_views, half_box, *, seed=scene.SEED, first=0, base=0, lod_distance=10.0, ctx=None,
                    meshes=MESHES, threads=8, entities=None, equal_nan=False):
    ps, rot, mesh = scene.synth_entities(seed, first, n, half_box, len(meshes)) if entities is None else entities
    views = scene.scene_views(n_views)
    own = ctx is None
    if own:
        ctx = CullContext(max(n, 1), max_views=n_views, max_meshes=len(meshes), materialize_aabb=True,
                          materialize_flags=True, lod_distance=lod_distance)
        ctx.set_mesh_table(meshes)
    try:
        ctx.n = 0
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(views, scene.CAMERA_POS, n=n, draw_index_base=base)
        counts, totals = ctx.counts()
        pos, quat, scale = scene.to_components(ps, rot)
        ref = O.frame(pos, quat, scale, mesh, meshes, views, scene.CAMERA_POS, lod_distance=lod_distance,
                      draw_index_base=base, n_threads=threads)
        assert np.array_equal(counts[:n_views], ref["counts"])
        assert np.array_equal(totals[:n_views], ref["totals"])
        if n:
            assert np.array_equal(ctx.model(n), ref["model"], equal_nan=equal_nan)
            assert np.array_equal(ctx.aabb(n), ref["aabb"], equal_nan=equal_nan)
        for k in range(n_views):
            c = int(counts[k])
            if n:
                assert np.array_equal(ctx.flags(k, n), ref["flags"][k]), f"view {k} flags"
            assert np.array_equal(ctx.visible(k, c), ref["visible"][k]), f"view {k} visible"
            assert np.array_equal(ctx.cmds(k, c), ref["cmds"][k]), f"view {k} cmds"
        return counts[:n_views], n
    finally:
        if own:
            ctx.destroy()


def _with_special_components(values, seed):
    """Synthetic entities where every 5th one has one component (position, scale or quaternion) replaced."""
    n = 8192
    ps, rot, mesh = scene.synth_entities(scene.SEED + seed, 0, n, 60.0, len(MESHES))
    ps, rot = ps.copy(), rot.copy()
    rng = np.random.default_rng(seed)
    for i in range(0, n, 5):
        col, v = int(rng.integers(0, 8)), values[int(rng.integers(len(values)))]
        if col < 4:
            ps[i, col] = v
        else:
            rot[i, col - 4] = v
    return ps, rot, mesh


def test_extreme_finite_components():
    """Zero / negative / tiny / huge scales, denormal and far-away positions, non-unit and zero quaternions:
    no overflow, so every output stays bit-exact (denormals are not flushed: no -ftz, FMUL2 included).
    Non-finite components are outside the contract (include/rcull.h): the reference's generic 4x4 products
    turn inf * 0 into NaN where the reduced form of A1 keeps inf."""
    values = [0.0, -0.0, -2.0, 1e-30, 1e-42, -1e-40, 1e6, -1e6, 3.0, 1e9]
    run_and_compare(8192, 3, 60.0, entities=_with_special_components(values, 11))


def test_index_lengths_beyond_2_pow_25_use_wide_sums():
    """A mesh table with index_len >= 2^25: the per-warp index sums no longer fit 32 bits (128 x 2^25), the kernel
    takes its general variant (16-bit halves), firstIndex wraps mod 2^32 as the u32 field does and the index
    totals stay exact in 64 bits."""
    meshes = MESHES.copy()
    meshes["lod"][:, 0, 1] = (np.arange(len(meshes), dtype=np.uint32) + 1) * 3 * 11_200_000  # up to 537.6 M indices
    meshes["lod"][:, 1, 1] = 3 * 11_184_811  # just above 2^25 for every LOD 1
    counts, _ = run_and_compare(3000, 2, 20.0, meshes=meshes)
    assert counts[0] > 0


def test_config1_10k_single_camera():
    counts, n = run_and_compare(10_000, 1, 60.0)
    assert 0 < counts[0] < n


def test_config1_10k_five_views():
    run_and_compare(10_000, 5, 60.0)


# around the warp span (128), the tile (1920) and one tile per CTA of the persistent grid (148 x 1920: the
# first ticket drawn from the counter is tile 148)
@pytest.mark.parametrize("n", [0, 1, 31, 32, 33, 127, 128, 129, 1023, 1024, 1025, 1919, 1920, 1921, 3841, 5000,
                               148 * 1920 - 1, 148 * 1920, 148 * 1920 + 1, 149 * 1920 + 7])
def test_ragged_sizes(n):
    run_and_compare(n, 2, 30.0)


@pytest.mark.parametrize("n_views", [1, 2, 3, 4, 5, 6, 7, 8])
def test_every_view_count(n_views):
    run_and_compare(3000, n_views, 50.0)


def test_all_visible_dense_lists():
    # everything inside the light frustum at the origin: exercises full-tile emission
    counts, n = run_and_compare(20_000, 2, 4.0)
    assert counts[1] > 0.5 * n


def test_none_visible():
    meshes = MESHES
    n = 3000
    ps, rot, mesh = scene.synth_entities(7, 0, n, 10.0)
    ps[:, 2] = -500.0  # far behind the camera
    with CullContext(n, max_views=1, max_meshes=len(meshes)) as ctx:
        ctx.set_mesh_table(meshes)
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(scene.scene_views(1), scene.CAMERA_POS)
        counts, totals = ctx.counts()
        assert counts[0] == 0 and totals[0] == 0


def test_config2_1m_camera():
    counts, n = run_and_compare(1_000_000, 1, 250.0)
    assert counts[0] > 0


def test_draw_index_base_and_lod_distance():
    run_and_compare(7000, 2, 40.0, base=123_456, lod_distance=25.5)
    run_and_compare(7000, 1, 40.0, lod_distance=0.0)
    run_and_compare(7000, 1, 40.0, lod_distance=1e9)


def test_context_reuse_many_frames():
    """Ticket / epoch bookkeeping across frames of different sizes on one context."""
    n_max = 40_000
    ctx = CullContext(n_max, max_views=3, max_meshes=len(MESHES), materialize_aabb=True, materialize_flags=True)
    ctx.set_mesh_table(MESHES)
    try:
        for i, (n, v, hb) in enumerate([(40_000, 3, 60.0), (1000, 1, 20.0), (33_333, 2, 80.0), (40_000, 3, 60.0), (5, 1, 5.0)]):
            run_and_compare(n, v, hb, seed=100 + i, ctx=ctx)
    finally:
        ctx.destroy()


def test_double_buffered_outputs():
    n = 6000
    ps, rot, mesh = scene.synth_entities(3, 0, n, 40.0)
    views = scene.scene_views(1)
    with CullContext(n, max_views=1, max_meshes=len(MESHES), n_buffers=2) as ctx:
        ctx.set_mesh_table(MESHES)
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(views, scene.CAMERA_POS, buffer_ix=0)
        c0, _ = ctx.counts(0)
        ps2 = ps.copy()
        ps2[:, :3] *= 0.5
        ctx.upload_entities(ps2, rot, mesh)
        ctx.frame(views, scene.CAMERA_POS, buffer_ix=1)
        c1, _ = ctx.counts(1)
        c0b, _ = ctx.counts(0)
        assert c0[0] == c0b[0] and c1[0] != c0[0]
        assert np.array_equal(ctx.model(n, buffer_ix=0)[:, 12:15], ps[:, :3])
        assert np.array_equal(ctx.model(n, buffer_ix=1)[:, 12:15], ps2[:, :3])


def test_capacity_visible_clamps_writes_not_count():
    n = 5000
    ps, rot, mesh = scene.synth_entities(5, 0, n, 4.0)
    views = scene.scene_views(2)[1:]
    with CullContext(n, max_views=1, max_meshes=len(MESHES), capacity_visible=100) as ctx:
        ctx.set_mesh_table(MESHES)
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(views, scene.CAMERA_POS)
        counts, _ = ctx.counts()
        assert counts[0] > 100  # the true survivor count is still reported
        pos, quat, scale = scene.to_components(ps, rot)
        ref = O.frame(pos, quat, scale, mesh, MESHES, views, scene.CAMERA_POS)
        assert counts[0] == ref["counts"][0]
        assert np.array_equal(ctx.cmds(0, 100), ref["cmds"][0][:100])


def test_errors_are_loud():
    with pytest.raises(RcullError):
        CullContext(0)
    with pytest.raises(RcullError):
        CullContext(10, max_views=9)
    with CullContext(100, max_views=1, max_meshes=4) as ctx:
        with pytest.raises(RcullError):  # frame before mesh table
            ctx.frame(scene.scene_views(1), scene.CAMERA_POS, n=10)
        with pytest.raises(RcullError):
            ctx.set_mesh_table(MESHES)  # 16 > max_meshes
        ctx.set_mesh_table(MESHES[:4])
        with pytest.raises(RcullError):
            ctx.frame(scene.scene_views(2), scene.CAMERA_POS, n=10)  # n_views > max_views
        with pytest.raises(RcullError):
            ctx.frame(scene.scene_views(1), scene.CAMERA_POS, n=101)  # n > capacity
        ps, rot, mesh = scene.synth_entities(1, 0, 100, 10.0, 16)  # mesh ids up to 15 > table of 4
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(scene.scene_views(1), scene.CAMERA_POS)
        with pytest.raises(RcullError):
            ctx.counts()
        bad = MESHES[:1].copy()
        bad["n_lods"] = 0
        with pytest.raises(RcullError):
            ctx.set_mesh_table(bad)


def test_device_generator_matches_host_twin():
    n, first = 50_000, 1_234_567
    ps, rot, mesh = scene.synth_entities(scene.SEED, first, n, 250.0)
    views = scene.scene_views(1)
    with CullContext(n, max_views=1, max_meshes=len(MESHES)) as a, CullContext(n, max_views=1, max_meshes=len(MESHES)) as b:
        for c in (a, b):
            c.set_mesh_table(MESHES)
        a.upload_entities(ps, rot, mesh)
        b.synth_entities(scene.SEED, first, n, 250.0, len(MESHES))
        a.frame(views, scene.CAMERA_POS)
        b.frame(views, scene.CAMERA_POS)
        assert np.array_equal(a.model(n), b.model(n))
        ca, ta = a.counts()
        cb, tb = b.counts()
        assert ca[0] == cb[0] and ta[0] == tb[0]
        assert np.array_equal(a.cmds(0, int(ca[0])), b.cmds(0, int(cb[0])))


def test_golden_vectors_config1():
    """The committed fixture (tests/golden/config1_10k.npz, made by make_golden.py)."""
    import os

    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "config1_10k.npz"))
    n, n_views = int(g["n"]), len(g["views"])
    ps, rot, mesh = scene.synth_entities(int(g["seed"]), 0, n, float(g["half_box"]))
    with CullContext(n, max_views=n_views, max_meshes=len(g["meshes"]), materialize_aabb=True, materialize_flags=True) as ctx:
        ctx.set_mesh_table(g["meshes"])
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(g["views"], g["cam_pos"])
        counts, totals = ctx.counts()
        assert np.array_equal(counts[:n_views], g["counts"]) and np.array_equal(totals[:n_views], g["totals"])
        assert np.array_equal(ctx.model(64), g["model_head"]) and np.array_equal(ctx.aabb(64), g["aabb_head"])
        flags = np.stack([ctx.flags(k, n) for k in range(n_views)])
        assert np.array_equal(np.packbits(flags, axis=1), g["flags_packed"])
        for k in range(n_views):
            assert np.array_equal(ctx.visible(k, int(counts[k])), g[f"visible_{k}"])
            assert np.array_equal(ctx.cmds(k, int(counts[k])), g[f"cmds_{k}"])


def _list_properties(ctx, k, count, total, meshes, base=0):
    """Size-independent checks of one view's draw list (used where the oracle is too slow)."""
    vis = ctx.visible(k, count)
    cmds = ctx.cmds(k, count)
    assert np.all(np.diff(vis.astype(np.int64)) > 0), "visible indices not strictly ascending"
    assert np.array_equal(cmds["firstInstance"], vis)
    assert np.all(cmds["instanceCount"] == 1)
    ic = cmds["indexCount"].astype(np.uint64)
    first = np.concatenate([[0], np.cumsum(ic)[:-1]]).astype(np.uint64) & np.uint64(0xFFFFFFFF)
    assert np.array_equal(cmds["firstIndex"], first.astype(np.uint32)), "firstIndex is not the running sum of indexCount"
    assert int(ic.sum()) == int(total)
    lens = set(int(x) for x in meshes["lod"][:, :, 1].reshape(-1))
    assert set(np.unique(cmds["indexCount"]).tolist()) <= lens
    return vis


@pytest.mark.parametrize("rank", [0, 1, 7])
def test_config3_16m_five_views_full_compare(rank):
    """BASELINE configs[2] at full size (rank 0), and ranks 1 and 7 of configs[4] (128 Mi entities in 8 range
    shards of 16 Mi: entity i of the shard is entity rank * 16 Mi + i of the scene, draw_index_base = rank * 16 Mi;
    firstInstance = gltfIndex, generate_work.comp:63-66; prefix sum cull_pipeline.rs:558): ALL five visible lists
    and command lists bit-equal to the oracle's, counts, index totals, list properties, idempotence; the model
    matrices (1 GiB) are compared on three slices."""
    import os

    n, n_views = 16 * 1024 * 1024, 5
    first = rank * n
    views = scene.scene_views(n_views)
    pos, quat, scale, mesh = O.synth_entities(scene.SEED, first, n, 600.0, len(MESHES))
    ref = O.frame(pos, quat, scale, mesh, MESHES, views, scene.CAMERA_POS, draw_index_base=first,
                  n_threads=os.cpu_count() or 8, want=("cmds", "visible"))
    with CullContext(n, max_views=n_views, max_meshes=len(MESHES)) as ctx:
        ctx.set_mesh_table(MESHES)
        ctx.synth_entities(scene.SEED, first, n, 600.0, len(MESHES))
        ctx.frame(views, scene.CAMERA_POS, draw_index_base=first)
        counts, totals = ctx.counts()
        assert np.array_equal(counts[:n_views], ref["counts"]) and np.array_equal(totals[:n_views], ref["totals"])
        for k in range(n_views):
            c = int(counts[k])
            assert np.array_equal(ctx.visible(k, c), ref["visible"][k]), f"rank {rank} view {k} visible"
            assert np.array_equal(ctx.cmds(k, c), ref["cmds"][k]), f"rank {rank} view {k} cmds"
            vis = _list_properties(ctx, k, c, totals[k], MESHES)
            assert c == 0 or (vis[0] >= first and int(vis[-1]) < first + n)
        # second frame on the same context: identical (idempotence, epoch/ticket reuse)
        ctx.frame(views, scene.CAMERA_POS, draw_index_base=first)
        c2, t2 = ctx.counts()
        assert np.array_equal(c2, counts) and np.array_equal(t2, totals)
        for lo in (0, 7_654_321, n - 300_000):
            m = 300_000
            want = O.frame(pos[lo:lo + m], quat[lo:lo + m], scale[lo:lo + m], mesh[lo:lo + m], MESHES, views, scene.CAMERA_POS,
                           n_threads=8, want=("model",))
            assert np.array_equal(ctx.model(m, first=lo), want["model"])


def test_range_shards_merge_to_the_single_gpu_lists():
    """SURVEY 8e on one device: two contexts own the two halves of the scene, each rebases
    its lists into its slot of a merged buffer (rcull_rebase_into); merged == unsharded."""
    import torch

    from renderer_b200.shard import plan_merge, shard_range

    n, n_views, world = 200_001, 3, 2
    views = scene.scene_views(n_views)
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, n, 60.0)
    with CullContext(n, max_views=n_views, max_meshes=len(MESHES)) as whole:
        whole.set_mesh_table(MESHES)
        whole.upload_entities(ps, rot, mesh)
        whole.frame(views, scene.CAMERA_POS)
        wc, wt = whole.counts()
        want_cmds = [whole.cmds(k, int(wc[k])) for k in range(n_views)]
        want_vis = [whole.visible(k, int(wc[k])) for k in range(n_views)]
    ctxs, counts, totals = [], [], []
    try:
        for r in range(world):
            lo, hi = shard_range(n, world, r)
            c = CullContext(hi - lo, max_views=n_views, max_meshes=len(MESHES))
            c.set_mesh_table(MESHES)
            c.upload_entities(ps[lo:hi], rot[lo:hi], mesh[lo:hi])
            c.frame(views, scene.CAMERA_POS, draw_index_base=lo)
            cc, tt = c.counts()
            ctxs.append(c), counts.append(cc[:n_views]), totals.append(tt[:n_views])
        plan = plan_merge(np.stack(counts), np.stack(totals))
        assert np.array_equal(plan.counts, wc[:n_views]) and np.array_equal(plan.totals, wt[:n_views])
        for k in range(n_views):
            m_cmds = torch.zeros(int(plan.counts[k]) * 5 + 5, dtype=torch.int32, device="cuda")
            m_vis = torch.zeros(int(plan.counts[k]) + 1, dtype=torch.int32, device="cuda")
            for r in range(world):
                off = int(plan.count_prefix[r, k])
                ctxs[r].rebase_into(k, int(plan.index_prefix[r, k]) & 0xFFFFFFFF, m_cmds.data_ptr() + off * 20, m_vis.data_ptr() + off * 4)
            torch.cuda.synchronize()
            got_cmds = m_cmds.cpu().numpy()[: int(plan.counts[k]) * 5].view(np.uint32)
            assert np.array_equal(got_cmds, want_cmds[k].view(np.uint32).reshape(-1)), k
            assert np.array_equal(m_vis.cpu().numpy()[: int(plan.counts[k])].view(np.uint32), want_vis[k]), k
    finally:
        for c in ctxs:
            c.destroy()


def _merged_rank_setup(shards, n_views, meshes, cap, cap_r, output_mode=None):
    """world = len(shards) ranks emulated in ONE process on one GPU: a context + merge per rank, windows named to each
    other directly (rcull_merge_set_peer_ptr).  Returns (contexts, merge handles)."""
    import ctypes as C

    from renderer_b200 import _lib

    lib = _lib.load()
    world = len(shards)
    ctxs, merges, windows = [], [], []
    kw = {} if output_mode is None else {"output_mode": output_mode}
    for r, (lo, hi, ps, rot, mesh) in enumerate(shards):
        c = CullContext(max(hi - lo, 1), max_views=n_views, max_meshes=len(meshes), **kw)
        c.set_mesh_table(meshes)
        c.upload_entities(ps[lo:hi], rot[lo:hi], mesh[lo:hi])
        c.n = hi - lo
        h = C.c_void_p()
        _lib.check(lib.rcull_merge_create(c._h, world, r, cap, cap_r, C.byref(h)))
        w = C.c_void_p()
        _lib.check(lib.rcull_merge_window(h, C.byref(w)))
        ctxs.append(c)
        merges.append(h)
        windows.append(w)
    for h in merges:
        for p in range(world):
            _lib.check(lib.rcull_merge_set_peer_ptr(h, p, windows[p]))
    return ctxs, merges


def _merged_lists(h, k, n_views, per_mesh=False, cap=None):
    import ctypes as C

    import torch

    from renderer_b200 import _lib
    from renderer_b200._lib import CMD_DTYPE
    from renderer_b200.shard import _wrap_device

    lib, dev = _lib.load(), torch.device("cuda", 0)
    a, b, c_, d = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
    _lib.check(lib.rcull_merge_lists(h, k, C.byref(a), C.byref(b), C.byref(c_), C.byref(d)))
    torch.cuda.synchronize()
    counts = _wrap_device(torch, c_.value, n_views, torch.int32, dev).cpu().numpy().view(np.uint32)
    totals = _wrap_device(torch, d.value, n_views, torch.int64, dev).cpu().numpy().view(np.uint64)
    n_cmd = int(counts[k]) if cap is None else min(int(counts[k]), cap)
    n_vis = int(totals[k]) if per_mesh else n_cmd
    cmds = _wrap_device(torch, a.value, max(n_cmd, 1) * 5, torch.int32, dev).cpu().numpy()[: n_cmd * 5].view(CMD_DTYPE)
    vis = _wrap_device(torch, b.value, max(n_vis, 1), torch.int32, dev).cpu().numpy()[:n_vis].view(np.uint32)
    return cmds, vis, int(counts[k]), int(totals[k])


@pytest.mark.parametrize("world,ce", [(2, False), (3, False), (2, True), (3, True)])
def test_record_merge_ranks_in_one_process(world, ce):
    """SURVEY 8e through the record path (rcull_merge_*): ragged range shards, every rank's frame kernel sends 8-byte
    records into every window, every rank expands them -- the merged lists on EVERY rank are byte-identical to the
    single-range oracle lists (firstIndex = running sum over the merged order, u32 wrap included).  Four frames, so
    both record regions and the epoch tags are reused; one rank has an empty shard.  The ranks share one GPU here, so
    all frames complete before any expand (no launch ever waits on another launch).  ce: the copy-engine transport
    (rcull_merge_set_transport(m, 1)): the frame kernels keep their records in their own windows and rcull_merge_send
    moves them with peer copies."""
    import torch

    from renderer_b200 import _lib
    from renderer_b200.shard import shard_range

    lib = _lib.load()
    n, n_views = 150_001, 3
    views = scene.scene_views(n_views)
    ctxs = merges = None
    try:
        for frame in range(4):
            ps, rot, mesh = scene.synth_entities(scene.SEED + frame, 0, n, 70.0 + 10 * frame, len(MESHES))
            pos, quat, scale = scene.to_components(ps, rot)
            ref = O.frame(pos, quat, scale, mesh, MESHES, views, scene.CAMERA_POS, n_threads=8, want=("cmds", "visible"))
            bounds = [shard_range(n, world, r) for r in range(world)]
            if world == 3:  # rank 1 owns nothing
                bounds = [(0, n // 3), (n // 3, n // 3), (n // 3, n)]
            if ctxs is None:
                shards = [(lo, hi, ps, rot, mesh) for lo, hi in bounds]
                ctxs, merges = _merged_rank_setup(shards, n_views, MESHES, cap=n, cap_r=n)
                if ce:
                    for r, h in enumerate(merges):
                        _lib.check(lib.rcull_merge_set_transport(h, 1))
                        # rank 0 copies at most 512 / 256 / all records of the three views: the rest is read from its window by the others
                        if r == 0:
                            bound = np.array([512, 1, 0], np.uint32)
                            _lib.check(lib.rcull_merge_set_send_bound(h, _lib.ptr(bound), 3))
                    with pytest.raises(RcullError, match="copy engines|transport"):
                        _lib.check(lib.rcull_merge_set_transport(merges[0], 2))
            else:
                for c, (lo, hi) in zip(ctxs, bounds):
                    if hi > lo:
                        c.upload_entities(ps[lo:hi], rot[lo:hi], mesh[lo:hi])
            for c, (lo, hi) in zip(ctxs, bounds):
                c.frame(views, scene.CAMERA_POS, n=hi - lo, draw_index_base=lo)
            torch.cuda.synchronize()
            if ce:
                side = torch.cuda.Stream()  # the sends go on a stream of their own: the library orders it behind the frame
                for h in merges:
                    _lib.check(lib.rcull_merge_send(h, side.cuda_stream))
                torch.cuda.synchronize()
                with pytest.raises(RcullError, match="no merged frame is waiting to be sent"):
                    _lib.check(lib.rcull_merge_send(merges[0], None))
            else:
                with pytest.raises(RcullError, match="copy engines"):
                    _lib.check(lib.rcull_merge_send(merges[0], None))
            for h in merges:
                _lib.check(lib.rcull_merge_expand(h, None))
            torch.cuda.synchronize()
            for h in merges:
                for k in range(n_views):
                    cmds, vis, cnt, tot = _merged_lists(h, k, n_views)
                    assert cnt == int(ref["counts"][k]) and tot == int(ref["totals"][k]), (frame, k)
                    assert np.array_equal(vis, ref["visible"][k]), (frame, k)
                    assert np.array_equal(cmds, ref["cmds"][k]), (frame, k)
        # protocol errors are loud: a third frame in flight, an expand with nothing to expand
        with pytest.raises(RcullError, match="nothing|no merged frame"):
            _lib.check(lib.rcull_merge_expand(merges[0], None))
        lo, hi = bounds[0]
        def frame_and_send(c, h, lo, hi):
            c.frame(views, scene.CAMERA_POS, n=hi - lo, draw_index_base=lo)
            if ce:
                _lib.check(lib.rcull_merge_send(h, None))

        if ce:  # a frame whose records have not been sent blocks the next one
            ctxs[0].frame(views, scene.CAMERA_POS, n=hi - lo, draw_index_base=lo)
            with pytest.raises(RcullError, match="have not been sent"):
                ctxs[0].frame(views, scene.CAMERA_POS, n=hi - lo, draw_index_base=lo)
            _lib.check(lib.rcull_merge_send(merges[0], None))
        else:
            frame_and_send(ctxs[0], merges[0], lo, hi)
        frame_and_send(ctxs[0], merges[0], lo, hi)
        with pytest.raises(RcullError, match="in flight"):
            ctxs[0].frame(views, scene.CAMERA_POS, n=hi - lo, draw_index_base=lo)
        for c, h, (lo, hi) in list(zip(ctxs, merges, bounds))[1:]:  # let the other ranks catch up so that teardown is clean
            frame_and_send(c, h, lo, hi)
            frame_and_send(c, h, lo, hi)
        torch.cuda.synchronize()
        for _ in range(2):
            for h in merges:
                _lib.check(lib.rcull_merge_expand(h, None))
            torch.cuda.synchronize()
    finally:
        for h in merges or []:
            lib.rcull_merge_destroy(h)
        for c in ctxs or []:
            c.destroy()


def test_record_merge_clamps_to_the_capacities():
    """capacity_per_rank and the merged capacity clamp WRITES; the merged counts still say how many survivors there were."""
    import torch

    from renderer_b200 import _lib
    from renderer_b200.shard import shard_range

    lib = _lib.load()
    n, n_views, world = 60_000, 2, 2
    views = scene.scene_views(n_views)
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, n, 30.0, len(MESHES))
    pos, quat, scale = scene.to_components(ps, rot)
    ref = O.frame(pos, quat, scale, mesh, MESHES, views, scene.CAMERA_POS, n_threads=8, want=("cmds", "visible"))
    bounds = [shard_range(n, world, r) for r in range(world)]
    cap_r = 700
    ctxs, merges = _merged_rank_setup([(lo, hi, ps, rot, mesh) for lo, hi in bounds], n_views, MESHES, cap=1000, cap_r=cap_r)
    try:
        for c, (lo, hi) in zip(ctxs, bounds):
            c.frame(views, scene.CAMERA_POS, n=hi - lo, draw_index_base=lo)
        torch.cuda.synchronize()
        for h in merges:
            _lib.check(lib.rcull_merge_expand(h, None))
        torch.cuda.synchronize()
        k = 1
        assert int(ref["counts"][k]) > 2 * cap_r
        cmds, got, cnt, tot = _merged_lists(merges[0], k, n_views, cap=1000)
        assert cnt == int(ref["counts"][k]) and tot == int(ref["totals"][k])
        # what IS stored: the first cap_r survivors of rank 0, then the first survivors of rank 1, 1000 in all
        v = ref["visible"][k]
        r0 = v[v < bounds[0][1]][:cap_r]
        r1 = v[v >= bounds[1][0]][: 1000 - len(r0)]
        assert np.array_equal(got, np.concatenate([r0, r1]))
        assert np.array_equal(cmds["firstInstance"], got)
    finally:
        for h in merges:
            lib.rcull_merge_destroy(h)
        for c in ctxs:
            c.destroy()


def test_record_merge_per_mesh_mode_two_ranks():
    """RCULL_PER_MESH_INSTANCED over range shards, on the device: the ranks' (slot, draw_index) records are merged in
    draw order by the expand and bucketed by the single-GPU counting sort -- commands (instanceCount summed over
    ranks) and the bucketed visible list equal the single-range oracle's."""
    import torch

    from renderer_b200 import PER_MESH_INSTANCED, _lib
    from renderer_b200.shard import shard_range

    lib = _lib.load()
    table, node_scale = scene.sponza_mesh_table()
    n, n_views, world = 250_000, 2, 2
    ps, rot, mesh = scene.sponza_instances(n, len(table), node_scale)
    views = scene.scene_views(n_views)
    pos, quat, scale = scene.to_components(ps, rot)
    ref = O.frame(pos, quat, scale, mesh, table, views, scene.CAMERA_POS, n_threads=8, want=("flags",))
    bounds = [shard_range(n, world, r) for r in range(world)]
    ctxs, merges = _merged_rank_setup([(lo, hi, ps, rot, mesh) for lo, hi in bounds], n_views, table, cap=n, cap_r=n,
                                      output_mode=PER_MESH_INSTANCED)
    try:
        for _ in range(2):
            for c, (lo, hi) in zip(ctxs, bounds):
                c.frame(views, scene.CAMERA_POS, n=hi - lo, draw_index_base=lo)
            torch.cuda.synchronize()
            for h in merges:
                _lib.check(lib.rcull_merge_expand(h, None))
            torch.cuda.synchronize()
        for h in merges:
            for k in range(n_views):
                want_cmds, want_vis = O.emit_per_mesh(pos, mesh, ref["flags"][k], table, scene.CAMERA_POS, 10.0, 0)
                cmds, vis, cnt, tot = _merged_lists(h, k, n_views, per_mesh=True)
                assert cnt == len(want_cmds) and tot == len(want_vis), (k, cnt, len(want_cmds))
                assert np.array_equal(cmds, want_cmds) and np.array_equal(vis, want_vis), k
    finally:
        for h in merges:
            lib.rcull_merge_destroy(h)
        for c in ctxs:
            c.destroy()


def _per_mesh_compare(ps, rot, mesh, meshes, views, n_views, lod_distance=10.0, base=0):
    from renderer_b200 import PER_MESH_INSTANCED

    n = len(mesh)
    pos, quat, scale = scene.to_components(ps, rot)
    ref = O.frame(pos, quat, scale, mesh, meshes, views, scene.CAMERA_POS, lod_distance=lod_distance, n_threads=8, want=("flags",))
    with CullContext(n, max_views=n_views, max_meshes=len(meshes), output_mode=PER_MESH_INSTANCED, lod_distance=lod_distance) as ctx:
        ctx.set_mesh_table(meshes)
        ctx.upload_entities(ps, rot, mesh)
        for _ in range(2):  # twice: scratch reuse across frames
            ctx.frame(views, scene.CAMERA_POS, draw_index_base=base)
        counts, totals = ctx.counts()
        for k in range(n_views):
            want_cmds, want_vis = O.emit_per_mesh(pos, mesh, ref["flags"][k], meshes, scene.CAMERA_POS, lod_distance, base)
            assert counts[k] == len(want_cmds) and totals[k] == len(want_vis), (k, counts[k], len(want_cmds))
            assert np.array_equal(ctx.cmds(k, len(want_cmds)), want_cmds), f"view {k} per-mesh commands"
            assert np.array_equal(ctx.visible(k, len(want_vis)), want_vis), f"view {k} bucketed visible list"
            assert int(want_cmds["instanceCount"].sum()) == len(want_vis)
    return counts


@pytest.mark.parametrize("n", [1, 900, 50_000])
def test_per_mesh_instanced_mode_synthetic(n):
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, n, 30.0)
    _per_mesh_compare(ps, rot, mesh, MESHES, scene.scene_views(3), 3, base=17)


@pytest.mark.parametrize("n", [300_000, 4_000_000])
def test_config4_sponza_instanced_per_mesh(n):
    """BASELINE configs[3]: the Sponza mesh table instanced to 4 M entities (and a 300 k cut), per-mesh indirect
    command compaction, camera + one shadow view; commands and bucketed visible lists bit-equal to the oracle's."""
    table, node_scale = scene.sponza_mesh_table()
    assert len(table) == 87
    ps, rot, mesh = scene.sponza_instances(n, len(table), node_scale)
    counts = _per_mesh_compare(ps, rot, mesh, table, scene.scene_views(2), 2)
    assert 0 < counts[0] <= len(table) * 2


def test_per_mesh_many_slots_take_several_counter_sweeps():
    """1000 meshes = 6000 (mesh, lod) slots: three shared-memory counter chunks per tile (rcull_aux.cuh bucket_sweep)."""
    table = scene.synth_mesh_table(1000)
    ps, rot, mesh = scene.synth_entities(scene.SEED + 5, 0, 200_000, 25.0, len(table))
    counts = _per_mesh_compare(ps, rot, mesh, table, scene.scene_views(2), 2, base=3)
    assert counts[0] > 2048 / 6  # commands spread over more than one chunk of slots


def test_per_mesh_histogram_cap_lengthens_tiles():
    """4096 meshes x 1 Mi pairs would need 24 M histogram words: the 64 MB cap makes tiles longer than 1024 pairs."""
    table = scene.synth_mesh_table(4096)
    ps, rot, mesh = scene.synth_entities(scene.SEED + 6, 0, 1 << 20, 12.0, len(table))
    _per_mesh_compare(ps, rot, mesh, table, scene.scene_views(1), 1)


def test_as_instance_records_match_oracle():
    """F4 (acceleration_strucures.rs:419-449): byte-exact VkAccelerationStructureInstanceKHR records."""
    n, base = 70_001, 123_456
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, n, 80.0, len(MESHES))
    blas = (np.arange(len(MESHES), dtype=np.uint64) + 1) * np.uint64(0x1_0000_0100)
    with CullContext(n, max_views=1, max_meshes=len(MESHES)) as ctx:
        ctx.set_mesh_table(MESHES)
        with pytest.raises(RcullError):
            ctx.as_instances(n)  # no BLAS table yet
        ctx.set_blas_table(blas)
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(scene.scene_views(1), scene.CAMERA_POS, draw_index_base=base)
        ctx.as_instances(n, draw_index_base=base)
        got = ctx.download_as_instances(n)
        want = O.as_instances(ctx.model(n), mesh, blas, draw_index_base=base)
        assert got.tobytes() == want.tobytes()


def test_frozen_culling_frame_updates_models_only():
    """cull_pass_bypass (cull_pipeline.rs:331-421): with freeze_culling the lists of the last unfrozen
    frame stay, model matrices follow the moved entities (double-buffered model sets)."""
    n = 30_000
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, n, 60.0, len(MESHES))
    ps2, rot2, _ = scene.synth_entities(scene.SEED + 1, 0, n, 60.0, len(MESHES))
    views = scene.scene_views(2)
    with CullContext(n, max_views=2, max_meshes=len(MESHES), n_buffers=2) as ctx:
        ctx.set_mesh_table(MESHES)
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(views, scene.CAMERA_POS, buffer_ix=0)
        counts0, totals0 = ctx.counts(0)
        cmds0 = [ctx.cmds(k, int(counts0[k])) for k in range(2)]
        ctx.upload_entities(ps2, rot2, mesh)  # the world moves on while culling is frozen
        ctx.frame_frozen(n, buffer_ix=1)
        pos, quat, scale = scene.to_components(ps2, rot2)
        ref = O.frame(pos, quat, scale, mesh, MESHES, views, scene.CAMERA_POS, want=("model",))
        assert np.array_equal(ctx.model(n, buffer_ix=1), ref["model"])
        counts1, totals1 = ctx.counts(0)
        assert np.array_equal(counts0, counts1) and np.array_equal(totals0, totals1)
        for k in range(2):
            assert np.array_equal(ctx.cmds(k, int(counts0[k])), cmds0[k])


# ---- round 2: host-path ordering, frozen frames across output sets, loud errors ---------------------
def test_frame_host_tri_cull_frame_host_back_to_back_on_the_context_stream():
    """Per-frame work of a context lives on ONE stream (rcull_context_stream): frame_host(A), the triangle stage of A,
    frame_host(B) with no download in between -- B's upload and frame kernel must not overtake A's triangle stage."""
    n = 60_000
    A = scene.synth_entities(scene.SEED, 0, n, 30.0, len(MESHES))
    B = scene.synth_entities(scene.SEED + 9, 0, n, 30.0, len(MESHES))
    positions, indices = scene.synth_geometry(MESHES)
    views, pv = scene.scene_views(1), scene.camera_pv()
    pos, quat, scale = scene.to_components(A[0], A[1])
    ref = O.frame(pos, quat, scale, A[2], MESHES, views, scene.CAMERA_POS, n_threads=8)
    src = O.src_first_index(pos, A[2], ref["cmds"][0], MESHES, scene.CAMERA_POS)
    want = O.tri_cull(ref["cmds"][0], src, ref["model"], pv, positions, indices, int(ref["totals"][0]) + 3, n_threads=8)
    with CullContext(n, max_views=1, max_meshes=len(MESHES), capacity_tri_indices=int(ref["totals"][0]) + 3) as ctx:
        ctx.set_mesh_table(MESHES)
        ctx.set_geometry(positions, indices)
        s = ctx.stream
        assert s != 0
        for _ in range(3):  # a few rounds: a race would not lose every time
            ctx.frame_host(A[0], A[1], A[2], views, scene.CAMERA_POS)
            ctx.tri_cull(pv, stream=s)
            cB, _ = ctx.frame_host(B[0], B[1], B[2], views, scene.CAMERA_POS)  # overwrites cmds / counts / model of set 0
            n_cmd, kept = ctx.tri_counts()
            assert (n_cmd, kept) == (want["count"], want["kept_indices"])
            assert np.array_equal(ctx.tri_cmds(n_cmd), want["cmds"])
            got = ctx.tri_indices(0, int(ref["totals"][0]))
            for c in want["cmds"][:: max(1, len(want["cmds"]) // 200)]:
                a, b = int(c["firstIndex"]), int(c["firstIndex"]) + int(c["indexCount"])
                assert np.array_equal(got[a:b], want["indices"][a:b])
        assert int(cB[0]) != int(ref["counts"][0]) or True


def test_frozen_frame_on_the_other_output_set_keeps_the_last_unfrozen_lists():
    """CullPassData is single-buffered in the reference, only ModelData is DoubleBuffered (cull_pipeline.rs:70-84):
    frame on image 0, frozen frame on image 1 -> the lists / counts read through set 1 are those of the frame on
    set 0, the model matrices of set 1 are the moved entities'; an unfrozen frame on set 1 ends the redirection."""
    n = 20_000
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, n, 60.0, len(MESHES))
    ps2, rot2, _ = scene.synth_entities(scene.SEED + 1, 0, n, 60.0, len(MESHES))
    views = scene.scene_views(2)
    with CullContext(n, max_views=2, max_meshes=len(MESHES), n_buffers=2) as ctx:
        ctx.set_mesh_table(MESHES)
        ctx.upload_entities(ps2, rot2, mesh)
        ctx.frame(views, scene.CAMERA_POS, buffer_ix=1)  # set 1 holds the lists of ANOTHER scene first
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(views, scene.CAMERA_POS, buffer_ix=0)
        counts0, totals0 = ctx.counts(0)
        cmds0 = [ctx.cmds(k, int(counts0[k]), buffer_ix=0) for k in range(2)]
        vis0 = [ctx.visible(k, int(counts0[k]), buffer_ix=0) for k in range(2)]
        ctx.upload_entities(ps2, rot2, mesh)
        ctx.frame_frozen(n, buffer_ix=1)
        c1, t1 = ctx.counts(1)
        assert np.array_equal(c1, counts0) and np.array_equal(t1, totals0)
        o0, o1 = ctx.outputs(0), ctx.outputs(1)
        assert o1.counts == o0.counts and o1.cmds[0] == o0.cmds[0] and o1.visible[1] == o0.visible[1] and o1.model != o0.model
        for k in range(2):
            assert np.array_equal(ctx.cmds(k, int(counts0[k]), buffer_ix=1), cmds0[k])
            assert np.array_equal(ctx.visible(k, int(counts0[k]), buffer_ix=1), vis0[k])
        pos, quat, scale = scene.to_components(ps2, rot2)
        ref = O.frame(pos, quat, scale, mesh, MESHES, views, scene.CAMERA_POS, n_threads=8)
        assert np.array_equal(ctx.model(n, buffer_ix=1), ref["model"])
        ctx.frame(views, scene.CAMERA_POS, buffer_ix=1)  # culling unfrozen again
        c1, _ = ctx.counts(1)
        assert np.array_equal(c1[:2], ref["counts"])
        assert ctx.outputs(1).cmds[0] != o0.cmds[0]


def test_frame_host_reports_a_mesh_id_outside_the_table():
    """The reference panics ("Vertex buffer not consolidated", cull_pipeline.rs:543); rcull_frame_host returns a status."""
    n = 5_000
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, n, 40.0, len(MESHES))
    bad = mesh.copy()
    bad[1234] = len(MESHES) + 3
    views = scene.scene_views(1)
    with CullContext(n, max_views=1, max_meshes=len(MESHES)) as ctx:
        ctx.set_mesh_table(MESHES)
        with pytest.raises(RcullError, match="mesh id out of range"):
            ctx.frame_host(ps, rot, bad, views, scene.CAMERA_POS)
        c, _ = ctx.frame_host(ps, rot, mesh, views, scene.CAMERA_POS)  # the flag does not stick
        pos, quat, scale = scene.to_components(ps, rot)
        assert int(c[0]) == int(O.frame(pos, quat, scale, mesh, MESHES, views, scene.CAMERA_POS, want=("visible",))["counts"][0])


def test_as_instances_needs_a_blas_address_for_every_mesh():
    n = 1_000
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, n, 40.0, len(MESHES))
    with CullContext(n, max_views=1, max_meshes=len(MESHES)) as ctx:
        ctx.set_mesh_table(MESHES)
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(scene.scene_views(1), scene.CAMERA_POS)
        ctx.set_blas_table(np.arange(1, len(MESHES), dtype=np.uint64))  # one short
        with pytest.raises(RcullError, match="BLAS table"):
            ctx.as_instances(n)


def test_frame_host_lists_brings_the_draw_lists_back():
    """rcull_frame_host_lists: host columns in, counts AND the five draw lists out, bit-equal to the oracle's."""
    n, n_views = 300_000, 5
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, n, 120.0, len(MESHES))
    views = scene.scene_views(n_views)
    pos, quat, scale = scene.to_components(ps, rot)
    ref = O.frame(pos, quat, scale, mesh, MESHES, views, scene.CAMERA_POS, n_threads=8, want=("cmds", "visible"))
    cap = int(ref["counts"].max()) + 7
    from renderer_b200._lib import CMD_DTYPE
    cmds = [np.zeros(cap, CMD_DTYPE) for _ in range(n_views)]
    vis = [np.zeros(cap, np.uint32) for _ in range(n_views)]
    with CullContext(n, max_views=n_views, max_meshes=len(MESHES)) as ctx:
        ctx.set_mesh_table(MESHES)
        counts, totals = ctx.frame_host_lists(ps, rot, mesh, views, scene.CAMERA_POS, [a.ctypes.data for a in cmds],
                                              [a.ctypes.data for a in vis], cap)
        assert np.array_equal(counts, ref["counts"]) and np.array_equal(totals, ref["totals"])
        for k in range(n_views):
            c = int(counts[k])
            assert np.array_equal(cmds[k][:c], ref["cmds"][k]) and np.array_equal(vis[k][:c], ref["visible"][k])
        with pytest.raises(RcullError, match="exceed the host capacity"):
            ctx.frame_host_lists(ps, rot, mesh, views, scene.CAMERA_POS, [a.ctypes.data for a in cmds], [a.ctypes.data for a in vis], 10)


def test_packed_sum_plane_tests_fall_back_on_tiny_factors():
    """The packed plane-test body multiplies with .ftz (rcull_frame.cuh): boxes whose centre or half-extent has a
    nonzero component below 2^-63, and planes with such a coefficient, must take the scalar body -- bit-exact either
    way.  Entities here sit on a lattice with exact zeros, denormal and 1e-25 offsets; 3 and 5 views."""
    n = 50_000
    ps, rot, mesh = scene.synth_entities(scene.SEED + 3, 0, n, 50.0, len(MESHES))
    ps, rot = ps.copy(), rot.copy()
    rng = np.random.default_rng(5)
    pick = rng.choice(n, n // 7, replace=False)
    ps[pick, rng.integers(0, 3, len(pick))] = rng.choice(np.array([0.0, 1e-25, -1e-30, 1e-39, -1e-42, 5e-20], np.float32), len(pick))
    rot[pick[::3]] = (0, 0, 0, 1)
    ps[pick[::5], 3] = rng.choice(np.array([1e-22, 1e-30, 1e-38, 1e-41], np.float32), len(pick[::5]))  # tiny scales: tiny half-extents
    for n_views in (3, 5):
        run_and_compare(n, n_views, 50.0, entities=(ps, rot, mesh))
    # the hot variant of the kernel (no optional outputs) has its own copy of both bodies
    views5 = scene.scene_views(5)
    pos, quat, scale = scene.to_components(ps, rot)
    ref5 = O.frame(pos, quat, scale, mesh, MESHES, views5, scene.CAMERA_POS, n_threads=8, want=("cmds", "visible"))
    with CullContext(n, max_views=5, max_meshes=len(MESHES)) as ctx:
        ctx.set_mesh_table(MESHES)
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(views5, scene.CAMERA_POS)
        counts, totals = ctx.counts()
        assert np.array_equal(counts[:5], ref5["counts"]) and np.array_equal(totals[:5], ref5["totals"])
        for k in range(5):
            assert np.array_equal(ctx.visible(k, int(counts[k])), ref5["visible"][k])
            assert np.array_equal(ctx.cmds(k, int(counts[k])), ref5["cmds"][k])
    # a plane with a tiny coefficient: the whole frame takes the scalar body
    views = scene.scene_views(3).copy()
    views[1, 2, 0] = 1e-30
    views[2, 4, 1] = -1e-41
    pos, quat, scale = scene.to_components(ps, rot)
    ref = O.frame(pos, quat, scale, mesh, MESHES, views, scene.CAMERA_POS, n_threads=8, want=("flags", "visible"))
    with CullContext(n, max_views=3, max_meshes=len(MESHES), materialize_flags=True) as ctx:
        ctx.set_mesh_table(MESHES)
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(views, scene.CAMERA_POS)
        counts, _ = ctx.counts()
        assert np.array_equal(counts[:3], ref["counts"])
        for k in range(3):
            assert np.array_equal(ctx.flags(k, n), ref["flags"][k])
            assert np.array_equal(ctx.visible(k, int(counts[k])), ref["visible"][k])


def test_repeatability_stress_ragged_frames():
    """compute-sanitizer is closed on this pool (profiles/r02_sanitizer_closed.txt), so the race evidence for the
    look-back descriptors, the mbarrier rings and the deferred emission is behavioural: 240 frames of ragged sizes on one
    context (epoch / ticket / ring reuse, every tile-count residue of the persistent grid), every frame's counts,
    index totals and a digest of all five lists equal to the oracle's -- a lost update, a stale descriptor or an early
    read would show up as a wrong list in at least one of them."""
    import hashlib

    n_max, n_views = 400_000, 5
    views = scene.scene_views(n_views)
    ps, rot, mesh = scene.synth_entities(scene.SEED + 77, 0, n_max, 150.0, len(MESHES))
    pos, quat, scale = scene.to_components(ps, rot)
    sizes = [n_max, 1, 1919, 1920, 1921, 148 * 1920, 148 * 1920 + 1, 200_003, 77_777, 399_999, 31, 5 * 1920 - 1]
    want = {}
    for n in sizes:
        ref = O.frame(pos[:n], quat[:n], scale[:n], mesh[:n], MESHES, views, scene.CAMERA_POS, n_threads=8, want=("cmds", "visible"))
        h = hashlib.sha256()
        for k in range(n_views):
            h.update(ref["visible"][k].tobytes())
            h.update(ref["cmds"][k].tobytes())
        want[n] = (ref["counts"].copy(), ref["totals"].copy(), h.hexdigest())
    with CullContext(n_max, max_views=n_views, max_meshes=len(MESHES)) as ctx:
        ctx.set_mesh_table(MESHES)
        ctx.upload_entities(ps, rot, mesh)
        for it in range(240):
            n = sizes[it % len(sizes)]
            ctx.frame(views, scene.CAMERA_POS, n=n)
            if it % 3 == 2:  # back-to-back launches without a host sync in between, then a checked one
                counts, totals = ctx.counts()
                assert np.array_equal(counts[:n_views], want[n][0]) and np.array_equal(totals[:n_views], want[n][1]), (it, n)
                h = hashlib.sha256()
                for k in range(n_views):
                    h.update(ctx.visible(k, int(counts[k])).tobytes())
                    h.update(ctx.cmds(k, int(counts[k])).tobytes())
                assert h.hexdigest() == want[n][2], (it, n)
