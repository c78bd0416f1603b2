"""rcull_frame_noalloc (include/rcull.h): the frame kernel for callers that cannot allocate.  Its scratch is a
file-scope device array inside librcull.so instead of buffers an rcull_ctx owns; the caller passes every input
and output as device memory.  This entry point has no counterpart in the reference,
so its oracle is the existing entry point: rcull_frame on the same inputs, compared bit for
bit -- model matrices with == on floats, like tests/test_gpu_parity.py.

Collected last on purpose: a failure here hides no other test.  The device buffers of
the test come from torch; the entry point itself allocates nothing, which the test checks by watching the free
device memory across the repeated calls."""
import ctypes as C

import numpy as np
import pytest

from renderer_b200 import CullContext, _lib, scene

pytestmark = pytest.mark.gpu

MESHES = scene.synth_mesh_table()
TILE = _lib.NOALLOC_TILE


def _padded(n):
    return (n + TILE - 1) // TILE * TILE


def _existing_entry_point(n, n_views, half_box, base, capacity_visible=0):
    """rcull_create + rcull_frame on the seeded scene: what rcull_frame_noalloc has to reproduce."""
    ps, rot, mesh = scene.synth_entities(scene.SEED + 7, 0, n, half_box, len(MESHES))
    views = scene.scene_views(n_views)
    with CullContext(max(n, 1), max_views=n_views, max_meshes=len(MESHES), capacity_visible=capacity_visible) as ctx:
        ctx.set_mesh_table(MESHES)
        ctx.upload_entities(ps, rot, mesh)
        ctx.frame(views, scene.CAMERA_POS, n=n, draw_index_base=base)
        counts, totals = ctx.counts()
        cap = capacity_visible or n
        out = {
            "counts": counts[:n_views].copy(),
            "totals": totals[:n_views].copy(),
            "model": ctx.model(n),
            "visible": [ctx.visible(k, min(int(counts[k]), cap)) for k in range(n_views)],
            "cmds": [ctx.cmds(k, min(int(counts[k]), cap)) for k in range(n_views)],
        }
    return (ps, rot, mesh, views), out


class _Buffers:
    """Device memory the CALLER owns, sized for the largest case and filled with garbage between calls."""

    def __init__(self, torch, n_max, cap_vis):
        dev = "cuda:0"
        self.torch = torch
        self.cap_vis = cap_vis
        self.ps = torch.zeros((_padded(n_max), 4), dtype=torch.float32, device=dev)
        self.rot = torch.zeros((_padded(n_max), 4), dtype=torch.float32, device=dev)
        self.mesh = torch.zeros(_padded(n_max), dtype=torch.int32, device=dev)
        self.model = torch.zeros((n_max + 1, 16), dtype=torch.float32, device=dev)
        self.cmds = [torch.zeros((cap_vis, 5), dtype=torch.int32, device=dev) for _ in range(_lib.MAX_VIEWS)]
        self.visible = [torch.zeros(cap_vis, dtype=torch.int32, device=dev) for _ in range(_lib.MAX_VIEWS)]
        self.counts = torch.zeros(_lib.MAX_VIEWS, dtype=torch.int32, device=dev)
        self.totals = torch.zeros(_lib.MAX_VIEWS, dtype=torch.int64, device=dev)
        self.error = torch.zeros(1, dtype=torch.int32, device=dev)
        # pinned staging, so that loading a case allocates nothing either
        self.h_ps = torch.zeros((_padded(n_max), 4), dtype=torch.float32).pin_memory()
        self.h_rot = torch.zeros((_padded(n_max), 4), dtype=torch.float32).pin_memory()
        self.h_mesh = torch.zeros(_padded(n_max), dtype=torch.int32).pin_memory()

    def load(self, ps, rot, mesh):
        n = len(mesh)
        # slots past n hold junk on purpose: the kernel must ignore them (mesh ids far outside the table included)
        self.h_ps.fill_(float("nan"))
        self.h_rot.fill_(3.0)
        self.h_mesh.fill_(0x7FFFFFFF)
        self.h_ps[:n] = self.torch.from_numpy(ps)
        self.h_rot[:n] = self.torch.from_numpy(rot)
        self.h_mesh[:n] = self.torch.from_numpy(mesh.view(np.int32))
        self.ps.copy_(self.h_ps)
        self.rot.copy_(self.h_rot)
        self.mesh.copy_(self.h_mesh)
        # outputs start as garbage: whatever must start at zero is the entry point's job
        self.model.fill_(-7.0)
        for t in self.cmds + self.visible:
            t.fill_(-1)
        self.counts.fill_(12345)
        self.totals.fill_(-1)
        self.error.fill_(99)
        self.torch.cuda.synchronize()

    def io(self):
        io = _lib.NoallocIO()
        io.pos_scale, io.rot, io.mesh, io.model = self.ps.data_ptr(), self.rot.data_ptr(), self.mesh.data_ptr(), self.model.data_ptr()
        for k in range(_lib.MAX_VIEWS):
            io.cmds[k] = self.cmds[k].data_ptr()
            io.visible[k] = self.visible[k].data_ptr()
        io.counts, io.index_totals, io.error = self.counts.data_ptr(), self.totals.data_ptr(), self.error.data_ptr()
        io.capacity_visible = self.cap_vis
        return io


def _call(lib, io, n, views, n_views, base, meshes=MESHES, lod_distance=10.0, stream=0):
    v = np.ascontiguousarray(views, dtype=np.float32)
    cam = np.ascontiguousarray(scene.CAMERA_POS, dtype=np.float32)
    m = np.ascontiguousarray(meshes, dtype=_lib.MESH_DTYPE)
    return lib.rcull_frame_noalloc(0, n, C.byref(io), _lib.ptr(m), len(m), _lib.ptr(v), n_views, _lib.ptr(cam),
                                   lod_distance, base, C.c_void_p(stream))


def _check(buf, ref, n, n_views):
    t = buf.torch
    t.cuda.synchronize()
    assert int(buf.error.item()) == 0
    counts = buf.counts.cpu().numpy().view(np.uint32)
    totals = buf.totals.cpu().numpy().view(np.uint64)
    assert np.array_equal(counts[:n_views], ref["counts"])
    assert np.array_equal(totals[:n_views], ref["totals"])
    assert np.array_equal(buf.model[:n].cpu().numpy(), ref["model"])
    for k in range(n_views):
        c = min(int(counts[k]), buf.cap_vis)
        assert np.array_equal(buf.visible[k][:c].cpu().numpy().view(np.uint32), ref["visible"][k]), f"view {k} visible"
        got = buf.cmds[k][:c].contiguous().cpu().numpy().view(_lib.CMD_DTYPE).reshape(-1)
        assert np.array_equal(got, ref["cmds"][k]), f"view {k} cmds"
        # nothing is written past the list
        assert bool((buf.visible[k][c:] == -1).all())


# (n, views, half_box, draw_index_base): a large frame, then a smaller one with fewer views that finds the
# scratch full of the first one's descriptors, a single partial tile, and the large one again
CASES = [(300_007, 5, 60.0, 0), (40_001, 2, 30.0, 1000), (7, 1, 4.0, 5), (1920 * 3, 3, 20.0, 0)]


def test_noalloc_matches_rcull_frame_on_calls_in_a_row():
    import torch

    lib = _lib.load()
    refs = [_existing_entry_point(n, v, hb, base) for n, v, hb, base in CASES]
    buf = _Buffers(torch, max(c[0] for c in CASES), max(c[0] for c in CASES))
    io = buf.io()
    # first round: every kernel variant gets loaded (lazy module loading may take device memory once)
    for (n, v, hb, base), (inp, ref) in zip(CASES, refs):
        buf.load(*inp[:3])
        assert _call(lib, io, n, inp[3], v, base) == 0, lib.rcull_last_error()
        _check(buf, ref, n, v)
    # second round: the same calls in a row again, now with the free device memory watched -- no cudaMalloc
    for (n, v, hb, base), (inp, ref) in zip(CASES, refs):
        buf.load(*inp[:3])  # ends with a synchronize; no torch work between the two readings below
        free_before = torch.cuda.mem_get_info()[0]
        assert _call(lib, io, n, inp[3], v, base) == 0, lib.rcull_last_error()
        assert _call(lib, io, n, inp[3], v, base) == 0, lib.rcull_last_error()  # back to back on one stream: ordered, same result
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_before, "rcull_frame_noalloc changed the amount of free device memory"
        _check(buf, ref, n, v)


def test_noalloc_on_a_side_stream_and_with_a_clamped_list():
    """A caller's own stream, and capacity_visible below the survivor count: counted, not stored (as rcull_frame)."""
    import torch

    lib = _lib.load()
    n, v, hb, base, cap = 50_000, 2, 4.0, 0, 1000
    inp, ref = _existing_entry_point(n, v, hb, base, capacity_visible=cap)
    assert int(ref["counts"][1]) > cap
    buf = _Buffers(torch, n, cap)
    buf.load(*inp[:3])
    s = torch.cuda.Stream()
    assert _call(lib, buf.io(), n, inp[3], v, base, stream=s.cuda_stream) == 0, lib.rcull_last_error()
    s.synchronize()
    _check(buf, ref, n, v)


def test_noalloc_empty_frame_clears_the_counts():
    import torch

    lib = _lib.load()
    buf = _Buffers(torch, 10, 10)
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, 10, 4.0, len(MESHES))
    buf.load(ps, rot, mesh)
    assert _call(lib, buf.io(), 0, scene.scene_views(2), 2, 0) == 0
    torch.cuda.synchronize()
    assert int(buf.counts[:2].abs().sum().item()) == 0 and int(buf.totals[:2].abs().sum().item()) == 0 and int(buf.error.item()) == 0


def test_noalloc_reports_a_mesh_id_outside_the_table():
    import torch

    lib = _lib.load()
    n = 5000
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, n, 30.0, len(MESHES))
    mesh = mesh.copy()
    mesh[1234] = len(MESHES) + 3
    buf = _Buffers(torch, n, n)
    buf.load(ps, rot, mesh)
    assert _call(lib, buf.io(), n, scene.scene_views(1), 1, 0) == 0
    torch.cuda.synchronize()
    assert int(buf.error.item()) != 0


def test_noalloc_rejects_what_the_static_scratch_cannot_hold():
    import torch

    lib = _lib.load()
    buf = _Buffers(torch, 10, 10)
    io = buf.io()
    views = scene.scene_views(1)
    assert _call(lib, io, _lib.NOALLOC_MAX_ENTITIES + 1, views, 1, 0) == -1
    assert b"RCULL_NOALLOC_MAX_ENTITIES" in lib.rcull_last_error()
    big = np.zeros(_lib.NOALLOC_MAX_MESHES + 1, _lib.MESH_DTYPE)
    big["n_lods"] = 1
    assert _call(lib, io, 10, views, 1, 0, meshes=big) == -1
    assert b"RCULL_NOALLOC_MAX_MESHES" in lib.rcull_last_error()
    assert _call(lib, io, 10, scene.scene_views(8), 9, 0) == -1
    io.pos_scale += 4  # misaligned column
    assert _call(lib, io, 10, views, 1, 0) == -1
    assert b"aligned" in lib.rcull_last_error()
