"""The C++ host side above the C ABI (include/rcull_systems.hpp), driven through
tests/cpp/systems_driver.cpp: the reference's component / resource / system
interface (Position, Rotation, Scale, GltfMesh, Camera, Light, ImageIndex;
project_camera, gpu_transform_cull, freeze_culling) restated in C++ because
there is no rustc on the box.  The driver only runs the systems; the check
against the oracle happens here."""
import os
import shutil
import signal
import struct
import subprocess

import numpy as np
import pytest

import oracle_lib as O
from renderer_b200 import scene
from renderer_b200._lib import CMD_DTYPE, MESH_DTYPE

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "systems_driver.cpp")
EXE = os.path.join(ROOT, "tests", "cpp", "systems_driver")
DEPS = [SRC, os.path.join(ROOT, "include", "rcull_systems.hpp"), os.path.join(ROOT, "include", "rcull.h")]
MESHES = scene.synth_mesh_table()
ASPECT = 16.0 / 9.0


def compile_driver():
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-I" + os.path.join(ROOT, "include"), "-o", EXE + ".new", SRC,
                           "-L" + os.path.join(ROOT, "renderer_b200"), "-lrcull", "-Wl,-rpath,$ORIGIN/../../renderer_b200"])
    os.replace(EXE + ".new", EXE)  # a failed build leaves the previous binary in place


@pytest.fixture(scope="module")
def driver():
    """g++ only: the header and the driver hold no device code.  __graft_entry__.build() compiles it too and
    the binary travels with the tree; it is rebuilt here when missing or older than its sources."""
    stale = os.path.exists(EXE) and any(os.path.getmtime(d) > os.path.getmtime(EXE) for d in DEPS)
    if not os.path.exists(EXE):
        compile_driver()
    elif stale and shutil.which("g++"):
        try:
            compile_driver()
        except subprocess.CalledProcessError:
            pass  # a copied tree can scramble mtimes: the shipped binary stays in use
    return EXE


def light_components(n_lights):
    pos = np.stack([scene.LIGHT_POS[k % 4] * np.float32(1.0 + 0.25 * (k // 4)) for k in range(n_lights)]) \
        if n_lights else np.zeros((0, 3), np.float32)
    rot = np.stack([scene.quat_look_at_lh(-p.astype(np.float64)) for p in pos]) if n_lights else np.zeros((0, 4), np.float32)
    return pos.astype(np.float32), rot.astype(np.float32)


def write_scene(path, pos, quat, scale, mesh, meshes, n_lights, base=0, lod_distance=10.0):
    lp, lr = light_components(n_lights)
    with open(path, "wb") as f:
        f.write(struct.pack("<QIIIff", len(scale), len(meshes), n_lights, base, ASPECT, lod_distance))
        f.write(scene.CAMERA_POS.astype("<f4").tobytes())
        f.write(scene.CAMERA_ROT.astype("<f4").tobytes())
        f.write(np.concatenate([lp, lr], axis=1).astype("<f4").tobytes())  # Light = (Position, Rotation)
        f.write(np.ascontiguousarray(meshes, MESH_DTYPE).tobytes())
        for a, t in ((pos, "<f4"), (quat, "<f4"), (scale, "<f4"), (mesh, "<u4")):
            f.write(np.ascontiguousarray(a).astype(t).tobytes())


class Result:
    def __init__(self, path, n, full):
        self.buf = open(path, "rb").read()
        self.off = 0
        self.n_views = int(self.take("<u4", 1)[0])
        self.counts = self.take("<u4", 8)
        self.totals = self.take("<u8", 8)
        self.planes = self.take("<f4", self.n_views * 24).reshape(self.n_views, 6, 4)
        if not full:
            assert self.off == len(self.buf)
            return
        self.model = self.take("<f4", n * 16).reshape(n, 16)
        self.aabb = self.take("<f4", n * 6).reshape(n, 6)
        self.flags, self.cmds, self.visible = [], [], []
        for k in range(self.n_views):
            c = int(self.counts[k])
            self.flags.append(self.take("u1", n))
            self.cmds.append(self.take(CMD_DTYPE, c))
            self.visible.append(self.take("<u4", c))
        self.frozen_model = self.take("<f4", n * 16).reshape(n, 16)
        self.counts_after_frozen = self.take("<u4", 8)
        assert self.off == len(self.buf), "result file has trailing bytes"

    def take(self, dtype, count):
        a = np.frombuffer(self.buf, dtype=dtype, count=count, offset=self.off)
        self.off += a.nbytes
        return a


@pytest.mark.parametrize("n_lights", [0, 4, 7])
def test_project_camera_and_light_frusta_match_the_harness_views(driver, tmp_path, n_lights):
    """Host-only systems (row A5): the C++ Camera / Light path gives the planes the parity tests feed
    the oracle (scene.scene_views), bit for bit."""
    pos, quat, scale, mesh = O.synth_entities(scene.SEED, 0, 3, 10.0, len(MESHES))
    write_scene(tmp_path / "s.bin", pos, quat, scale, mesh, MESHES, n_lights)
    subprocess.check_call([driver, "planes", str(tmp_path / "s.bin"), str(tmp_path / "r.bin")])
    r = Result(tmp_path / "r.bin", 3, full=False)
    assert r.n_views == 1 + n_lights
    assert np.array_equal(r.planes.view(np.uint32), scene.scene_views(1 + n_lights).view(np.uint32))
    # and the oracle's own restatement of src/ecs.rs:66-91 / shadow_mapping.rs:324-329
    assert np.array_equal(r.planes[0], O.project_camera(ASPECT, scene.CAMERA_POS, scene.CAMERA_ROT)[3])
    lp, lr = light_components(n_lights)
    for k in range(n_lights):
        assert np.array_equal(r.planes[1 + k], O.light_view(lp[k], lr[k])[3])


def test_no_device_panics_like_the_reference(driver, tmp_path):
    """panic = "abort" (Cargo.toml:133): without a CUDA device RcullRes::new aborts with the library's
    message -- no CPU fallback behind the systems."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    pos, quat, scale, mesh = O.synth_entities(scene.SEED, 0, 100, 10.0, len(MESHES))
    write_scene(tmp_path / "s.bin", pos, quat, scale, mesh, MESHES, 1)
    p = subprocess.run([driver, "frame", str(tmp_path / "s.bin"), str(tmp_path / "r.bin")], capture_output=True, text=True)
    assert p.returncode == -signal.SIGABRT, (p.returncode, p.stderr)
    assert "rcull panic: rcull_create" in p.stderr


def test_truncated_scene_panics(driver, tmp_path):
    (tmp_path / "s.bin").write_bytes(b"\x01\x00\x00")
    p = subprocess.run([driver, "planes", str(tmp_path / "s.bin"), str(tmp_path / "r.bin")], capture_output=True, text=True)
    assert p.returncode == -signal.SIGABRT and "truncated" in p.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("n,n_lights,half_box,base", [(10_000, 0, 60.0, 0), (50_001, 4, 60.0, 0), (1, 1, 2.0, 7), (0, 2, 10.0, 0),
                                                      (300_000, 2, 120.0, 1_000_000)])
def test_gpu_transform_cull_system_matches_oracle(driver, tmp_path, n, n_lights, half_box, base):
    """AoS components in, the five systems' outputs back: bit-exact against the oracle, through the
    C++ layer only (pack -> rcull_frame_host -> downloads); then a frozen frame."""
    pos, quat, scale, mesh = O.synth_entities(scene.SEED, 0, n, half_box, len(MESHES))
    write_scene(tmp_path / "s.bin", pos, quat, scale, mesh, MESHES, n_lights, base=base)
    subprocess.check_call([driver, "frame", str(tmp_path / "s.bin"), str(tmp_path / "r.bin")], timeout=300)
    r = Result(tmp_path / "r.bin", n, full=True)
    v = 1 + n_lights
    views = scene.scene_views(v)
    assert r.n_views == v and np.array_equal(r.planes.view(np.uint32), views.view(np.uint32))
    ref = O.frame(pos, quat, scale, mesh, MESHES, views, scene.CAMERA_POS, draw_index_base=base, n_threads=8)
    assert np.array_equal(r.counts[:v], ref["counts"])
    assert np.array_equal(r.totals[:v], ref["totals"])
    assert np.array_equal(r.model, ref["model"])
    assert np.array_equal(r.aabb, ref["aabb"])
    for k in range(v):
        assert np.array_equal(r.flags[k], ref["flags"][k]), f"view {k} flags"
        assert np.array_equal(r.cmds[k], ref["cmds"][k]), f"view {k} cmds"
        assert np.array_equal(r.visible[k], ref["visible"][k]), f"view {k} visible"
    frozen = O.frame(pos, quat, scale * np.float32(2.0), mesh, MESHES, views, scene.CAMERA_POS, want=("model",))
    assert np.array_equal(r.frozen_model, frozen["model"])
    assert np.array_equal(r.counts_after_frozen[:v], ref["counts"])


def write_stage2_tail(path, positions, indices, capacity, blas):
    positions = np.ascontiguousarray(positions, "<f4").reshape(-1)
    with open(path, "ab") as f:
        f.write(struct.pack("<QQQ", len(positions) // 3, len(indices), capacity))
        f.write(positions.tobytes())
        f.write(np.ascontiguousarray(indices, "<u4").tobytes())
        f.write(np.ascontiguousarray(blas, "<u8").tobytes())


@pytest.mark.gpu
@pytest.mark.parametrize("n,half_box,base", [(5_000, 40.0, 0), (1, 3.0, 0), (20_000, 30.0, 250_000)])
def test_gpu_cull_triangles_and_as_instances_systems_match_oracle(driver, tmp_path, n, half_box, base):
    """The second half of cull_pass (generate_work dispatches + compact_draw_stream) and the TLAS instance loop
    through the C++ systems: commands, count, culled index ranges and the 64-byte records are bit-exact."""
    positions, indices = scene.synth_geometry(MESHES)
    blas = (np.arange(len(MESHES), dtype=np.uint64) + 1) * np.uint64(0x1_0000_0100)
    pos, quat, scale, mesh = O.synth_entities(scene.SEED, 0, n, half_box, len(MESHES))
    views = scene.scene_views(1)
    ref = O.frame(pos, quat, scale, mesh, MESHES, views, scene.CAMERA_POS, draw_index_base=base, n_threads=8)
    total = int(ref["totals"][0])
    cap = max(total, 3)
    write_scene(tmp_path / "s.bin", pos, quat, scale, mesh, MESHES, 0, base=base)
    write_stage2_tail(tmp_path / "s.bin", positions, indices, cap, blas)
    subprocess.check_call([driver, "stage2", str(tmp_path / "s.bin"), str(tmp_path / "r.bin")], timeout=300)

    buf, off = open(tmp_path / "r.bin", "rb").read(), 0

    def take(dtype, count):
        nonlocal off
        a = np.frombuffer(buf, dtype=dtype, count=count, offset=off)
        off += a.nbytes
        return a

    assert int(take("<u4", 1)[0]) == 1
    counts, totals, pv = take("<u4", 8), take("<u8", 8), take("<f4", 16)
    assert counts[0] == ref["counts"][0] and totals[0] == total
    assert np.array_equal(pv.view(np.uint32), scene.camera_pv().view(np.uint32))
    tri_count, kept = int(take("<u4", 1)[0]), int(take("<u8", 1)[0])
    tri_cmds, culled = take(CMD_DTYPE, tri_count), take("<u4", total)
    instances = take(O.AS_INSTANCE_DTYPE, n)
    assert off == len(buf), "result file has trailing bytes"

    src = O.src_first_index(pos, mesh, ref["cmds"][0], MESHES, scene.CAMERA_POS, draw_index_base=base)
    want = O.tri_cull(ref["cmds"][0], src, ref["model"], pv, positions, indices, cap, draw_index_base=base, n_threads=8)
    assert (tri_count, kept) == (want["count"], want["kept_indices"])
    assert np.array_equal(tri_cmds, want["cmds"])
    for c in want["cmds"]:  # only the ranges the commands own are defined
        a, b = int(c["firstIndex"]), int(c["firstIndex"]) + int(c["indexCount"])
        assert np.array_equal(culled[a:b], want["indices"][a:b])
    want_inst = O.as_instances(ref["model"], mesh, blas, draw_index_base=base)
    assert np.array_equal(instances["transform"], want_inst["transform"])  # == on floats, like the model matrices
    for field in ("custom_index_and_mask", "sbt_offset_and_flags", "blas"):
        assert np.array_equal(instances[field], want_inst[field]), field


def _read_mgpu_result(path, n_views):
    from renderer_b200._lib import CMD_DTYPE

    b = open(path, "rb").read()
    v, n_gpus, tl = struct.unpack_from("<III", b, 0)
    assert v == n_views
    transport = b[12 : 12 + tl].decode()
    off = 12 + tl
    ranks = []
    for _ in range(n_gpus):
        views = []
        for _ in range(n_views):
            count, total = struct.unpack_from("<IQ", b, off)
            off += 12
            cmds = np.frombuffer(b, CMD_DTYPE, count, off)
            off += 20 * count
            vis = np.frombuffer(b, "<u4", count, off)
            off += 4 * count
            views.append((count, total, cmds, vis))
        ranks.append(views)
    assert off == len(b)
    return transport, ranks


def _gpu_count():
    try:
        import torch

        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.gpu
@pytest.mark.parametrize("n_gpus,n,n_lights", [(1, 40_000, 2), (2, 120_001, 4), (2, 3, 0), (4, 200_000, 1), (8, 300_007, 4)])
def test_sharded_system_over_the_gpus_of_one_process_matches_oracle(driver, tmp_path, n_gpus, n, n_lights):
    """SURVEY 8e without Python in the data plane: the C++ host stands up n_gpus ranks in ONE process through
    rcull_mgpu_create (peer access, NVLS multicast when the driver offers it), shards the query by contiguous range,
    runs two merged frames, and every rank's merged lists are bit-equal to the single-range oracle's."""
    if _gpu_count() < n_gpus:
        pytest.skip(f"needs {n_gpus} GPUs")
    pos, quat, scale, mesh = O.synth_entities(scene.SEED, 0, n, 80.0, len(MESHES))
    write_scene(tmp_path / "s.bin", pos, quat, scale, mesh, MESHES, n_lights)
    subprocess.check_call([driver, f"mgpu{n_gpus}", str(tmp_path / "s.bin"), str(tmp_path / "r.bin")], timeout=300)
    v = 1 + n_lights
    transport, ranks = _read_mgpu_result(tmp_path / "r.bin", v)
    assert len(ranks) == n_gpus and transport
    ref = O.frame(pos, quat, scale, mesh, MESHES, scene.scene_views(v), scene.CAMERA_POS, n_threads=8, want=("cmds", "visible"))
    for r in range(n_gpus):
        for k in range(v):
            count, total, cmds, vis = ranks[r][k]
            assert count == int(ref["counts"][k]) and total == int(ref["totals"][k]), (r, k)
            assert np.array_equal(vis, ref["visible"][k]) and np.array_equal(cmds, ref["cmds"][k]), (r, k, transport)
