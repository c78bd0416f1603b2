"""CPU tests of the C-ABI boundary: librcull.so loads, exports every symbol that
include/rcull.h declares, agrees with the header on struct layouts, fails loudly
without a GPU, and its SASS has the properties the parity argument relies on."""
import ctypes as C
import os
import re
import shutil
import subprocess

import numpy as np
import pytest

import renderer_b200
from renderer_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "rcull.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rcull_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound():
    lib = renderer_b200.load()
    names = declared_functions()
    assert len(names) >= 20
    for name in names:
        assert hasattr(lib, name), f"{name} declared in rcull.h but not exported by librcull.so"
        assert name in _lib.SYMBOLS, f"{name} declared in rcull.h but not bound in renderer_b200/_lib.py"
    assert set(_lib.SYMBOLS) == set(names)


def test_struct_layouts_match_the_header():
    assert _lib.MESH_DTYPE.itemsize == 80  # rcull_mesh
    assert _lib.CMD_DTYPE.itemsize == 20  # VkDrawIndexedIndirectCommand stride (generate_work.comp:9-15)
    assert C.sizeof(_lib.View) == 96
    assert C.sizeof(_lib.Config) == 56
    assert C.sizeof(_lib.TriOutputs) == 32
    assert C.sizeof(_lib.Outputs) == 8 * (2 + 3 * _lib.MAX_VIEWS + 2)


def test_version_and_error_string():
    lib = renderer_b200.load()
    assert lib.rcull_version() >= 1
    assert isinstance(lib.rcull_last_error(), bytes)


def test_pack_components_is_host_only_and_exact():
    lib = renderer_b200.load()
    n = 1000
    rng = np.random.default_rng(0)
    pos, rot, scale = rng.standard_normal((n, 3), np.float32), rng.standard_normal((n, 4), np.float32), rng.random(n, np.float32)
    ps, rt = np.empty((n, 4), np.float32), np.empty((n, 4), np.float32)
    assert lib.rcull_pack_components(n, _lib.ptr(pos), _lib.ptr(rot), _lib.ptr(scale), _lib.ptr(ps), _lib.ptr(rt)) == 0
    assert np.array_equal(ps[:, :3], pos) and np.array_equal(ps[:, 3], scale) and np.array_equal(rt, rot)
    assert lib.rcull_pack_components(n, None, None, None, None, None) < 0  # null arguments are an error


def _have_gpu():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_have_gpu(), reason="checks the no-device failure mode")
def test_create_fails_loudly_without_a_device():
    with pytest.raises(renderer_b200.RcullError) as e:
        renderer_b200.CullContext(1000)
    assert e.value.code == -3  # RCULL_ERR_NO_DEVICE: there is no CPU path
    assert b"no CPU path" in renderer_b200.load().rcull_last_error() or "no CPU path" in str(e.value)


def test_invalid_config_is_rejected_before_touching_a_device():
    lib = renderer_b200.load()
    h = C.c_void_p()
    cfg = _lib.Config(device=0, max_views=9, capacity_entities=10, capacity_visible=0, max_meshes=1, output_mode=0,
                      n_buffers=1, materialize_aabb=0, materialize_flags=0, lod_distance=10.0)
    assert lib.rcull_create(C.byref(cfg), C.byref(h)) == -1 and not h.value
    cfg.max_views, cfg.n_buffers = 1, 4
    assert lib.rcull_create(C.byref(cfg), C.byref(h)) == -1
    assert lib.rcull_create(None, C.byref(h)) == -1


def test_noalloc_entry_point_checks_its_limits_before_touching_a_device():
    """rcull_frame_noalloc (no counterpart in the reference): what its 128 KiB of static scratch cannot hold is rejected
    with a message naming the limit -- argument checks come first, so this needs no GPU; with valid arguments and no device
    it fails like every other entry point (there is no CPU path)."""
    lib = renderer_b200.load()
    assert C.sizeof(_lib.NoallocIO) == 8 * (4 + 2 * _lib.MAX_VIEWS + 3) + 8
    io = _lib.NoallocIO()
    io.pos_scale = io.rot = io.mesh = io.model = 0x1000
    for k in range(_lib.MAX_VIEWS):
        io.cmds[k] = io.visible[k] = 0x1000
    io.counts = io.index_totals = io.error = 0x1000
    io.capacity_visible = 16
    meshes = np.zeros(1, _lib.MESH_DTYPE)
    meshes["n_lods"] = 1
    views = np.zeros((1, 6, 4), np.float32)
    cam = np.zeros(3, np.float32)

    def call(n, io_=io, m=meshes, n_meshes=1, n_views=1):
        return lib.rcull_frame_noalloc(0, n, C.byref(io_) if io_ is not None else None, _lib.ptr(m), n_meshes, _lib.ptr(views), n_views,
                                       _lib.ptr(cam), 10.0, 0, None)

    assert call(10, io_=None) == -1
    assert call(_lib.NOALLOC_MAX_ENTITIES + 1) == -1 and b"RCULL_NOALLOC_MAX_ENTITIES" in lib.rcull_last_error()
    assert call(10, n_meshes=_lib.NOALLOC_MAX_MESHES + 1) == -1 and b"RCULL_NOALLOC_MAX_MESHES" in lib.rcull_last_error()
    assert call(10, n_views=9) == -1
    io.rot = 0x1004
    assert call(10) == -1 and b"aligned" in lib.rcull_last_error()
    io.rot = 0x1000
    if not _have_gpu():
        assert call(10) == -3  # RCULL_ERR_NO_DEVICE


def test_merge_transport_entry_points_reject_null_handles():
    lib = renderer_b200.load()
    assert lib.rcull_merge_set_transport(None, 1) == -1
    assert lib.rcull_merge_send(None, None) == -1
    assert lib.rcull_merge_set_send_bound(None, None, 0) == -1
    assert lib.rcull_merge_expand(None, None) == -1


CUOBJDUMP = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"


@pytest.mark.skipif(not os.path.exists(CUOBJDUMP), reason="cuobjdump not available")
def test_sass_of_the_frame_kernel():
    """The frame kernel must (a) contain no fused multiply-add -- the reference never
    contracts a*b+c and ptxas contracts packed f32x2 mul+add even under --fmad=false --
    (b) be built for sm_100a, (c) use the TMA engine for loads and stores."""
    out = subprocess.run([CUOBJDUMP, "-sass", _lib.LIB_PATH], capture_output=True, text=True, check=True).stdout
    assert "sm_100a" in out or "sm_100" in out
    funcs = {}
    cur = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
        elif cur and re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
            funcs[cur].append(line)
    frame = {k: v for k, v in funcs.items() if "transform_cull_compact" in k}
    assert len(frame) == 16  # 8 view counts x {hot, materialising}
    for name, body in frame.items():
        text = "\n".join(body)
        assert not re.search(r"\bFFMA2?\b", text), f"fused multiply-add in {name}"
        assert "UBLKCP" in text, f"no bulk TMA load in {name}"
        assert "UTMASTG" in text, f"no TMA tensor store in {name}"
        assert "FMUL2" in text and "SYNCS" in text


def test_rust_ffi_file_is_generated_from_the_current_header():
    """integration/rust/rcull_ffi.rs (source only: no rustc here) declares every entry point of rcull.h and is
    what gen_ffi.py produces from the header as it stands."""
    import importlib.util

    rust = os.path.join(os.path.dirname(HEADER), "..", "integration", "rust")
    spec = importlib.util.spec_from_file_location("gen_ffi", os.path.join(rust, "gen_ffi.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    text = open(os.path.join(rust, "rcull_ffi.rs")).read()
    assert text == gen.generate(), "run python integration/rust/gen_ffi.py"
    declared = re.findall(r"pub fn (rcull_\w+)\(", text)
    assert sorted(declared) == declared_functions()
    for name, (res, args) in _lib.SYMBOLS.items():  # same arity as the ctypes binding
        params = re.search(r"pub fn %s\((.*?)\)( -> [^;]+)?;" % name, text)
        assert len([p for p in params.group(1).split(",") if p.strip()]) == len(args), name
        assert (params.group(2) is None) == (res is None), name


@pytest.mark.skipif(not os.path.exists(CUOBJDUMP), reason="cuobjdump not available")
def test_sass_of_the_bucketing_chain_waits_for_its_predecessor():
    """The per-mesh bucketing kernels are programmatic dependent launches: each must execute griddepcontrol.wait (SASS
    ACQBULK) before it reads what the kernel in front of it wrote; the command pass lives in bucket_colscan now."""
    out = subprocess.run([CUOBJDUMP, "-sass", _lib.LIB_PATH], capture_output=True, text=True, check=True).stdout
    funcs, cur = {}, None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
        elif cur:
            funcs[cur].append(line)
    for name in ("bucket_hist", "bucket_colscan", "bucket_scatter"):
        body = [v for k, v in funcs.items() if name in k]
        assert len(body) == 1, name
        assert any("ACQBULK" in l for l in body[0]), f"{name} does not wait for its predecessor grid"
    assert not any("bucket_cmds" in k for k in funcs)
