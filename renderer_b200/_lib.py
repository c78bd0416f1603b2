"""ctypes binding of librcull.so (C ABI in include/rcull.h).

The library is the product; this module only declares its signatures.  Loading
fails loudly when the shared object is missing -- there is no Python or CPU
fallback for any compute entry point.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

MAX_VIEWS = 8
MAX_LODS = 6

PER_ENTITY = 0
PER_MESH_INSTANCED = 1

LIB_PATH = os.environ.get("RCULL_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "librcull.so")  # RCULL_LIB: A/B builds of the same ABI


class RcullError(RuntimeError):
    """A librcull call returned a negative status (the reference would panic)."""

    def __init__(self, code: int, message: str):
        super().__init__(f"rcull error {code}: {message}")
        self.code = code


class Config(C.Structure):
    _fields_ = [
        ("device", C.c_int32),
        ("max_views", C.c_uint32),
        ("capacity_entities", C.c_uint64),
        ("capacity_visible", C.c_uint64),
        ("max_meshes", C.c_uint32),
        ("output_mode", C.c_uint32),
        ("n_buffers", C.c_uint32),
        ("materialize_aabb", C.c_uint32),
        ("materialize_flags", C.c_uint32),
        ("lod_distance", C.c_float),
        ("capacity_tri_indices", C.c_uint64),
    ]


class View(C.Structure):
    _fields_ = [("planes", (C.c_float * 4) * 6)]


class TriOutputs(C.Structure):
    _fields_ = [("cmds", C.c_void_p), ("count", C.c_void_p), ("indices", C.c_void_p), ("kept_indices", C.c_void_p)]


class Outputs(C.Structure):
    _fields_ = [
        ("model", C.c_void_p),
        ("aabb", C.c_void_p),
        ("flags", C.c_void_p * MAX_VIEWS),
        ("cmds", C.c_void_p * MAX_VIEWS),
        ("visible", C.c_void_p * MAX_VIEWS),
        ("counts", C.c_void_p),
        ("index_totals", C.c_void_p),
    ]


class NoallocIO(C.Structure):
    """rcull_noalloc_io: device pointers the caller owns (rcull_frame_noalloc)."""

    _fields_ = [
        ("pos_scale", C.c_void_p),
        ("rot", C.c_void_p),
        ("mesh", C.c_void_p),
        ("model", C.c_void_p),
        ("cmds", C.c_void_p * MAX_VIEWS),
        ("visible", C.c_void_p * MAX_VIEWS),
        ("counts", C.c_void_p),
        ("index_totals", C.c_void_p),
        ("error", C.c_void_p),
        ("capacity_visible", C.c_uint64),
    ]


NOALLOC_TILE = 1920
NOALLOC_MAX_ENTITIES = 1 << 20
NOALLOC_MAX_MESHES = 256

# numpy twins of the POD structs
MESH_DTYPE = np.dtype(
    [
        ("aabb_min", "<f4", (3,)),
        ("aabb_max", "<f4", (3,)),
        ("n_lods", "<u4"),
        ("lod", "<u4", (MAX_LODS, 2)),  # (first_index, index_len)
        ("vertex_offset", "<i4"),
    ]
)
CMD_DTYPE = np.dtype(
    [
        ("indexCount", "<u4"),
        ("instanceCount", "<u4"),
        ("firstIndex", "<u4"),
        ("vertexOffset", "<i4"),
        ("firstInstance", "<u4"),
    ]
)
# VkAccelerationStructureInstanceKHR (acceleration_strucures.rs:439-449)
AS_INSTANCE_DTYPE = np.dtype(
    [("transform", "<f4", (3, 4)), ("custom_index_and_mask", "<u4"), ("sbt_offset_and_flags", "<u4"), ("blas", "<u8")]
)
assert MESH_DTYPE.itemsize == 80 and CMD_DTYPE.itemsize == 20 and AS_INSTANCE_DTYPE.itemsize == 64

# every symbol include/rcull.h declares: name -> (restype, argtypes)
_P = C.c_void_p
_u32, _u64, _f = C.c_uint32, C.c_uint64, C.c_float
SYMBOLS = {
    "rcull_last_error": (C.c_char_p, []),
    "rcull_version": (_u32, []),
    "rcull_create": (C.c_int, [C.POINTER(Config), C.POINTER(_P)]),
    "rcull_destroy": (None, [_P]),
    "rcull_set_mesh_table": (C.c_int, [_P, _P, _u32]),
    "rcull_upload_entities": (C.c_int, [_P, _u64, _u64, _P, _P, _P, _P]),
    "rcull_context_stream": (C.c_int, [_P, C.POINTER(_P)]),
    "rcull_entity_ptrs": (C.c_int, [_P, C.POINTER(_P), C.POINTER(_P), C.POINTER(_P)]),
    "rcull_pack_components": (C.c_int, [_u64, _P, _P, _P, _P, _P]),
    "rcull_frame": (C.c_int, [_P, _u64, _P, _u32, _P, _u32, _u32, _P]),
    "rcull_frame_frozen": (C.c_int, [_P, _u64, _u32, _P]),
    "rcull_frame_host": (C.c_int, [_P, _u64, _P, _P, _P, _P, _u32, _P, _u32, _u32, _P, _P]),
    "rcull_frame_host_lists": (C.c_int, [_P, _u64, _P, _P, _P, _P, _u32, _P, _u32, _u32, _P, _P, _P, _P, _u64]),
    "rcull_get_outputs": (C.c_int, [_P, _u32, C.POINTER(Outputs)]),
    "rcull_download_counts": (C.c_int, [_P, _u32, _P, _P]),
    "rcull_download_model": (C.c_int, [_P, _u32, _u64, _u64, _P]),
    "rcull_download_aabb": (C.c_int, [_P, _u32, _u64, _u64, _P]),
    "rcull_download_flags": (C.c_int, [_P, _u32, _u32, _u64, _u64, _P]),
    "rcull_download_cmds": (C.c_int, [_P, _u32, _u32, _u64, _u64, _P]),
    "rcull_download_visible": (C.c_int, [_P, _u32, _u32, _u64, _u64, _P]),
    "rcull_get_slot_counts": (C.c_int, [_P, _u32, C.POINTER(_P), C.POINTER(_u32)]),
    "rcull_download_slot_counts": (C.c_int, [_P, _u32, _P, _u32]),
    "rcull_set_geometry": (C.c_int, [_P, _P, _u64, _P, _u64]),
    "rcull_tri_cull": (C.c_int, [_P, _P, _u32, _u32, _P]),
    "rcull_get_tri_outputs": (C.c_int, [_P, C.POINTER(TriOutputs)]),
    "rcull_download_tri": (C.c_int, [_P, _P, _P, _P, _u64, _P, _u64, _u64]),
    "rcull_set_blas_table": (C.c_int, [_P, _P, _u32]),
    "rcull_as_instances": (C.c_int, [_P, _u64, _u32, _u32, _P]),
    "rcull_get_as_instances": (C.c_int, [_P, C.POINTER(_P)]),
    "rcull_download_as_instances": (C.c_int, [_P, _u64, _u64, _P]),
    "rcull_rebase_into": (C.c_int, [_P, _u32, _u32, _u64, _P, _P, _P]),
    "rcull_merge_window_bytes": (C.c_int, [_P, _u32, _u64, _P]),
    "rcull_merge_create": (C.c_int, [_P, _u32, _u32, _u64, _u64, C.POINTER(_P)]),
    "rcull_merge_create_on": (C.c_int, [_P, _u32, _u32, _u64, _u64, _P, C.POINTER(_P)]),
    "rcull_merge_destroy": (None, [_P]),
    "rcull_merge_ipc_handle": (C.c_int, [_P, _P]),
    "rcull_merge_open_peer": (C.c_int, [_P, _u32, _P]),
    "rcull_merge_set_peer_ptr": (C.c_int, [_P, _u32, _P]),
    "rcull_merge_window": (C.c_int, [_P, C.POINTER(_P)]),
    "rcull_merge_set_multicast_ptr": (C.c_int, [_P, _P]),
    "rcull_merge_set_overlap": (C.c_int, [_P, _u32]),
    "rcull_merge_set_transport": (C.c_int, [_P, _u32]),
    "rcull_merge_send": (C.c_int, [_P, _P]),
    "rcull_merge_set_send_bound": (C.c_int, [_P, _P, _u32]),
    "rcull_merge_expand": (C.c_int, [_P, _P]),
    "rcull_merge_lists": (C.c_int, [_P, _u32, C.POINTER(_P), C.POINTER(_P), C.POINTER(_P), C.POINTER(_P)]),
    "rcull_merge_download": (C.c_int, [_P, _u32, _P, _P, _P, _P, _u64]),
    "rcull_mgpu_create": (C.c_int, [C.POINTER(Config), _u32, _P, _u64, _u64, _u32, C.POINTER(_P)]),
    "rcull_mgpu_destroy": (None, [_P]),
    "rcull_mgpu_rank": (C.c_int, [_P, _u32, C.POINTER(_P), C.POINTER(_P), C.POINTER(_P)]),
    "rcull_mgpu_transport": (C.c_char_p, [_P]),
    "rcull_mgpu_last_error": (C.c_char_p, []),
    "rcull_mgpu_frame": (C.c_int, [_P, _P, _P, _P, _u32, _P]),
    "rcull_mgpu_synchronize": (C.c_int, [_P]),
    "rcull_frame_noalloc": (C.c_int, [C.c_int32, _u64, C.POINTER(NoallocIO), _P, _u32, _P, _u32, _P, _f, _u32, _P]),
    "rcull_synth_entities": (C.c_int, [_P, _u64, _u64, _u64, _f, _u32, _P]),
    "rcull_launch_count": (_u64, [_P]),
    "rcull_frustum_planes": (None, [_P, C.POINTER(View)]),
    "rcull_project_camera": (None, [_f, _P, _P, _P, _P, _P, C.POINTER(View)]),
    "rcull_light_view": (None, [_P, _P, _P, _P, _P, C.POINTER(View)]),
}

_lib = None


def load() -> C.CDLL:
    """dlopen librcull.so and type every entry point; raises if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make -C renderer_b200/csrc` "
            "(or __graft_entry__.build()); renderer_b200 has no CPU fallback"
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError = header/library drift
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        raise RcullError(rc, load().rcull_last_error().decode("utf-8", "replace"))


def ptr(a: np.ndarray | None) -> C.c_void_p:
    if a is None:
        return C.c_void_p(0)
    assert a.flags["C_CONTIGUOUS"]
    return C.c_void_p(a.ctypes.data)
