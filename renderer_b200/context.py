"""Host-side wrapper of a librcull context: one context = one GPU.

Mirrors the resources the reference's systems hold for this stage --
ModelData.model_buffer (src/renderer.rs:1225-1244) and CullPassData
(src/renderer/systems/cull_pipeline.rs:74-84) -- as device buffers owned by the
library.  Every compute call goes through the C ABI; nothing here computes.
"""
from __future__ import annotations

import ctypes as C
from typing import Sequence

import numpy as np

from . import _lib
from ._lib import CMD_DTYPE, MESH_DTYPE, Config, Outputs, RcullError, TriOutputs, View, check, ptr


def views_array(views: Sequence[np.ndarray] | np.ndarray) -> np.ndarray:
    """[n_views][6][4] float32, rcull_view layout (left,right,bottom,top,near,far)."""
    a = np.ascontiguousarray(np.asarray(views, dtype=np.float32))
    if a.ndim == 2:
        a = a[None]
    if a.shape[1:] != (6, 4):
        raise ValueError(f"views must be [n][6][4], got {a.shape}")
    return a


class CullContext:
    """rcull_create / rcull_destroy with explicit destroy(), like the reference's
    `.destroy(device)` convention (cull_pipeline.rs:276-286)."""

    def __init__(
        self,
        capacity_entities: int,
        max_views: int = 1,
        max_meshes: int = 1024,
        *,
        device: int = 0,
        capacity_visible: int = 0,
        output_mode: int = _lib.PER_ENTITY,
        n_buffers: int = 1,
        materialize_aabb: bool = False,
        materialize_flags: bool = False,
        lod_distance: float = 10.0,
        capacity_tri_indices: int = 0,
    ):
        self._lib = _lib.load()
        cfg = Config(
            device=device,
            max_views=max_views,
            capacity_entities=capacity_entities,
            capacity_visible=capacity_visible,
            max_meshes=max_meshes,
            output_mode=output_mode,
            n_buffers=n_buffers,
            materialize_aabb=int(materialize_aabb),
            materialize_flags=int(materialize_flags),
            lod_distance=lod_distance,
            capacity_tri_indices=capacity_tri_indices,
        )
        self.cfg = cfg
        self._h = C.c_void_p()
        check(self._lib.rcull_create(C.byref(cfg), C.byref(self._h)))
        self.n = 0

    # -- lifetime ---------------------------------------------------------------
    def destroy(self) -> None:
        if self._h:
            self._lib.rcull_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.destroy()

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass

    # -- inputs -----------------------------------------------------------------
    def set_mesh_table(self, meshes: np.ndarray) -> None:
        meshes = np.ascontiguousarray(meshes, dtype=MESH_DTYPE)
        check(self._lib.rcull_set_mesh_table(self._h, ptr(meshes), len(meshes)))

    def upload_entities(self, pos_scale: np.ndarray, rot: np.ndarray, mesh: np.ndarray, first: int = 0, stream: int = 0) -> None:
        pos_scale = np.ascontiguousarray(pos_scale, dtype=np.float32).reshape(-1, 4)
        rot = np.ascontiguousarray(rot, dtype=np.float32).reshape(-1, 4)
        mesh = np.ascontiguousarray(mesh, dtype=np.uint32).reshape(-1)
        n = len(mesh)
        if len(pos_scale) != n or len(rot) != n:
            raise ValueError("component columns differ in length")
        check(self._lib.rcull_upload_entities(self._h, first, n, ptr(pos_scale), ptr(rot), ptr(mesh), C.c_void_p(stream)))
        self.n = max(self.n, first + n)

    def upload_components(self, position: np.ndarray, rotation: np.ndarray, scale: np.ndarray, mesh: np.ndarray) -> None:
        """AoS bevy components (Position, Rotation, Scale; components.rs:5-10) -> SoA columns."""
        position = np.ascontiguousarray(position, dtype=np.float32).reshape(-1, 3)
        rotation = np.ascontiguousarray(rotation, dtype=np.float32).reshape(-1, 4)
        scale = np.ascontiguousarray(scale, dtype=np.float32).reshape(-1)
        n = len(scale)
        ps = np.empty((n, 4), np.float32)
        rt = np.empty((n, 4), np.float32)
        check(self._lib.rcull_pack_components(n, ptr(position), ptr(rotation), ptr(scale), ptr(ps), ptr(rt)))
        self.upload_entities(ps, rt, mesh)

    def entity_ptrs(self) -> tuple[int, int, int]:
        a, b, c = C.c_void_p(), C.c_void_p(), C.c_void_p()
        check(self._lib.rcull_entity_ptrs(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def synth_entities(self, seed: int, first_global: int, n: int, half_box: float, n_meshes: int, stream: int = 0) -> None:
        check(self._lib.rcull_synth_entities(self._h, seed, first_global, n, half_box, n_meshes, C.c_void_p(stream)))
        self.n = n

    # -- the frame --------------------------------------------------------------
    def frame(self, views, cam_pos, n: int | None = None, draw_index_base: int = 0, buffer_ix: int = 0, stream: int = 0) -> None:
        v = views_array(views)
        cam = np.ascontiguousarray(cam_pos, dtype=np.float32).reshape(3)
        n = self.n if n is None else n
        check(self._lib.rcull_frame(self._h, n, ptr(v), len(v), ptr(cam), draw_index_base, buffer_ix, C.c_void_p(stream)))
        self._last = (n, len(v))

    def frame_frozen(self, n: int | None = None, buffer_ix: int = 0, stream: int = 0) -> None:
        """FREEZE_CULLING frame (cull_pass_bypass): model matrices only, draw lists untouched."""
        check(self._lib.rcull_frame_frozen(self._h, self.n if n is None else n, buffer_ix, C.c_void_p(stream)))

    def frame_host(self, pos_scale, rot, mesh, views, cam_pos, draw_index_base: int = 0, buffer_ix: int = 0):
        """Host buffers in, counts out (the end-to-end call)."""
        v = views_array(views)
        cam = np.ascontiguousarray(cam_pos, dtype=np.float32).reshape(3)
        n = len(mesh)
        counts = np.zeros(_lib.MAX_VIEWS, np.uint32)
        totals = np.zeros(_lib.MAX_VIEWS, np.uint64)
        check(
            self._lib.rcull_frame_host(
                self._h, n, ptr(pos_scale), ptr(rot), ptr(mesh), ptr(v), len(v), ptr(cam), draw_index_base, buffer_ix,
                ptr(counts), ptr(totals),
            )
        )
        self.n = n
        self._last = (n, len(v))
        return counts[: len(v)], totals[: len(v)]

    def frame_host_lists(self, pos_scale, rot, mesh, views, cam_pos, cmds_out, visible_out, capacity: int,
                         draw_index_base: int = 0, buffer_ix: int = 0):
        """frame_host + the draw lists back: cmds_out / visible_out are sequences (one per view) of host
        addresses (int; pinned memory for full PCIe speed) each holding `capacity` entries."""
        v = views_array(views)
        cam = np.ascontiguousarray(cam_pos, dtype=np.float32).reshape(3)
        n = len(mesh)
        counts = np.zeros(_lib.MAX_VIEWS, np.uint32)
        totals = np.zeros(_lib.MAX_VIEWS, np.uint64)
        ca = (C.c_void_p * len(v))(*[C.c_void_p(int(a)) for a in cmds_out])
        va = (C.c_void_p * len(v))(*[C.c_void_p(int(a)) for a in visible_out])
        check(
            self._lib.rcull_frame_host_lists(
                self._h, n, ptr(pos_scale), ptr(rot), ptr(mesh), ptr(v), len(v), ptr(cam), draw_index_base, buffer_ix,
                ptr(counts), ptr(totals), C.cast(ca, C.c_void_p), C.cast(va, C.c_void_p), capacity,
            )
        )
        self.n = n
        self._last = (n, len(v))
        return counts[: len(v)], totals[: len(v)]

    @property
    def stream(self) -> int:
        """The context's own stream (rcull_context_stream): pass it to the asynchronous calls that follow a
        frame_host so that all per-frame work is ordered on one stream."""
        s = C.c_void_p()
        check(self._lib.rcull_context_stream(self._h, C.byref(s)))
        return int(s.value or 0)

    # -- outputs ----------------------------------------------------------------
    def outputs(self, buffer_ix: int = 0) -> Outputs:
        o = Outputs()
        check(self._lib.rcull_get_outputs(self._h, buffer_ix, C.byref(o)))
        return o

    def counts(self, buffer_ix: int = 0):
        c = np.zeros(self.cfg.max_views, np.uint32)
        t = np.zeros(self.cfg.max_views, np.uint64)
        check(self._lib.rcull_download_counts(self._h, buffer_ix, ptr(c), ptr(t)))
        return c, t

    def model(self, n: int | None = None, first: int = 0, buffer_ix: int = 0) -> np.ndarray:
        n = self.n if n is None else n
        out = np.empty((n, 16), np.float32)
        check(self._lib.rcull_download_model(self._h, buffer_ix, first, n, ptr(out)))
        return out

    def aabb(self, n: int | None = None, first: int = 0, buffer_ix: int = 0) -> np.ndarray:
        n = self.n if n is None else n
        out = np.empty((n, 6), np.float32)
        check(self._lib.rcull_download_aabb(self._h, buffer_ix, first, n, ptr(out)))
        return out

    def flags(self, view: int, n: int | None = None, first: int = 0, buffer_ix: int = 0) -> np.ndarray:
        n = self.n if n is None else n
        out = np.empty(n, np.uint8)
        check(self._lib.rcull_download_flags(self._h, buffer_ix, view, first, n, ptr(out)))
        return out

    def cmds(self, view: int, n: int, first: int = 0, buffer_ix: int = 0) -> np.ndarray:
        out = np.empty(n, CMD_DTYPE)
        check(self._lib.rcull_download_cmds(self._h, buffer_ix, view, first, n, ptr(out)))
        return out

    def visible(self, view: int, n: int, first: int = 0, buffer_ix: int = 0) -> np.ndarray:
        out = np.empty(n, np.uint32)
        check(self._lib.rcull_download_visible(self._h, buffer_ix, view, first, n, ptr(out)))
        return out

    # -- stage 2: per-triangle cull (generate_work.comp) + compact_draw_stream ------------
    def set_geometry(self, positions: np.ndarray, indices: np.ndarray) -> None:
        """Consolidated position (vec3) and index (u32) buffers (consolidate_mesh_buffers.rs:26-35)."""
        pos = np.ascontiguousarray(positions, dtype=np.float32).reshape(-1, 3)
        idx = np.ascontiguousarray(indices, dtype=np.uint32).reshape(-1)
        check(self._lib.rcull_set_geometry(self._h, ptr(pos), len(pos), ptr(idx), len(idx)))

    def tri_cull(self, pv: np.ndarray, draw_index_base: int = 0, buffer_ix: int = 0, stream: int = 0) -> None:
        """One call = cull_pass' generate_work dispatch loop + the compact_draw_stream dispatch."""
        m = np.ascontiguousarray(pv, dtype=np.float32).reshape(16)
        check(self._lib.rcull_tri_cull(self._h, ptr(m), draw_index_base, buffer_ix, C.c_void_p(stream)))

    def tri_outputs(self) -> TriOutputs:
        o = TriOutputs()
        check(self._lib.rcull_get_tri_outputs(self._h, C.byref(o)))
        return o

    def tri_counts(self) -> tuple[int, int]:
        count = np.zeros(1, np.uint32)
        kept = np.zeros(1, np.uint64)
        check(self._lib.rcull_download_tri(self._h, ptr(count), ptr(kept), None, 0, None, 0, 0))
        return int(count[0]), int(kept[0])

    def tri_cmds(self, n: int) -> np.ndarray:
        count = np.zeros(1, np.uint32)
        out = np.zeros(n, CMD_DTYPE)
        check(self._lib.rcull_download_tri(self._h, ptr(count), None, ptr(out), n, None, 0, 0))
        return out

    def tri_indices(self, first: int, n: int) -> np.ndarray:
        count = np.zeros(1, np.uint32)
        out = np.zeros(n, np.uint32)
        check(self._lib.rcull_download_tri(self._h, ptr(count), None, None, 0, ptr(out), first, n))
        return out

    def slot_counts(self, view: int, n_slots: int) -> np.ndarray:
        """Per-mesh instanced mode: survivors per (mesh, lod) slot of one view after the last frame."""
        out = np.zeros(n_slots, np.uint32)
        check(self._lib.rcull_download_slot_counts(self._h, view, ptr(out), n_slots))
        return out

    # -- acceleration-structure instance records (acceleration_strucures.rs:419-445) -------
    def set_blas_table(self, addresses: np.ndarray) -> None:
        a = np.ascontiguousarray(addresses, dtype=np.uint64)
        check(self._lib.rcull_set_blas_table(self._h, ptr(a), len(a)))

    def as_instances(self, n: int | None = None, draw_index_base: int = 0, buffer_ix: int = 0, stream: int = 0) -> None:
        check(self._lib.rcull_as_instances(self._h, self.n if n is None else n, draw_index_base, buffer_ix, C.c_void_p(stream)))

    def download_as_instances(self, n: int, first: int = 0) -> np.ndarray:
        out = np.zeros(n, _lib.AS_INSTANCE_DTYPE)
        check(self._lib.rcull_download_as_instances(self._h, first, n, ptr(out)))
        return out

    def rebase_into(self, view: int, index_prefix: int, dst_cmds_ptr: int, dst_visible_ptr: int, buffer_ix: int = 0, stream: int = 0) -> None:
        check(self._lib.rcull_rebase_into(self._h, buffer_ix, view, index_prefix, C.c_void_p(dst_cmds_ptr), C.c_void_p(dst_visible_ptr), C.c_void_p(stream)))

    @property
    def launch_count(self) -> int:
        return int(self._lib.rcull_launch_count(self._h))


__all__ = ["CullContext", "RcullError", "views_array", "View"]
