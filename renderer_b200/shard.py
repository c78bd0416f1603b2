"""Range sharding across GPUs and the one exchange step.

Entities are split into contiguous ranges, one per rank; every rank runs the
single-GPU frame on its range with global draw indices (draw_index_base = lo)
and LOCAL firstIndex.  The merge is:

  1. all_gather of per-rank, per-view (count, index_total);
  2. exclusive prefixes over ranks -> every rank's slot in the merged lists and
     the amount to add to its firstIndex;
  3. each rank rebases its list straight into its slot of the merged buffers
     (rcull_rebase_into, a device kernel) and the slices are exchanged in place
     with one broadcast per (rank, view, buffer) -- NCCL has no allgatherv.

Because slots are ordered by rank and ranges are ascending, the merged list is
byte-identical to the single-GPU list.  The planning arithmetic is plain numpy
and backend-agnostic; torch.distributed (nccl on GPUs, gloo in CPU tests) only
moves bytes.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


def shard_range(n_total: int, world: int, rank: int) -> tuple[int, int]:
    """Contiguous range [lo, hi) of rank; ragged when world does not divide n_total."""
    return n_total * rank // world, n_total * (rank + 1) // world


@dataclass
class MergePlan:
    count_prefix: np.ndarray  # [world, views] u64: slot start (in commands) of each rank
    index_prefix: np.ndarray  # [world, views] u64: sum of earlier ranks' index totals
    counts: np.ndarray  # [views] merged survivor counts
    totals: np.ndarray  # [views] merged index totals


def plan_merge(counts_all: np.ndarray, totals_all: np.ndarray) -> MergePlan:
    """counts_all/totals_all: [world, views] as gathered from every rank."""
    c = np.asarray(counts_all, np.uint64)
    t = np.asarray(totals_all, np.uint64)
    cp = np.zeros_like(c)
    tp = np.zeros_like(t)
    cp[1:] = np.cumsum(c, axis=0)[:-1]
    tp[1:] = np.cumsum(t, axis=0)[:-1]
    return MergePlan(cp, tp, c.sum(axis=0), t.sum(axis=0))


class ShardMerger:
    """Owns the merged draw lists of one rank and runs the exchange.

    merged_cmds[k] is an int32 tensor of 5 * capacity words (20-B commands),
    merged_visible[k] an int32 tensor of capacity words, on `device`.
    """

    def __init__(self, n_views: int, capacity: int, device, group=None):
        import torch
        import torch.distributed as dist

        self.torch, self.dist = torch, dist
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.n_views = n_views
        self.capacity = capacity
        self.device = device
        self.merged_cmds = [torch.empty(capacity * 5, dtype=torch.int32, device=device) for _ in range(n_views)]
        self.merged_visible = [torch.empty(capacity, dtype=torch.int32, device=device) for _ in range(n_views)]
        self._meta = torch.zeros(2 * n_views, dtype=torch.int64, device=device)
        self._meta_flat = torch.zeros(self.world * 2 * n_views, dtype=torch.int64, device=device)  # gloo wants a flat output
        self._meta_all = self._meta_flat.view(self.world, 2 * n_views)
        self.plan: MergePlan | None = None
        self.bytes_received = 0

    def gather_meta(self, counts_dev, totals_dev) -> MergePlan:
        """counts_dev/totals_dev: device tensors (or numpy) with this rank's per-view
        counts and index totals.  One small all_gather + one host read."""
        torch = self.torch
        v = self.n_views
        self._meta[:v] = torch.as_tensor(counts_dev, device=self.device)[:v].to(torch.int64)
        self._meta[v:] = torch.as_tensor(totals_dev, device=self.device)[:v].to(torch.int64)
        self.dist.all_gather_into_tensor(self._meta_flat, self._meta, group=self.group)
        m = self._meta_all.cpu().numpy().astype(np.uint64)
        self.plan = plan_merge(m[:, :v], m[:, v:])
        if int(self.plan.counts.max(initial=0)) > self.capacity:
            raise ValueError(f"merged draw list of {int(self.plan.counts.max())} commands exceeds capacity {self.capacity}")
        return self.plan

    def exchange(self, counts_all: np.ndarray) -> None:
        """In-place allgatherv: rank r's slot of every merged list is broadcast from r.
        Call after this rank wrote its own slot (rebase step)."""
        dist, plan = self.dist, self.plan
        works = []
        recv = 0
        for k in range(self.n_views):
            for r in range(self.world):
                n = int(counts_all[r, k])
                if n == 0:
                    continue
                off = int(plan.count_prefix[r, k])
                src = dist.get_global_rank(self.group, r) if self.group is not None else r
                works.append(dist.broadcast(self.merged_cmds[k][off * 5 : (off + n) * 5], src=src, group=self.group, async_op=True))
                works.append(dist.broadcast(self.merged_visible[k][off : off + n], src=src, group=self.group, async_op=True))
                if r != self.rank:
                    recv += n * 24
        for w in works:
            w.wait()
        self.bytes_received = recv

    def merge_from_context(self, ctx, buffer_ix: int = 0, stream: int = 0) -> MergePlan:
        """GPU path: meta exchange, device-side rebase into the merged slot, exchange."""
        torch = self.torch
        o = ctx.outputs(buffer_ix)
        v = self.n_views
        counts = _wrap_device(torch, o.counts, v, torch.int32, self.device)
        totals = _wrap_device(torch, o.index_totals, v, torch.int64, self.device)
        plan = self.gather_meta(counts, totals)
        m = self._meta_all.cpu().numpy().astype(np.uint64)
        for k in range(v):
            off = int(plan.count_prefix[self.rank, k])
            ctx.rebase_into(
                k,
                int(plan.index_prefix[self.rank, k]) & 0xFFFFFFFF,
                self.merged_cmds[k].data_ptr() + off * 20,
                self.merged_visible[k].data_ptr() + off * 4,
                buffer_ix=buffer_ix,
                stream=stream,
            )
        self.exchange(m[:, :v])
        return plan


# ---- per-mesh instanced mode (BASELINE configs[3]) over range shards: SURVEY 8e, last paragraph ----
# Every rank buckets its own survivors by (mesh, lod) slot; the merged output is the single-GPU one: one
# command per non-empty slot with instanceCount summed over ranks, and the visible list slot-major with the
# ranks' buckets of a slot in rank order (ranges ascend with the rank, so that IS draw_index order).  No
# reduction collective: an all_gather of the per-slot counts (S words per rank) gives every rank all it
# needs to place every bucket.
@dataclass
class PerMeshMergePlan:
    counts_all: np.ndarray  # [world, S] as gathered
    slot_counts: np.ndarray  # [S] merged survivors per slot
    slot_start: np.ndarray  # [S] exclusive prefix over slots = firstInstance of the slot's command
    dest: np.ndarray  # [world, S] start of rank r's bucket of slot s in the MERGED visible list
    src: np.ndarray  # [world, S] start of that bucket in rank r's OWN bucketed visible list
    rank_base: np.ndarray  # [world] start of rank r's list in the rank-ordered staging buffer
    total: int  # merged survivors


def _exclusive(a: np.ndarray, axis: int) -> np.ndarray:
    out = np.zeros_like(a)
    sl_dst = [slice(None)] * a.ndim
    sl_src = [slice(None)] * a.ndim
    sl_dst[axis], sl_src[axis] = slice(1, None), slice(None, -1)
    out[tuple(sl_dst)] = np.cumsum(a, axis=axis)[tuple(sl_src)]
    return out


def plan_merge_per_mesh(slot_counts_all: np.ndarray) -> PerMeshMergePlan:
    """slot_counts_all: [world, S] survivors of every rank in every (mesh, lod) slot, as gathered."""
    c = np.asarray(slot_counts_all, np.uint64)
    totals = c.sum(axis=0)
    start = _exclusive(totals, 0)
    per_rank = c.sum(axis=1)
    return PerMeshMergePlan(c, totals, start, start[None, :] + _exclusive(c, 0), _exclusive(c, 1), _exclusive(per_rank, 0), int(per_rank.sum()))


def per_mesh_commands(meshes: np.ndarray, plan: PerMeshMergePlan, max_lods: int = 6) -> np.ndarray:
    """The merged command list: one command per non-empty slot, ascending slot = mesh * 6 + lod
    (the keep rule of compact_draw_stream.comp:41 on instanceCount, as in the single-GPU mode)."""
    from ._lib import CMD_DTYPE

    keep = np.nonzero(plan.slot_counts)[0]
    mesh, lod = keep // max_lods, keep % max_lods
    out = np.zeros(len(keep), CMD_DTYPE)
    out["indexCount"] = meshes["lod"][mesh, lod, 1]
    out["instanceCount"] = plan.slot_counts[keep].astype(np.uint32)
    out["firstIndex"] = meshes["lod"][mesh, lod, 0]
    out["vertexOffset"] = meshes["vertex_offset"][mesh]
    out["firstInstance"] = plan.slot_start[keep].astype(np.uint32)
    return out


def per_mesh_permutation(plan: PerMeshMergePlan) -> np.ndarray:
    """perm with merged[j] = staged[perm[j]], staged = the ranks' own bucketed lists back to back in rank order."""
    seg_len = plan.counts_all.astype(np.int64).reshape(-1)  # one segment per (rank, slot) bucket
    seg_dst = plan.dest.astype(np.int64).reshape(-1)
    seg_src = (plan.rank_base.astype(np.int64)[:, None] + plan.src.astype(np.int64)).reshape(-1)
    within = np.arange(int(seg_len.sum()), dtype=np.int64) - np.repeat(np.cumsum(seg_len) - seg_len, seg_len)
    perm = np.empty(plan.total, np.int64)
    perm[np.repeat(seg_dst, seg_len) + within] = np.repeat(seg_src, seg_len) + within
    return perm


class PerMeshShardMerger:
    """REFERENCE of the per-mesh merge order for the CPU (gloo, world 2 / 3) tests -- not the GPU path.  On GPUs the
    merge is rcull_merge_expand: the ranks' (slot, draw_index) records arrive in every window, are concatenated in rank
    order (= draw order) and bucketed by the single-GPU counting sort, all on the device (RecordMerger).  This class
    states the same result with collectives and a permutation: all_gather of the slot counts, in-place allgatherv of the
    ranks' bucketed visible lists into a rank-ordered staging buffer, then the slot-major permutation."""

    def __init__(self, n_slots: int, capacity: int, device, group=None):
        import torch
        import torch.distributed as dist

        self.torch, self.dist, self.group = torch, dist, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.n_slots, self.capacity, self.device = n_slots, capacity, device
        self._counts = torch.zeros(n_slots, dtype=torch.int64, device=device)
        self._counts_flat = torch.zeros(self.world * n_slots, dtype=torch.int64, device=device)
        self.staged = torch.empty(capacity, dtype=torch.int32, device=device)
        self.plan: PerMeshMergePlan | None = None

    def gather_counts(self, slot_counts) -> PerMeshMergePlan:
        torch = self.torch
        self._counts[:] = torch.as_tensor(slot_counts, device=self.device).to(torch.int64)
        self.dist.all_gather_into_tensor(self._counts_flat, self._counts, group=self.group)
        self.plan = plan_merge_per_mesh(self._counts_flat.view(self.world, self.n_slots).cpu().numpy())
        if self.plan.total > self.capacity:
            raise ValueError(f"merged visible list of {self.plan.total} entries exceeds capacity {self.capacity}")
        return self.plan

    def merge_visible(self, own_visible):
        """own_visible: this rank's bucketed visible list (int32 tensor on `device`).  Returns the merged list."""
        torch, dist, plan = self.torch, self.dist, self.plan
        bounds = np.append(plan.rank_base, plan.total).astype(np.int64)
        lo, hi = int(bounds[self.rank]), int(bounds[self.rank + 1])
        self.staged[lo:hi] = own_visible[: hi - lo]
        works = []
        for r in range(self.world):
            a, b = int(bounds[r]), int(bounds[r + 1])
            if b > a:
                src = dist.get_global_rank(self.group, r) if self.group is not None else r
                works.append(dist.broadcast(self.staged[a:b], src=src, group=self.group, async_op=True))
        for w in works:
            w.wait()
        perm = torch.as_tensor(per_mesh_permutation(plan), device=self.device)
        return self.staged[: plan.total].index_select(0, perm)


def _wrap_device(torch, ptr: int, n: int, dtype, device):
    """Zero-copy torch view of a librcull device buffer (for NCCL plumbing only)."""
    itemsize = torch.empty((), dtype=dtype).element_size()
    kind = "f" if dtype.is_floating_point else "i"

    class _Holder:
        __cuda_array_interface__ = {
            "shape": (n,),
            "typestr": "<%s%d" % (kind, itemsize),
            "data": (int(ptr), False),
            "version": 3,
            "strides": None,
        }

    return torch.as_tensor(_Holder(), device=device)


class RecordMerger:
    """The merge over peer memory (rcull_merge_*, include/rcull.h): once attached, every frame of `ctx` sends its
    survivors as 8-byte records straight from the frame kernel into every rank's window (NVLink peer stores, or one
    multimem.st through an NVLS multicast mapping), and expand() rebuilds the merged draw lists on this GPU.
    Python only bootstraps the windows (IPC handles over torch.distributed, or torch symmetric memory); nothing in a
    step touches the host or calls a collective."""

    def __init__(self, ctx, capacity: int, capacity_per_rank: int, device, group=None, symmetric: bool = False):
        import ctypes as C

        import torch
        import torch.distributed as dist

        from . import _lib

        self.torch, self.dist, self.C, self._lib = torch, dist, C, _lib.load()
        self.ctx, self.group, self.capacity, self.device = ctx, group, capacity, device
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self._h = C.c_void_p()
        self.transport = "cuda-ipc peer stores"
        self._symm = None
        if symmetric and self.world > 1:
            # Symmetric memory: every rank maps every rank's window AND one NVLS multicast address that stands for all
            # of them -- a record then leaves the GPU once and the NVSwitch replicates it.
            try:
                import torch.distributed._symmetric_memory as symm

                nbytes = C.c_uint64()
                _lib.check(self._lib.rcull_merge_window_bytes(ctx._h, self.world, capacity_per_rank, C.byref(nbytes)))
                buf = symm.empty(int(nbytes.value), dtype=torch.uint8, device=device)
                g = group if group is not None else dist.group.WORLD
                hdl = symm.rendezvous(buf, g.group_name)
                ptrs = [int(p) for p in hdl.buffer_ptrs]
                _lib.check(self._lib.rcull_merge_create_on(ctx._h, self.world, self.rank, capacity, capacity_per_rank, C.c_void_p(ptrs[self.rank]), C.byref(self._h)))
                for p in range(self.world):
                    _lib.check(self._lib.rcull_merge_set_peer_ptr(self._h, p, C.c_void_p(ptrs[p])))
                mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
                if mc:
                    _lib.check(self._lib.rcull_merge_set_multicast_ptr(self._h, C.c_void_p(mc)))
                self.transport = "symmetric memory, " + ("NVLS multicast stores" if mc else "peer stores")
                self._symm = (buf, hdl)
            except Exception as e:  # no symmetric-memory support in this torch / on this fabric: CUDA IPC below
                if self._h:
                    self._lib.rcull_merge_destroy(self._h)
                self._symm = None
                self._h = C.c_void_p()
                self.transport = f"cuda-ipc peer stores (symmetric memory unavailable: {type(e).__name__})"
        if not self._h:
            _lib.check(self._lib.rcull_merge_create(ctx._h, self.world, self.rank, capacity, capacity_per_rank, C.byref(self._h)))
        if self.world > 1 and self._symm is not None:
            dist.barrier(group=group)  # every mailbox is cleared before anyone stores into it
        elif self.world > 1:  # one-time exchange of the IPC handles of the windows
            handle = (C.c_ubyte * 64)()
            _lib.check(self._lib.rcull_merge_ipc_handle(self._h, handle))
            handles = [None] * self.world
            dist.all_gather_object(handles, bytes(handle), group=group)
            for p, h in enumerate(handles):
                if p != self.rank:
                    buf = (C.c_ubyte * 64).from_buffer_copy(h)
                    _lib.check(self._lib.rcull_merge_open_peer(self._h, p, buf))
            dist.barrier(group=group)

    def destroy(self):
        if self._h:
            if self.world > 1:
                self.torch.cuda.synchronize(self.device)
                self.dist.barrier(group=self.group)  # nobody may still be storing into this window
            self._lib.rcull_merge_destroy(self._h)
            self._h = self.C.c_void_p()

    def set_overlap(self, expand_sms: int) -> None:
        """expand_sms > 0: the frame kernel leaves that many SMs free and expand() launches as many whole-SM CTAs (issue
        it on a second stream after the frame's event: it then runs beside the next frame); 0: full-chip expand."""
        from ._lib import check

        check(self._lib.rcull_merge_set_overlap(self._h, int(expand_sms)))

    def set_transport(self, copy_engines: bool) -> None:
        """True: the frame kernel keeps its records in its own window and send() moves them to the peers with the copy
        engines (no SM store crosses NVLink); False (default): the frame kernel stores into every window itself."""
        from ._lib import check

        check(self._lib.rcull_merge_set_transport(self._h, 1 if copy_engines else 0))
        self.copy_engines = bool(copy_engines)
        base = self.transport.split(" + ")[0]
        self.transport = base + (" + copy engines (rcull_merge_send)" if copy_engines else "")

    def set_send_bound(self, records_per_view) -> None:
        """Copy-engine transport: estimated survivors per view (+ margin) that a send copies; what a frame produces past
        it stays in this rank's window and the receivers read it from there.  None / empty removes the bounds."""
        from ._lib import check, ptr

        b = np.ascontiguousarray(records_per_view if records_per_view is not None else [], dtype=np.uint32)
        check(self._lib.rcull_merge_set_send_bound(self._h, ptr(b) if len(b) else None, len(b)))

    def send(self, stream: int) -> None:
        """Copy-engine transport: queue the peer copies of the last frame's records on `stream` (the library makes the
        stream wait for that frame, and a later frame for these copies)."""
        from ._lib import check

        check(self._lib.rcull_merge_send(self._h, self.C.c_void_p(stream)))

    def expand(self, stream: int | None = None) -> None:
        """Enqueue the expand of the oldest unexpanded merged frame (default: the current torch stream)."""
        from ._lib import check

        if stream is None:
            stream = self.torch.cuda.current_stream(self.device).cuda_stream
        check(self._lib.rcull_merge_expand(self._h, self.C.c_void_p(stream)))

    def lists(self, view: int):
        """(cmds_ptr, visible_ptr, counts_ptr, totals_ptr) device addresses of the merged outputs."""
        from ._lib import check

        C = self.C
        a, b, c, d = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
        check(self._lib.rcull_merge_lists(self._h, view, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        return a.value, b.value, c.value, d.value

    def download(self, view: int, per_mesh: bool = False):
        """Blocking read-back of one merged view (tests): cmds, visible, count, index total."""
        torch = self.torch
        from ._lib import CMD_DTYPE

        cp, vp, np_, tp = self.lists(view)
        torch.cuda.synchronize(self.device)
        mv = self.ctx.cfg.max_views
        counts = _wrap_device(torch, np_, mv, torch.int32, self.device).cpu().numpy().view(np.uint32)
        totals = _wrap_device(torch, tp, mv, torch.int64, self.device).cpu().numpy().view(np.uint64)
        n = int(min(counts[view], self.capacity))
        n_vis = int(min(totals[view], self.capacity)) if per_mesh else n
        cmds = _wrap_device(torch, cp, max(n, 1) * 5, torch.int32, self.device).cpu().numpy()[: n * 5].view(CMD_DTYPE)
        vis = _wrap_device(torch, vp, max(n_vis, 1), torch.int32, self.device).cpu().numpy()[:n_vis].view(np.uint32)
        return cmds, vis, int(counts[view]), int(totals[view])
