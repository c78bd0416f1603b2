#!/usr/bin/env python
"""bench.py -- entities/s through transform + cull + compact (BASELINE.json metric).

One "step" = one pass of the hot path (rcull_frame; plus the draw-list merge when
N > 1) over one batch of synthetic entities.  Workload (config.workload):
  16m_5view  BASELINE configs[2]: 16 Mi entities PER GPU, camera + 4 shadow views
             (the configuration north_star's 70 %-of-roofline target is quoted on;
             at N > 1 it becomes configs[4]: range shards + NCCL merge, weak scaling)
  1m_1view   BASELINE configs[1]: 1 M entities, camera only (fits in L2: flushed, see flush_l2)
Prints ONE JSON line (rank 0).  `--impl reference` times the CPU oracle -- the
reference's Rust/bevy code cannot be built on this box -- on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "16m_5view": dict(n=16 * 1024 * 1024, n_views=5, half_box=600.0),
    "16m_1view": dict(n=16 * 1024 * 1024, n_views=1, half_box=600.0),
    "1m_1view": dict(n=1_000_000, n_views=1, half_box=250.0),
    # BASELINE configs[1] through BOTH stages of the reference's cull pass: the frame kernel, then the
    # per-triangle cull (generate_work.comp) of the camera survivors and compact_draw_stream
    "1m_tri": dict(n=1_000_000, n_views=1, half_box=250.0, tri=True, coherent=True),
    "1m_tri_random": dict(n=1_000_000, n_views=1, half_box=250.0, tri=True, coherent=False),  # random index triples: gather worst case
}
METRIC = "entities/sec transformed+culled+compacted"
UNIT = "entities/s"
SM_ISSUE_SLOTS = 148 * 4  # warp instructions per clock the chip can issue (one per SM sub-partition)


L2_NOTE = "flushed between iterations (256 MB write, then a 256 MB read so that the timed kernel does not pay for writing the flush's dirty lines back)"


def flush_l2(buf):
    """buf: [2, 256 MB] uint8.  The write evicts the workload from L2; the read of the other half replaces
    the write's dirty lines with clean ones (otherwise their write-back is charged to the next kernel)."""
    import torch

    buf[0].zero_()
    buf[1].view(torch.int64).sum()


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def load_counters():
    """ncu-derived constants (profiles/r02_counters.json, written from the committed captures): DRAM traffic per
    launch and executed warp instructions per unit of work.  They are cross-references measured under the profiler,
    NOT measured in this run -- every use carries its `source` label."""
    for name in ("r02_counters.json",):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                return json.load(f)
        except Exception:
            pass
    return {}


def oracle_views(n_views: int):
    """The views of renderer_b200.scene.scene_views, computed by the ORACLE's project_camera / light_view (bit-equal to
    the library's host helpers, tests/test_oracle.py): the reference arm maps no product library."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    from renderer_b200 import scene  # constants and numpy helpers only: nothing here loads librcull.so

    vs = [O.project_camera(16.0 / 9.0, scene.CAMERA_POS, scene.CAMERA_ROT)[3]]
    for k in range(n_views - 1):
        pos = np.ascontiguousarray(scene.LIGHT_POS[k % len(scene.LIGHT_POS)] * (1.0 + 0.25 * (k // len(scene.LIGHT_POS))), np.float32)
        vs.append(O.light_view(pos, scene.quat_look_at_lh(-pos.astype(np.float64)))[3])
    return np.stack(vs).astype(np.float32)


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region.  NVML is polled in-process
    (the timed region of the 16 M workload lasts tens of milliseconds: too short for an
    nvidia-smi subprocess); nvidia-smi is the fallback if NVML cannot be loaded."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index: int):
        self.index, self.samples, self._stop = index, [], threading.Event()
        self._t = threading.Thread(target=self._run, daemon=True)
        self.max_mhz = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
            self.how = "nvml"
        except Exception:
            self._nv = None
            self.how = "nvidia-smi"

    def _run(self):
        while not self._stop.is_set():
            try:
                if self._nv is not None:
                    mhz = float(self._nv.nvmlDeviceGetClockInfo(self._h, self._nv.NVML_CLOCK_SM))
                    mask = int(self._nv.nvmlDeviceGetCurrentClocksEventReasons(self._h))
                    self.samples.append((mhz, mask))
                    self._stop.wait(0.0005)
                else:
                    q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
                    out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                         capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                    bits = [0x8, 0x40, 0x20, 0x4]
                    mask = sum(b for b, x in zip(bits, out[2:6]) if x.strip().lower().startswith("active"))
                    self.max_mhz = float(out[1])
                    self.samples.append((float(out[0]), mask))
            except Exception:
                self._stop.wait(0.01)

    def __enter__(self):
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        sm = [s[0] for s in self.samples]
        mask = 0
        for s in self.samples:
            mask |= s[1]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(v for k, v in self.REASONS.items() if mask & k), "samples": len(self.samples), "how": self.how}


def cpu_baseline(wl, steps: int, warmup: int, threads: int, fused_steps: int = 3):
    """The oracle (kind "port") on the WHOLE workload (the same entities, views and mesh table the GPU arm runs; the
    sample is bounded through the step count), all host threads.  Inputs and planes come from the oracle's own
    generator and project_camera / light_view, so this leg maps no product library."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O  # CPU-baseline leg: the one place bench.py may execute oracle/
    from renderer_b200 import scene  # numpy-only helpers (mesh table, constants)

    n = wl["n"]
    meshes = scene.synth_mesh_table()
    pos, quat, scale, mesh = O.synth_entities(scene.SEED, 0, n, wl["half_box"], len(meshes))
    views = oracle_views(wl["n_views"])
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        r = O.frame(pos, quat, scale, mesh, meshes, views, scene.CAMERA_POS, n_threads=threads, want=("cmds", "visible"))
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    t = float(np.mean(times))
    # the same oracle functions on a fused schedule (one pass per entity, parallel emit) --
    # the stronger CPU baseline, reported beside the reference-shaped one
    fused = []
    for i in range(1 + fused_steps):
        t0 = time.perf_counter()
        O.frame(pos, quat, scale, mesh, meshes, views, scene.CAMERA_POS, n_threads=threads, want=("cmds", "visible"), fused=True)
        if i >= 1:
            fused.append(time.perf_counter() - t0)
    tf = float(np.mean(fused))
    # "1 thread and all online cores" -- one untimed + one timed reference-shaped step on a single thread
    n1 = min(n, 1 << 19)
    one = []
    for i in range(2):
        t0 = time.perf_counter()
        O.frame(pos[:n1], quat[:n1], scale[:n1], mesh[:n1], meshes, views, scene.CAMERA_POS, n_threads=1, want=("cmds", "visible"))
        one.append(time.perf_counter() - t0)
    return {"value": n / t, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"all {n} entities of the workload, {wl['n_views']} views, {steps} steps, reference-shaped passes + serial emit",
            "ms_per_step": t * 1e3, "visible": [int(c) for c in r["counts"]],
            "one_thread": {"value": n1 / one[1], "unit": UNIT, "sample": f"first {n1} entities, reference-shaped, 1 thread"},
            "fused": {"value": n / tf, "unit": UNIT, "ms_per_step": tf * 1e3, "steps": fused_steps,
                      "note": "oracle_frame_fused: one pass per entity, per-thread counts + prefix, parallel emit; same results"}}


def gpu_arm(args, wl):
    import torch
    import torch.distributed as dist

    from renderer_b200 import CullContext, scene
    from renderer_b200.shard import RecordMerger, ShardMerger, _wrap_device

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run --nproc-per-node {args.gpus}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: librcull has no CPU path (use --impl reference for the CPU oracle)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    n, n_views = wl["n"], wl["n_views"]
    lo = rank * n  # weak scaling: every rank owns n entities of an (n * world)-entity scene
    meshes = scene.synth_mesh_table()
    views = scene.scene_views(n_views)
    tri = bool(wl.get("tri")) and world == 1
    tri_cap = 0
    if tri:  # size the culled index buffer from the scene itself (the reference fixes it at 20 M triples)
        with CullContext(n, max_views=n_views, max_meshes=len(meshes), device=local_rank) as probe:
            probe.set_mesh_table(meshes)
            probe.synth_entities(scene.SEED, lo, n, wl["half_box"], len(meshes))
            probe.frame(views, scene.CAMERA_POS, draw_index_base=lo)
            tri_cap = int(probe.counts()[1][0]) + 3
    # N > 1, record merge: the frame kernel sends its survivors itself; then frame, expand, frame, expand on one stream
    # (--pipeline queues the expand of frame i behind frame i + 1; --expand-sms splits the chip between the two kernels)
    beside = world > 1 and args.merge == "p2p" and args.expand_sms > 0
    # records travel by copy engine (rcull_merge_send) or by SM stores from the frame kernel; auto: copy engines beyond 4 GPUs, where
    # the stores of 7 peers' worth of records slow the frame kernel down more than the copies cost (N = 8: 0.68 against 0.79 ms
    # per step; N = 2: 0.49 against 0.47)
    ce = world > 1 and args.merge == "p2p" and (args.records == "ce" or (args.records == "auto" and world > 4))
    pipelined = world > 1 and args.merge == "p2p" and (args.pipeline or ce) and not beside
    ctx = CullContext(n, max_views=n_views, max_meshes=len(meshes), device=local_rank, capacity_tri_indices=tri_cap)
    ctx.set_mesh_table(meshes)
    pv = None
    if tri:
        ctx.set_geometry(*scene.synth_geometry(meshes, coherent=bool(wl.get("coherent"))))
        pv = scene.camera_pv()
    stream = torch.cuda.Stream(device=dev, priority=-1)
    sh = stream.cuda_stream
    ctx.synth_entities(scene.SEED, lo, n, wl["half_box"], len(meshes), stream=sh)
    stream.synchronize()

    merger = None
    cap = cap_r = 0
    own_counts = own_totals = None
    if world > 1:
        ctx.frame(views, scene.CAMERA_POS, draw_index_base=lo, stream=sh)  # sizes the merged lists; also this rank's own lists
        stream.synchronize()
        own_counts, own_totals = ctx.counts()
        c0 = torch.tensor(own_counts[:n_views].astype(np.int64), device=dev)
        c_sum, c_max = c0.clone(), c0.clone()
        dist.all_reduce(c_sum)
        dist.all_reduce(c_max, op=dist.ReduceOp.MAX)
        cap = int(c_sum.max().item() * 1.05) + 1024
        cap_r = int(c_max.max().item() * 1.05) + 1024
        # transport: windows in torch symmetric memory with an NVLS multicast mapping when the fabric offers one (a record
        # then leaves the GPU once; measured equal to or better than one peer store per rank at N = 2 and 4), else CUDA IPC
        merger = RecordMerger(ctx, cap, cap_r, dev, symmetric=args.transport != "ipc") if args.merge == "p2p" else ShardMerger(n_views, cap, dev)
        if beside:
            merger.set_overlap(args.expand_sms)
        if ce:
            merger.set_transport(True)
            if args.send_bound:  # copy what a frame of this scene produces (+ 5 %) view by view instead of the rank's whole slot in one piece
                merger.set_send_bound([int(c * 1.05) + 1024 for c in own_counts[:n_views]])  # (measured slower: a peer copy costs ~13 us before it moves a byte)
    copy_stream = torch.cuda.Stream(device=dev) if ce else None
    # expand beside the next frame: the chip is split -- the frame kernel's grid leaves expand_sms SMs, the expand's CTAs
    # (one whole SM each) run there, on a second stream
    expand_stream = torch.cuda.Stream(device=dev, priority=0) if beside else None
    frame_ev = [torch.cuda.Event() for _ in range(2)]

    flush = torch.zeros((2, 256 * 1024 * 1024), dtype=torch.uint8, device=dev) if n * 36 < 256e6 else None
    record = world > 1 and args.merge == "p2p"
    pending = [0]  # merged frames issued and not expanded yet
    unsent = [None]  # copy-engine transport, send queued behind the expand: the timing event of the frame that still has to be sent

    def step(ev0=None, sent_ev=None):
        if ev0 is not None:
            ev0.record(stream)
        ctx.frame(views, scene.CAMERA_POS, draw_index_base=lo, stream=sh)
        if ce and args.send_early:  # peer copies right behind the frame (the library orders the streams): they then also run beside the expand of the frame before
            merger.send(copy_stream.cuda_stream)
            if sent_ev is not None:
                sent_ev.record(copy_stream)
        unsent[0] = sent_ev if (ce and not args.send_early) else None
        pending[0] += 1

    def send_behind_expand():
        # the copies of frame i start when the expand of frame i - 1 is through, so that they share the memory system with the
        # next frame only (the expand is HBM-bound and lost 60 % beside 280 MB of copy traffic; the frame is latency-bound)
        e = frame_ev[pending[0] & 1]
        e.record(stream)
        copy_stream.wait_event(e)
        merger.send(copy_stream.cuda_stream)
        if unsent[0] is not None:
            unsent[0].record(copy_stream)
        unsent[0] = None

    def do_merge(final=False):
        if beside:
            e = frame_ev[pending[0] & 1]
            e.record(stream)
            expand_stream.wait_event(e)
            merger.expand(expand_stream.cuda_stream)
            pending[0] = 0
            if final:
                stream.wait_stream(expand_stream)  # the timed region ends when the last expand has
        elif record:
            # pipelined: frame i + 1 is already queued when frame i is expanded; the last step expands both
            late = ce and not args.send_early
            while pending[0] > (0 if (final or not pipelined) else 1):
                if late and pending[0] == 1:
                    send_behind_expand()  # the newest frame has to be sent before it can be expanded
                    late = False
                merger.expand(sh)
                pending[0] -= 1
            if late:
                send_behind_expand()
        else:
            pending[0] = 0
            merger.merge_from_context(ctx, stream=sh)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(stream):
        for i in range(args.warmup):
            step()
            if merger:
                do_merge(final=i == args.warmup - 1)
            if tri:
                ctx.tri_cull(pv, draw_index_base=lo, stream=sh)
            if flush is not None:
                flush_l2(flush)
        sync_all()
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
        sent_evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)] if ce else None
        launches0 = ctx.launch_count
        with ClockSampler(local_rank) as clocks:
            for i in range(args.steps):
                step(ev[i][0], sent_evs[i] if ce else None)
                ev[i][1].record(stream)
                if merger:
                    do_merge(final=i == args.steps - 1)  # the timed region holds K frames AND K expands
                if tri:
                    ctx.tri_cull(pv, draw_index_base=lo, stream=sh)
                ev[i][2].record(stream)
                if flush is not None:
                    flush_l2(flush)
            sync_all()
        launches = ctx.launch_count - launches0
    t_frame = sum(e[0].elapsed_time(e[1]) for e in ev) / 1e3  # seconds, all steps
    t_step = (ev[0][0].elapsed_time(ev[-1][2]) if (pipelined or beside) else sum(e[0].elapsed_time(e[2]) for e in ev)) / 1e3
    counts, totals = ctx.counts()  # this rank's own counts (the frame kernel writes them in every mode)
    counts = counts[:n_views].astype(np.int64)

    # ---- end to end: host (pinned) buffers in, counts out, through rcull_frame_host ----
    hp = torch.empty((n, 4), dtype=torch.float32).pin_memory()
    hr = torch.empty((n, 4), dtype=torch.float32).pin_memory()
    hm = torch.empty((n,), dtype=torch.int32).pin_memory()
    p_ps, p_rot, p_mesh = ctx.entity_ptrs()
    hp.copy_(_wrap_device(torch, p_ps, n * 4, torch.float32, dev).view(n, 4))
    hr.copy_(_wrap_device(torch, p_rot, n * 4, torch.float32, dev).view(n, 4))
    hm.copy_(_wrap_device(torch, p_mesh, n, torch.int32, dev))
    torch.cuda.synchronize()
    a_ps, a_rot, a_mesh = hp.numpy(), hr.numpy(), hm.numpy().view(np.uint32)
    e2e_steps = max(1, min(args.steps, 10))

    def e2e_step():
        c, _ = ctx.frame_host(a_ps, a_rot, a_mesh, views, scene.CAMERA_POS, draw_index_base=lo)
        if merger:
            if record:
                if ce:
                    merger.send(ctx.stream)
                merger.expand(ctx.stream)  # frame_host has completed on the context's stream: the records are out
            else:
                merger.merge_from_context(ctx, stream=sh)
            torch.cuda.synchronize()
        return c

    for _ in range(2):
        c_e2e = e2e_step()
    sync_all()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        c_e2e = e2e_step()
    sync_all()
    t_e2e = time.perf_counter() - t0
    assert np.array_equal(np.asarray(c_e2e, np.int64), counts), "end-to-end path disagrees with the resident path"

    # ---- end to end WITH the results back: the same call + the draw lists of every view into pinned host memory ----
    e2e_lists = None
    if world == 1:
        cap_l = int(counts.max()) + 1024
        h_cmds = [torch.empty((cap_l, 5), dtype=torch.int32).pin_memory() for _ in range(n_views)]
        h_vis = [torch.empty((cap_l,), dtype=torch.int32).pin_memory() for _ in range(n_views)]
        pc, pvis = [t.data_ptr() for t in h_cmds], [t.data_ptr() for t in h_vis]
        for _ in range(2):
            c_l, _ = ctx.frame_host_lists(a_ps, a_rot, a_mesh, views, scene.CAMERA_POS, pc, pvis, cap_l, draw_index_base=lo)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            c_l, _ = ctx.frame_host_lists(a_ps, a_rot, a_mesh, views, scene.CAMERA_POS, pc, pvis, cap_l, draw_index_base=lo)
        t_l = time.perf_counter() - t0
        assert np.array_equal(np.asarray(c_l, np.int64), counts)
        k_chk = n_views - 1  # spot check of what came back: ascending draw indices, firstInstance == visible index
        v_chk = h_vis[k_chk].numpy()[: int(counts[k_chk])].view(np.uint32)
        assert np.all(np.diff(v_chk.astype(np.int64)) > 0) and np.array_equal(h_cmds[k_chk].numpy()[: int(counts[k_chk]), 4].view(np.uint32), v_chk)
        e2e_lists = {"value": n * e2e_steps / t_l, "unit": UNIT, "h2d_bytes_per_step": n * 36, "d2h_bytes_per_step": 96 + 24 * int(counts.sum()),
                     "steps": e2e_steps, "ms_per_step": t_l / e2e_steps * 1e3,
                     "api": "rcull_frame_host_lists (pinned host columns in; counts, index totals and the draw lists of every view out)"}

    # ---- N > 1: what the merge delivered.  (a) every rank's OWN slice of every merged list, bit for bit, against the
    # oracle run on that rank's shard (the shards with draw_index_base >= 16 Mi included), plus the
    # merged counts / totals against the sum of the ranks' oracle counts; (b) the merged lists against an NCCL
    # allgatherv of per-rank lists produced by a second, merge-free context. ----
    merge_verified = oracle_verified = None
    if record and not args.no_verify:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as O  # checker

        pos, quat, scale, mesh_h = O.synth_entities(scene.SEED, lo, n, wl["half_box"], len(meshes))
        ref = O.frame(pos, quat, scale, mesh_h, meshes, views, scene.CAMERA_POS, draw_index_base=lo, n_threads=max(1, (os.cpu_count() or 8) // world * 2),
                      want=("cmds", "visible"), fused=True)
        del pos, quat, scale, mesh_h
        ok = bool(np.array_equal(ref["counts"].astype(np.int64), counts))
        meta = torch.tensor(np.concatenate([ref["counts"].astype(np.int64), ref["totals"].astype(np.int64)]), dtype=torch.int64, device=dev)
        meta_all = torch.zeros(world * 2 * n_views, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(meta_all, meta)
        meta_all = meta_all.view(world, 2, n_views).cpu().numpy()
        torch.cuda.synchronize()
        for k in range(n_views):
            cp, vp, cntp, totp = merger.lists(k)
            cnt_all = _wrap_device(torch, cntp, n_views, torch.int32, dev).cpu().numpy().view(np.uint32)
            tot_all = _wrap_device(torch, totp, n_views, torch.int64, dev).cpu().numpy().view(np.uint64)
            ok = ok and int(cnt_all[k]) == int(meta_all[:, 0, k].sum()) and int(tot_all[k]) == int(meta_all[:, 1, k].sum())
            first = int(meta_all[:rank, 0, k].sum())
            idx_prefix = int(meta_all[:rank, 1, k].sum()) & 0xFFFFFFFF
            c = int(ref["counts"][k])
            got_v = _wrap_device(torch, vp + 4 * first, max(c, 1), torch.int32, dev)[:c].cpu().numpy().view(np.uint32)
            got_c = _wrap_device(torch, cp + 20 * first, max(c, 1) * 5, torch.int32, dev)[: c * 5].cpu().numpy().view(np.uint32).reshape(c, 5)
            want_c = ref["cmds"][k].view(np.uint32).reshape(c, 5).copy()
            want_c[:, 2] += np.uint32(idx_prefix)  # the rank prefix of the merged running sum (u32 wrap like the reference's)
            ok = ok and bool(np.array_equal(got_v, ref["visible"][k])) and bool(np.array_equal(got_c, want_c))
        flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        oracle_verified = bool(flag.item())
        assert oracle_verified, "merged draw lists differ from the oracle run on the ranks' shards"
        del ref
        # (b) transport cross-check: NCCL allgatherv of a merge-free context's lists of the same shard
        with CullContext(n, max_views=n_views, max_meshes=len(meshes), device=local_rank) as plain:
            plain.set_mesh_table(meshes)
            plain.synth_entities(scene.SEED, lo, n, wl["half_box"], len(meshes), stream=sh)
            plain.frame(views, scene.CAMERA_POS, draw_index_base=lo, stream=sh)
            sync_all()
            ref_m = ShardMerger(n_views, cap, dev)
            with torch.cuda.stream(stream):
                ref_m.merge_from_context(plain, stream=sh)
            sync_all()
            ok = True
            for k in range(n_views):
                cp, vp, _, _ = merger.lists(k)
                cnt = int(ref_m.plan.counts[k])
                ok = ok and torch.equal(_wrap_device(torch, cp, max(cnt, 1) * 5, torch.int32, dev)[: cnt * 5], ref_m.merged_cmds[k][: cnt * 5])
                ok = ok and torch.equal(_wrap_device(torch, vp, max(cnt, 1), torch.int32, dev)[:cnt], ref_m.merged_visible[k][:cnt])
            flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            merge_verified = bool(flag.item())
            assert merge_verified, "merged draw lists of the record merge differ from the NCCL allgatherv of the ranks' lists"
            del ref_m

    # ---- reduce over ranks: time = max, work = sum ----
    times = torch.tensor([t_step, t_frame, t_e2e], dtype=torch.float64, device=dev)
    vis = torch.tensor(counts, dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
        dist.all_reduce(vis)
    t_step, t_frame, t_e2e = (float(x) for x in times.tolist())
    vis = vis.cpu().numpy()
    n_total = n * world

    if rank == 0:
        peak, peak_kind = load_peaks()
        v_frac = counts / float(n)  # this rank's visible fractions (rank 0's launch is the one timed)
        if record:  # 8-byte records instead of 24-byte list entries, and every rank's records land in this GPU's window during the frame
            bytes_per_entity = 36.0 + 64.0 + 8.0 * float(vis.sum()) / n
        else:
            bytes_per_entity = 36.0 + 64.0 + 24.0 * float(v_frac.sum())
        merge_report = None
        if world > 1:
            t_expand = sum(e[1].elapsed_time(e[2]) for e in ev) / args.steps / 1e3
            inbound = 8.0 * (float(vis.sum()) - float(counts.sum())) if record else 24.0 * (float(vis.sum()) - float(counts.sum()))
            merge_report = {
                "records": ("copy engines" if ce else "stores from the frame kernel") if record else None,
                # copy-engine transport: from the end of a frame to the last peer copy of its records (rank 0; they run under the next frame)
                "send_ms_per_step": (sum(e[1].elapsed_time(s_) for e, s_ in zip(ev, sent_evs)) / args.steps) if ce else None,
                "expand_ms_per_step" if record else "ms_per_step": t_expand * 1e3,
                "step_minus_frame_ms": (t_step - t_frame) / args.steps * 1e3,
                "expand_queued_behind_next_frame": pipelined, "expand_beside_next_frame": beside,
                "record_bytes_per_survivor": 8 if record else 24,
                "nvlink_inbound_bytes_per_gpu_per_frame": int(inbound),
                "nvlink_outbound_bytes_per_gpu_per_frame": int((8.0 if record else 24.0) * float(counts.sum()) * (1 if "multicast" in getattr(merger, "transport", "") else world - 1)),
                # records travel while the frame kernel runs: the link is busy for the frame's duration
                "nvlink_inbound_gbs_per_gpu": inbound / (t_frame / args.steps) / 1e9 if record else inbound / max(t_expand, 1e-9) / 1e9,
                "nvlink_peak_gbs": 770.0,
                "nvlink_frac": (inbound / (t_frame / args.steps) / 1e9 if record else inbound / max(t_expand, 1e-9) / 1e9) / 770.0,
                "expand_hbm_gbs": (32.0 * float(vis.sum()) / max(t_expand, 1e-9) / 1e9) if record else None,
                "transport": getattr(merger, "transport", "nccl"),
                "verified_against_oracle_per_rank": oracle_verified, "verified_against_nccl_allgatherv": merge_verified,
                "note": "peak = measured peer copy per direction; inbound = the other ranks' records landing in this GPU's window during the frame; expand = 8 B read + 24 B written per merged survivor, local"}
        t_launch = t_frame / args.steps
        achieved = n * bytes_per_entity / t_launch / 1e9
        line = {
            "metric": METRIC, "value": n_total * args.steps / t_step, "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_step / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {
                "workload": args.workload, "entities_per_gpu": n, "entities_total": n_total, "n_views": n_views,
                "half_box": wl["half_box"], "visible_fraction_per_view": [round(float(x), 6) for x in vis / float(n_total)],
                "sharding": "contiguous range per rank" if world > 1 else "none",
                "merge": (((("records: the frame kernel leaves an 8-byte {draw_index, slot} record per survivor in its own window and the copy engines move each rank's region to every peer while the next frame runs (rcull_merge_send: peer copies, no exchange kernel, no collective)") if ce else
                            "records: the frame kernel stores an 8-byte {draw_index, slot} record per survivor into every rank's window over NVLink as it emits (no exchange kernel, no collective)") + "; merge_expand rebuilds the merged commands on every GPU"
                           + (f"; the expand of frame i runs beside frame i+1 on {args.expand_sms} SMs of its own (second stream)" if beside else
                              "; the expand of frame i is queued behind frame i+1" if pipelined else "; frame, expand, frame, expand"))
                          if args.merge == "p2p" else "NCCL all_gather(counts) + in-place broadcast allgatherv") if world > 1 else "none",
                "l2": L2_NOTE if flush is not None else "inputs (36 B x n) exceed L2",
            },
            "e2e": {"value": n_total * e2e_steps / t_e2e, "unit": UNIT, "h2d_bytes_per_step": n * 36,
                    "d2h_bytes_per_step": 96, "steps": e2e_steps, "ms_per_step": t_e2e / e2e_steps * 1e3,
                    "api": "rcull_frame_host (pinned host columns in, counts + index totals out)",
                    "result_resident": True,
                    "note": "model matrices and draw lists stay on the device, as the reference's consumers (indirect draws) read them there; e2e_lists also brings the lists back"},
            "e2e_lists": e2e_lists,
            "gpu_launches": int(launches),
            "merge": None if world == 1 else merge_report,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "frac_of_nominal_8000": achieved / 8000.0, "traffic": None, "peak_kind": peak_kind, "kernel": f"transform_cull_compact<{n_views}>",
                         "bytes_per_entity": bytes_per_entity, "ms_per_launch": t_launch * 1e3},
            "clocks": clocks.summary(),
        }
        counters = load_counters()
        if tri:
            n_cmd, kept = ctx.tri_counts()
            line["tri_stage"] = tri_stage_report(wl, (t_step - t_frame) / args.steps, int(totals[0]) // 3, kept, int(counts[0]), n_cmd,
                                                 line["clocks"].get("sm_mhz"), counters)
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(wl, steps=3, warmup=1, threads=os.cpu_count() or 1)
        if world == 1 and not args.no_also:
            line["also"] = [resident_timing(w, args.steps, local_rank) for w in ("16m_1view", "16m_5view", "1m_1view") if w != args.workload]
            line["also"].append(resident_tri("1m_tri", args.steps, local_rank, counters))
            line["also"].append(resident_per_mesh("sponza_4m", args.steps, local_rank))
            line["also"] += resident_small_kernels(args.steps, local_rank)
        tr = (counters.get("dram_bytes_per_launch") or {}).get(args.workload)
        if tr:
            line["roofline"]["traffic"] = tr["bytes"]
            line["roofline"]["traffic_source"] = tr["source"] + " (ncu --set full capture; a cross-reference, not measured in this run)"
        print(json.dumps(line), flush=True)
    if merger is not None and hasattr(merger, "destroy"):
        merger.destroy()
    ctx.destroy()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def resident_timing(name, steps, device):
    """Kernel-resident timing of another workload of the table (no e2e / CPU legs): the same CUDA-event
    protocol as the main line, reported under "also" so one run shows the 1-view and 5-view rooflines together."""
    import torch

    from renderer_b200 import CullContext, scene

    wl = WORKLOADS[name]
    n, n_views = wl["n"], wl["n_views"]
    meshes = scene.synth_mesh_table()
    views = scene.scene_views(n_views)
    dev = torch.device("cuda", device)
    stream = torch.cuda.Stream(device=dev)
    flush = torch.zeros((2, 256 * 1024 * 1024), dtype=torch.uint8, device=dev) if n * 36 < 256e6 else None
    with CullContext(n, max_views=n_views, max_meshes=len(meshes), device=device) as ctx:
        ctx.set_mesh_table(meshes)
        ctx.synth_entities(scene.SEED, 0, n, wl["half_box"], len(meshes), stream=stream.cuda_stream)
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(steps)]
        with torch.cuda.stream(stream):
            for _ in range(5):
                ctx.frame(views, scene.CAMERA_POS, stream=stream.cuda_stream)
                if flush is not None:
                    flush_l2(flush)
            torch.cuda.synchronize()
            for i in range(steps):
                ev[i][0].record(stream)
                ctx.frame(views, scene.CAMERA_POS, stream=stream.cuda_stream)
                ev[i][1].record(stream)
                if flush is not None:
                    flush_l2(flush)
            torch.cuda.synchronize()
        t = sum(a.elapsed_time(b) for a, b in ev) / steps / 1e3
        counts, _ = ctx.counts()
    peak, _ = load_peaks()
    bpe = 36.0 + 64.0 + 24.0 * float(counts[:n_views].sum()) / n
    return {"workload": name, "entities": n, "n_views": n_views, "ms_per_step": t * 1e3, "value": n / t, "unit": UNIT,
            "bytes_per_entity": bpe, "achieved_gbs": n * bpe / t / 1e9, "roofline_frac": n * bpe / t / 1e9 / peak,
            "l2": L2_NOTE if flush is not None else "inputs (36 B x n) exceed L2"}


def tri_stage_report(wl, t_tri, tris, kept, commands_in, commands_out, sm_mhz, counters):
    """Stage 2 (tri_cull + compact_draw_stream).  The stage is bound by FP32 ISSUE, not by HBM (ncu: 370 GB/s of DRAM
    traffic, 75 % of the issue slots busy): its roofline is the chip's issue rate -- one warp instruction per clock and
    SM sub-partition -- and `achieved` is the executed warp instructions (per-triangle count from the committed ncu
    capture, labelled) over the time measured here."""
    peak, _ = load_peaks()
    # 12 B index triple read per triangle, 12 B written per kept one, 20 + 4 B command in, 20 B out
    alg = 12.0 * tris + 4.0 * kept + 24.0 * commands_in + 20.0 * commands_out
    out = {"kernels": "tri_cull + compact_draw_stream", "ms_per_step": t_tri * 1e3, "triangles_in": tris, "triangles_kept": kept // 3,
           "commands_in": commands_in, "commands_out": commands_out, "gtri_per_s": tris / t_tri / 1e9,
           "algorithmic_gbs": alg / t_tri / 1e9, "hbm_frac": alg / t_tri / 1e9 / peak,
           "meshes": "grid sheets, row-major cell order (vertex reuse like an indexed glTF mesh)" if wl.get("coherent") else "random index triples (no vertex reuse)",
           "note": "e2e and cpu_baseline cover the frame stage only"}
    ipt = (counters.get("warp_inst_per_unit") or {}).get("tri_cull")
    if ipt and sm_mhz:
        issue_peak = SM_ISSUE_SLOTS * sm_mhz * 1e6
        ach = ipt["value"] * tris / t_tri
        out["roofline"] = {"bound": "fp32_issue", "achieved": ach / 1e9, "peak": issue_peak / 1e9, "unit": "G warp-inst/s", "frac": ach / issue_peak,
                           "warp_inst_per_triangle": ipt["value"], "warp_inst_source": ipt["source"],
                           "peak_kind": f"148 SMs x 4 sub-partitions x {sm_mhz:.0f} MHz (SM clock sampled in this run)"}
    return out


def resident_tri(name, steps, device, counters):
    """BASELINE configs[1] through both stages, resident: frame kernel, then tri_cull + compact_draw_stream."""
    import torch

    from renderer_b200 import CullContext, scene

    wl = WORKLOADS[name]
    n = wl["n"]
    meshes = scene.synth_mesh_table()
    views, pv = scene.scene_views(1), scene.camera_pv()
    dev = torch.device("cuda", device)
    stream = torch.cuda.Stream(device=dev)
    sh = stream.cuda_stream
    flush = torch.zeros((2, 256 * 1024 * 1024), dtype=torch.uint8, device=dev)
    with CullContext(n, max_views=1, max_meshes=len(meshes), device=device) as probe:
        probe.set_mesh_table(meshes)
        probe.synth_entities(scene.SEED, 0, n, wl["half_box"], len(meshes))
        probe.frame(views, scene.CAMERA_POS)
        tri_cap = int(probe.counts()[1][0]) + 3
    with CullContext(n, max_views=1, max_meshes=len(meshes), device=device, capacity_tri_indices=tri_cap) as ctx:
        ctx.set_mesh_table(meshes)
        ctx.set_geometry(*scene.synth_geometry(meshes, coherent=bool(wl.get("coherent"))))
        ctx.synth_entities(scene.SEED, 0, n, wl["half_box"], len(meshes), stream=sh)
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps)]
        with ClockSampler(device) as clocks, torch.cuda.stream(stream):
            for i in range(5 + steps):
                j = i - 5
                if j >= 0:
                    ev[j][0].record(stream)
                ctx.frame(views, scene.CAMERA_POS, stream=sh)
                if j >= 0:
                    ev[j][1].record(stream)
                ctx.tri_cull(pv, stream=sh)
                if j >= 0:
                    ev[j][2].record(stream)
                flush_l2(flush)
            torch.cuda.synchronize()
        t_frame = sum(e[0].elapsed_time(e[1]) for e in ev) / steps / 1e3
        t_tri = sum(e[1].elapsed_time(e[2]) for e in ev) / steps / 1e3
        counts, totals = ctx.counts()
        n_cmd, kept = ctx.tri_counts()
    rep = tri_stage_report(wl, t_tri, int(totals[0]) // 3, kept, int(counts[0]), n_cmd, clocks.summary().get("sm_mhz"), counters)
    rep.update({"workload": name, "entities": n, "n_views": 1, "frame_ms": t_frame * 1e3, "l2": L2_NOTE})
    return rep


def resident_per_mesh(name, steps, device):
    """BASELINE configs[3]: the Sponza mesh table instanced to 4 M entities, per-mesh indirect command compaction
    (output_mode PER_MESH_INSTANCED): frame kernel + the bucketing passes, resident."""
    import torch

    from renderer_b200 import PER_MESH_INSTANCED, CullContext, scene

    n, n_views = 4_000_000, 1
    table, node_scale = scene.sponza_mesh_table()
    ps, rot, mesh = scene.sponza_instances(n, len(table), node_scale)
    views = scene.scene_views(n_views)
    dev = torch.device("cuda", device)
    stream = torch.cuda.Stream(device=dev)
    sh = stream.cuda_stream
    flush = torch.zeros((2, 256 * 1024 * 1024), dtype=torch.uint8, device=dev)
    with CullContext(n, max_views=n_views, max_meshes=len(table), device=device, output_mode=PER_MESH_INSTANCED) as ctx:
        ctx.set_mesh_table(table)
        ctx.upload_entities(ps, rot, mesh)
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(steps)]
        l0 = 0
        with torch.cuda.stream(stream):
            for i in range(5 + steps):
                j = i - 5
                if j == 0:
                    l0 = ctx.launch_count
                if j >= 0:
                    ev[j][0].record(stream)
                ctx.frame(views, scene.CAMERA_POS, stream=sh)
                if j >= 0:
                    ev[j][1].record(stream)
                flush_l2(flush)
            torch.cuda.synchronize()
        t = sum(a.elapsed_time(b) for a, b in ev) / steps / 1e3
        launches = (ctx.launch_count - l0) / steps
        counts, totals = ctx.counts()
    peak, _ = load_peaks()
    # 36 B in + 64 B model out per entity, 4 B per survivor (bucketed visible index), 20 B per non-empty slot
    alg = n * 100.0 + 4.0 * float(totals[0]) + 20.0 * float(counts[0])
    return {"workload": name, "entities": n, "meshes": len(table), "n_views": n_views, "mode": "per-mesh instanced (BASELINE configs[3])",
            "ms_per_step": t * 1e3, "value": n / t, "unit": UNIT, "launches_per_frame": launches, "commands": int(counts[0]), "instances": int(totals[0]),
            "bytes_per_entity": alg / n, "achieved_gbs": alg / t / 1e9, "roofline_frac": alg / t / 1e9 / peak, "bound": "hbm", "l2": L2_NOTE}


def resident_small_kernels(steps, device):
    """The two small per-entity kernels beside the frame kernel: model_only (frozen-culling frames, 32 B in + 64 B out
    per entity) and as_instances (acceleration-structure instance records, 68 B in + 64 B out), 16 Mi entities."""
    import torch

    from renderer_b200 import CullContext, scene

    n = 16 * 1024 * 1024
    meshes = scene.synth_mesh_table()
    dev = torch.device("cuda", device)
    stream = torch.cuda.Stream(device=dev)
    sh = stream.cuda_stream
    peak, _ = load_peaks()
    out = []
    with CullContext(n, max_views=1, max_meshes=len(meshes), device=device) as ctx:
        ctx.set_mesh_table(meshes)
        ctx.set_blas_table(np.arange(1, len(meshes) + 1, dtype=np.uint64) << np.uint64(12))
        ctx.synth_entities(scene.SEED, 0, n, 600.0, len(meshes), stream=sh)
        ctx.frame(scene.scene_views(1), scene.CAMERA_POS, stream=sh)
        for name, fn, bpe in (("model_only_16m", lambda: ctx.frame_frozen(n, stream=sh), 96.0), ("as_instances_16m", lambda: ctx.as_instances(n, stream=sh), 132.0)):
            ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(steps)]
            with torch.cuda.stream(stream):
                for i in range(3 + steps):
                    j = i - 3
                    if j >= 0:
                        ev[j][0].record(stream)
                    fn()
                    if j >= 0:
                        ev[j][1].record(stream)
                torch.cuda.synchronize()
            t = sum(a.elapsed_time(b) for a, b in ev) / steps / 1e3
            out.append({"workload": name, "entities": n, "ms_per_step": t * 1e3, "value": n / t, "unit": UNIT, "bytes_per_entity": bpe,
                        "achieved_gbs": n * bpe / t / 1e9, "roofline_frac": n * bpe / t / 1e9 / peak, "bound": "hbm",
                        "l2": "inputs exceed L2"})
    return out


def reference_arm(args, wl):
    """The reference's CPU path for this stage on the box's host cores: the oracle (the reference itself is Rust
    nightly + Vulkan and cannot be built here), all host threads, on the SAME workload as the GPU arm -- every entity,
    every view; the bound is the step count.  Maps no product library."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n = wl["n"]
    b = cpu_baseline(wl, steps=args.steps, warmup=args.warmup, threads=threads)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": b["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": b["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "entities_per_gpu": n, "entities_total": n * args.gpus, "n_views": wl["n_views"], "half_box": wl["half_box"],
                   "visible_fraction_per_view": [round(c / float(n), 6) for c in b["visible"]],
                   "sharding": "contiguous range per rank" if args.gpus > 1 else "none",
                   "entities_per_step": n,
                   "note": "CPU oracle (C restatement of the reference's bevy systems; Rust toolchain absent), all host threads"
                           + ("; one rank's 16 Mi-entity shard per step (the bounded sample of the N-GPU workload)" if args.gpus > 1 else "")},
        "cpu_baseline": {k: b[k] for k in ("value", "unit", "cores", "kind", "sample", "fused", "one_thread")},
        "e2e": {"value": b["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--merge", default="p2p", choices=["p2p", "nccl"], help="N > 1: peer-store merge kernel or NCCL broadcasts")
    ap.add_argument("--pipeline", action="store_true", help="N > 1: queue the expand of frame i behind frame i + 1 instead of frame, expand, frame, expand (measured slower: with two record regions the next frame then waits for the peers' expand of the frame before)")
    ap.add_argument("--expand-sms", type=int, default=0, help="N > 1: > 0 splits the chip -- the frame kernel leaves that many SMs to the expand of the previous frame, which runs beside it on a second stream; 0: the expand runs between frames")
    ap.add_argument("--records", default="auto", choices=["auto", "push", "ce"], help="N > 1: push = the frame kernel stores its records into every rank's window; ce = it keeps them local and the copy engines move them under the next frame (implies --pipeline); auto = ce beyond 4 GPUs")
    ap.add_argument("--send-early", action="store_true", help="--records ce: queue a frame's peer copies right behind the frame instead of behind the expand of the frame before")
    ap.add_argument("--send-bound", action="store_true", help="--records ce: per-view copies of an estimate of the survivors (rcull_merge_set_send_bound) instead of one copy of the rank's whole slot")
    ap.add_argument("--transport", default="auto", choices=["auto", "ipc"], help="N > 1: auto = windows in torch symmetric memory, records through the NVLS multicast mapping when there is one (falls back to CUDA IPC peer stores); ipc = CUDA IPC peer stores")
    ap.add_argument("--no-verify", action="store_true", help="N > 1: skip the oracle / NCCL verification of the merged lists")
    ap.add_argument("--workload", default="16m_5view", choices=sorted(WORKLOADS))
    ap.add_argument("--no-also", action="store_true", help="N = 1: skip the resident timings of the other entity-stage workloads")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiler runs)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    args.steps = max(args.steps, 1)
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        reference_arm(args, wl)
    else:
        gpu_arm(args, wl)


if __name__ == "__main__":
    main()
