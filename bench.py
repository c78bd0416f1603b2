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
CPU_SAMPLE = 2 * 1024 * 1024  # entities per CPU-baseline step (bounded sample of the workload)


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


def cpu_baseline(wl, steps: int, warmup: int, threads: int):
    """The oracle (kind "port") on a bounded sample of the workload, all host threads."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O  # CPU-baseline leg: the one place bench.py may execute oracle/
    from renderer_b200 import scene

    n = min(CPU_SAMPLE, wl["n"])
    ps, rot, mesh = scene.synth_entities(scene.SEED, 0, n, wl["half_box"])
    pos, quat, scale = scene.to_components(ps, rot)
    meshes = scene.synth_mesh_table()
    views = scene.scene_views(wl["n_views"])
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
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        O.frame(pos, quat, scale, mesh, meshes, views, scene.CAMERA_POS, n_threads=threads, want=("cmds", "visible"), fused=True)
        if i >= warmup:
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
            "sample": f"first {n} entities of the workload, {wl['n_views']} views, {steps} steps, reference-shaped passes + serial emit",
            "ms_per_step": t * 1e3, "visible": [int(c) for c in r["counts"]],
            "one_thread": {"value": n1 / one[1], "unit": UNIT, "sample": f"first {n1} entities, reference-shaped, 1 thread"},
            "fused": {"value": n / tf, "unit": UNIT, "ms_per_step": tf * 1e3,
                      "note": "oracle_frame_fused: one pass per entity, per-thread counts + prefix, parallel emit; same results"}}


def gpu_arm(args, wl):
    import torch
    import torch.distributed as dist

    from renderer_b200 import CullContext, scene
    from renderer_b200.shard import PeerMerger, ShardMerger, _wrap_device

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
    # N > 1 with the peer-store merge: two output sets, so the merge of frame i (its own stream) overlaps frame i + 1
    pipelined = world > 1 and args.merge == "p2p" and not args.no_overlap
    ctx = CullContext(n, max_views=n_views, max_meshes=len(meshes), device=local_rank, capacity_tri_indices=tri_cap,
                      n_buffers=2 if pipelined else 1)
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
    if world > 1:
        ctx.frame(views, scene.CAMERA_POS, draw_index_base=lo, stream=sh)
        stream.synchronize()
        c0, _ = ctx.counts()
        tot = torch.tensor([int(c0[:n_views].max())], dtype=torch.int64, device=dev)
        dist.all_reduce(tot)
        cap = int(tot.item() * 1.05) + 1024
        merger = PeerMerger(ctx, n_views, cap, dev, symmetric=world >= 8) if args.merge == "p2p" else ShardMerger(n_views, cap, dev)

    flush = torch.zeros((2, 256 * 1024 * 1024), dtype=torch.uint8, device=dev) if n * 36 < 256e6 else None

    merge_stream = torch.cuda.Stream(device=dev, priority=0) if pipelined else None  # below the frame stream: frame CTAs are placed first
    frame_done = [torch.cuda.Event() for _ in range(2)]
    step_no = [0]

    def step(ev0=None):
        if not pipelined:
            if ev0 is not None:
                ev0.record(stream)
            ctx.frame(views, scene.CAMERA_POS, draw_index_base=lo, stream=sh)
            return
        b = step_no[0] & 1
        if step_no[0] >= 2:
            stream.wait_event(merger.ev_scatter_done[b])  # merge i-2 has read output set b
        if ev0 is not None:
            ev0.record(stream)
        ctx.frame(views, scene.CAMERA_POS, draw_index_base=lo, buffer_ix=b, stream=sh)
        frame_done[b].record(stream)

    def do_merge():
        if pipelined:
            b = step_no[0] & 1
            step_no[0] += 1
            merge_stream.wait_event(frame_done[b])
            with torch.cuda.stream(merge_stream):
                merger.exchange(b)
        elif args.merge == "p2p":
            merger.merge()
        else:
            merger.merge_from_context(ctx, stream=sh)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            step()
            if merger:
                do_merge()
            if tri:
                ctx.tri_cull(pv, draw_index_base=lo, stream=sh)
            if flush is not None:
                flush_l2(flush)
        sync_all()
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
        launches0 = ctx.launch_count
        with ClockSampler(local_rank) as clocks:
            for i in range(args.steps):
                step(ev[i][0])
                ev[i][1].record(stream)
                if merger:
                    do_merge()
                if tri:
                    ctx.tri_cull(pv, draw_index_base=lo, stream=sh)
                if pipelined and i == args.steps - 1:
                    stream.wait_stream(merge_stream)  # the timed region ends when the last merge has
                ev[i][2].record(stream)
                if flush is not None:
                    flush_l2(flush)
            sync_all()
        launches = ctx.launch_count - launches0
    t_frame = sum(e[0].elapsed_time(e[1]) for e in ev) / 1e3  # seconds, all steps
    t_step = (ev[0][0].elapsed_time(ev[-1][2]) if pipelined else sum(e[0].elapsed_time(e[2]) for e in ev)) / 1e3
    counts, totals = ctx.counts()
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
            if pipelined:
                merger.exchange(0)  # frame_host wrote output set 0 and has already completed
            elif args.merge == "p2p":
                merger.merge(0)
            else:
                do_merge()
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

    # ---- N > 1: the merged lists of the peer-store exchange against an NCCL allgatherv of the same frame ----
    merge_verified = None
    if world > 1 and args.merge == "p2p":
        with torch.cuda.stream(stream):
            ctx.frame(views, scene.CAMERA_POS, draw_index_base=lo, buffer_ix=0, stream=sh)
            if pipelined:
                merger.exchange(0)
            else:
                merger.merge(0)
        sync_all()
        ref = ShardMerger(n_views, cap, dev)
        with torch.cuda.stream(stream):
            ref.merge_from_context(ctx, stream=sh)
        sync_all()
        ok = True
        for k in range(n_views):
            cp, vp, _, _ = merger.lists(k)
            cnt = int(ref.plan.counts[k])
            ok = ok and torch.equal(_wrap_device(torch, cp, max(cnt, 1) * 5, torch.int32, dev)[: cnt * 5], ref.merged_cmds[k][: cnt * 5])
            ok = ok and torch.equal(_wrap_device(torch, vp, max(cnt, 1), torch.int32, dev)[:cnt], ref.merged_visible[k][:cnt])
        flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        merge_verified = bool(flag.item())
        assert merge_verified, "merged draw lists of the peer-store exchange differ from the NCCL allgatherv of the same frame"

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
        bytes_per_entity = 36.0 + 64.0 + 24.0 * float(v_frac.sum())
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
                "merge": ((("one fused exchange kernel per rank: counts to every peer's mailbox, rebased lists stored into every rank's allocation over NVLink, done tags; on its own stream, the exchange of frame i overlaps frame i+1 (two output sets)" if pipelined else "NCCL all_gather(meta) + one peer-store kernel per rank (NVLink P2P) + 1-word all_reduce")) if args.merge == "p2p" else "NCCL all_gather(counts) + in-place broadcast allgatherv") if world > 1 else "none",
                "l2": L2_NOTE if flush is not None else "inputs (36 B x n) exceed L2",
            },
            "e2e": {"value": n_total * e2e_steps / t_e2e, "unit": UNIT, "h2d_bytes_per_step": n * 36,
                    "d2h_bytes_per_step": 96, "steps": e2e_steps, "ms_per_step": t_e2e / e2e_steps * 1e3,
                    "api": "rcull_frame_host (pinned host columns in, counts + index totals out)"},
            "gpu_launches": int(launches),
            "merge": None if world == 1 else {
                # pipelined: what the step costs beyond the frame kernel (the exchange itself runs under the next frame)
                ("exposed_ms_per_step" if pipelined else "ms_per_step"): (t_step - t_frame) / args.steps * 1e3,
                "overlapped_with_next_frame": pipelined,
                "bytes_delivered_to_peers_per_gpu": int(24 * int(counts.sum()) * (world - 1)),
                "bytes_leaving_each_gpu": int(24 * int(counts.sum()) * (1 if "multicast" in getattr(merger, "transport", "") else world - 1)),
                "nvlink_gbs_per_gpu": None if pipelined else 24.0 * float(counts.sum()) * (world - 1) / max((t_step - t_frame) / args.steps, 1e-9) / 1e9,
                "nvlink_peak_gbs": 770.0, "transport": getattr(merger, "transport", "nccl"), "verified_against_nccl_allgatherv": merge_verified, "note": "rank 0's own lists delivered to every peer (one multicast store each when the transport says so); peak = measured peer copy per direction"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "frac_of_nominal_8000": achieved / 8000.0, "traffic": None, "peak_kind": peak_kind, "kernel": f"transform_cull_compact<{n_views}>",
                         "bytes_per_entity": bytes_per_entity, "ms_per_launch": t_launch * 1e3},
            "clocks": clocks.summary(),
        }
        if tri:
            n_cmd, kept = ctx.tri_counts()
            t_tri = (t_step - t_frame) / args.steps
            tris = int(totals[0]) // 3
            line["tri_stage"] = {
                "kernels": "tri_cull + compact_draw_stream", "ms_per_step": t_tri * 1e3, "triangles_in": tris, "triangles_kept": kept // 3,
                "commands_in": int(counts[0]), "commands_out": n_cmd, "gtri_per_s": tris / t_tri / 1e9,
                # 12 B index triple read per triangle, 12 B written per kept one, 20 + 4 B command in, 20 B out
                "algorithmic_gbs": (12.0 * tris + 4.0 * kept + 24.0 * int(counts[0]) + 20.0 * n_cmd) / t_tri / 1e9,
                "meshes": "grid sheets, row-major cell order (vertex reuse like an indexed glTF mesh)" if wl.get("coherent") else "random index triples (no vertex reuse)",
                "note": "e2e and cpu_baseline cover the frame stage only"}
        if world == 1:
            line["cpu_baseline"] = cpu_baseline(wl, steps=3, warmup=1, threads=os.cpu_count() or 1)
        if world == 1 and not args.no_also:
            line["also"] = [resident_timing(w, args.steps, local_rank) for w in ("16m_1view", "16m_5view", "1m_1view") if w != args.workload]
        prof = os.path.join(ROOT, "profiles", "r01_traffic.json")
        if os.path.exists(prof):
            try:
                line["roofline"]["traffic"] = json.load(open(prof)).get(args.workload)
            except Exception:
                pass
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


def reference_arm(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    b = cpu_baseline(wl, steps=args.steps, warmup=min(args.warmup, 2), threads=threads)
    n_sample = min(CPU_SAMPLE, wl["n"])
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": b["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": min(args.warmup, 2), "ms_per_step": b["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "entities_per_step": n_sample, "n_views": wl["n_views"], "half_box": wl["half_box"],
                   "note": "CPU oracle (C restatement of the reference's bevy systems; Rust toolchain absent), all host threads"},
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
    ap.add_argument("--no-overlap", action="store_true", help="N > 1: run frame and merge back to back instead of pipelined")
    ap.add_argument("--workload", default="16m_5view", choices=sorted(WORKLOADS))
    ap.add_argument("--no-also", action="store_true", help="N = 1: skip the resident timings of the other entity-stage workloads")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        reference_arm(args, wl)
    else:
        gpu_arm(args, wl)


if __name__ == "__main__":
    main()
