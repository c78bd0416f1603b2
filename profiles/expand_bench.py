#!/usr/bin/env python
"""Single-GPU proxy for the N-rank record merge at full volume: `world` ranks emulated in one process on one GPU
(every rank's frame kernel stores its records into all `world` windows, all local), then rank 0's expand is timed
 (a) alone, in both shapes, and (b) beside rank 0's next frame kernel on a second stream.  No NVLink is involved: this
measures the expand kernel and its interference with the frame kernel at the survivor volume of the real N-GPU run.
Every frame of every rank completes before any expand is launched (no launch ever waits on another launch).

    python profiles/expand_bench.py [--world 8] [--n 16777216] [--views 5] [--steps 10]
"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch

    from renderer_b200 import CullContext, _lib, scene

    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--n", type=int, default=16 * 1024 * 1024)
    ap.add_argument("--views", type=int, default=5)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--sms", default="8,16,24", help="SM counts given to the expand in the split-chip runs")
    args = ap.parse_args()
    lib = _lib.load()
    meshes = scene.synth_mesh_table()
    views = scene.scene_views(args.views)
    dev = torch.device("cuda", 0)
    world, n = args.world, args.n
    # only rank 0 keeps a full context; the other ranks share ONE context that is re-synthesised per rank (their frames
    # only have to deliver records), so the proxy fits one GPU
    ctx0 = CullContext(n, max_views=args.views, max_meshes=len(meshes))
    ctxs = [ctx0] + [CullContext(n, max_views=args.views, max_meshes=len(meshes)) for _ in range(world - 1)]
    for r, c in enumerate(ctxs):
        c.set_mesh_table(meshes)
        c.synth_entities(scene.SEED, r * n, n, 600.0, len(meshes))
    ctx0.frame(views, scene.CAMERA_POS)
    own = int(ctx0.counts()[0][: args.views].max())
    cap_r, cap = int(own * 1.1) + 1024, int(own * 1.1 * world) + 1024
    merges, windows = [], []
    for r, c in enumerate(ctxs):
        h, w = C.c_void_p(), C.c_void_p()
        _lib.check(lib.rcull_merge_create(c._h, world, r, cap, cap_r, C.byref(h)))
        _lib.check(lib.rcull_merge_window(h, C.byref(w)))
        merges.append(h)
        windows.append(w)
    for h in merges:
        for p in range(world):
            _lib.check(lib.rcull_merge_set_peer_ptr(h, p, windows[p]))
    sa, sb = torch.cuda.Stream(device=dev, priority=-1), torch.cuda.Stream(device=dev, priority=0)

    def frames():
        for r, c in enumerate(ctxs):
            c.frame(views, scene.CAMERA_POS, draw_index_base=r * n, stream=sa.cuda_stream)
        torch.cuda.synchronize()

    def expands(skip0=False):
        for r, h in enumerate(merges):
            if r or not skip0:
                _lib.check(lib.rcull_merge_expand(h, C.c_void_p(sa.cuda_stream)))
        torch.cuda.synchronize()

    out = {"world": world, "entities_per_rank": n, "views": args.views}
    frames(); expands(); frames(); expands()  # warm
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))

    def set_sms(r_sms):
        for h in merges:
            _lib.check(lib.rcull_merge_set_overlap(h, r_sms))

    def other_frames():
        for r, c in enumerate(ctxs):
            if r:
                c.frame(views, scene.CAMERA_POS, draw_index_base=r * n, stream=sa.cuda_stream)
        torch.cuda.synchronize()

    for r_sms in [0] + [int(x) for x in args.sms.split(",")]:
        set_sms(r_sms)
        frames(); expands()
        t_e, t_f, t_both = [], [], []
        for _ in range(args.steps):  # (a) the expand alone (full chip, or r_sms whole-SM CTAs), (b) the frame alone on the other SMs
            frames()
            e0.record(sa)
            _lib.check(lib.rcull_merge_expand(merges[0], C.c_void_p(sa.cuda_stream)))
            e1.record(sa)
            torch.cuda.synchronize()
            t_e.append(e0.elapsed_time(e1))
            expands(skip0=True)
            e0.record(sa)
            ctx0.frame(views, scene.CAMERA_POS, draw_index_base=0, stream=sa.cuda_stream)
            e1.record(sa)
            torch.cuda.synchronize()
            t_f.append(e0.elapsed_time(e1))
            other_frames()
            expands()
        row = {"expand_alone_ms": float(np.mean(t_e)), "frame_alone_ms": float(np.mean(t_f))}
        if r_sms:
            t_f = []
            for _ in range(args.steps):  # (c) rank 0's expand on stream b beside rank 0's next frame on stream a
                frames()
                expands(skip0=True)
                e0.record(sa)
                sb.wait_event(e0)
                _lib.check(lib.rcull_merge_expand(merges[0], C.c_void_p(sb.cuda_stream)))
                ctx0.frame(views, scene.CAMERA_POS, draw_index_base=0, stream=sa.cuda_stream)
                e1.record(sa)
                sa.wait_stream(sb)
                e2.record(sa)
                torch.cuda.synchronize()
                t_f.append(e0.elapsed_time(e1))
                t_both.append(e0.elapsed_time(e2))
                other_frames()
                expands()
            row.update({"frame_ms_beside_expand": float(np.mean(t_f)), "frame_and_expand_ms": float(np.mean(t_both))})
        out[f"expand_sms_{r_sms}"] = row
    set_sms(0)
    frames()
    expands()
    cnt = np.zeros(8, np.uint32)
    cp = C.c_void_p()
    _lib.check(lib.rcull_merge_lists(merges[0], 0, None, None, C.byref(cp), None))
    from renderer_b200.shard import _wrap_device
    out["merged_survivors"] = int(_wrap_device(torch, cp.value, args.views, torch.int32, dev).cpu().numpy().view(np.uint32).astype(np.int64).sum())
    out["expand_algorithmic_mb"] = out["merged_survivors"] * 32 / 1e6
    print(json.dumps(out))
    for h in merges:
        lib.rcull_merge_destroy(h)
    for c in ctxs:
        c.destroy()


if __name__ == "__main__":
    main()
