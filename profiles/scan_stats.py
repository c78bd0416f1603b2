#!/usr/bin/env python
"""Diagnostic (RCULL_SCAN_STATS builds only): per-tile resolution latency of the look-back and the time
compute warps spend blocked, from the counters the instrumented kernel accumulates.
usage: RCULL_LIB=<variant .so> python profiles/scan_stats.py [n_views] [frames]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
from renderer_b200 import CullContext, scene, _lib

n_views = int(sys.argv[1]) if len(sys.argv) > 1 else 5
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 10
n = 16 * 1024 * 1024
lib = C.CDLL(_lib.LIB_PATH)
lib.rcull_debug_stats.argtypes = [C.POINTER(C.c_uint64), C.c_int]
meshes = scene.synth_mesh_table()
views = scene.scene_views(n_views)
with CullContext(n, max_views=n_views, max_meshes=len(meshes)) as ctx:
    ctx.set_mesh_table(meshes)
    ctx.synth_entities(scene.SEED, 0, n, 600.0, len(meshes))
    for _ in range(3):
        ctx.frame(views, scene.CAMERA_POS)
    torch.cuda.synchronize()
    out = (C.c_uint64 * 16)()
    assert lib.rcull_debug_stats(out, 1) == 0
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(frames):
        ctx.frame(views, scene.CAMERA_POS)
    t1.record(); torch.cuda.synchronize()
    assert lib.rcull_debug_stats(out, 0) == 0
s = [int(x) for x in out]
ghz = 1.965
tiles, ctas = max(s[0], 1), max(s[10], 1)
print(f"{os.path.basename(_lib.LIB_PATH)} V={n_views}: {t0.elapsed_time(t1)/frames:.4f} ms/frame (default stream incl. atomics overhead)")
print(f"  tiles/frame {tiles/frames:.0f}  CTAs {ctas/frames:.0f}  kernel per CTA {s[7]/ctas/ghz/1e3:.1f} us  tile period {s[7]/tiles/ghz/1e3:.2f} us")
print(f"  publish->resolved {s[1]/tiles/ghz/1e3:.2f} us   activate->resolved {s[9]/tiles/ghz/1e3:.2f} us   passes/tile {s[2]/tiles:.2f}  good rounds/tile {s[3]/tiles:.2f}")
print(f"  per CTA: blocked on offsets {s[5]/ctas/ghz/1e3:.1f} us in {s[8]/ctas:.1f} waits; blocked on input {s[6]/ctas/ghz/1e3:.1f} us")
