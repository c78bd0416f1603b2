import ctypes as C, sys, os
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np, torch
from renderer_b200 import CullContext, scene, _lib
from renderer_b200.shard import shard_range
lib = _lib.load()
MESHES = scene.synth_mesh_table()
n, n_views, world = 120_001, 2, 2
dev = torch.device("cuda", 0)
streams = [torch.cuda.Stream(device=dev) for _ in range(world)]
ctxs, merges = [], []
for r in range(world):
    lo, hi = shard_range(n, world, r)
    c = CullContext(hi - lo, max_views=n_views, max_meshes=len(MESHES), n_buffers=2)
    c.set_mesh_table(MESHES)
    h = C.c_void_p()
    _lib.check(lib.rcull_merge_create(c._h, world, r, n, C.byref(h)))
    ctxs.append(c); merges.append(h)
bases = []
for h in merges:
    b = C.c_void_p(); _lib.check(lib.rcull_merge_base(h, C.byref(b))); bases.append(b)
for h in merges:
    for p in range(world):
        _lib.check(lib.rcull_merge_set_peer_ptr(h, p, bases[p]))
views = scene.scene_views(n_views)
ps, rot, mesh = scene.synth_entities(scene.SEED, 0, n, 50.0)
for r, c in enumerate(ctxs):
    lo, hi = shard_range(n, world, r)
    c.upload_entities(ps[lo:hi], rot[lo:hi], mesh[lo:hi])
    c.frame(views, scene.CAMERA_POS, draw_index_base=lo, buffer_ix=0, stream=streams[r].cuda_stream)
torch.cuda.synchronize(); print("frames done", flush=True)
order = [0, 1] if len(sys.argv) < 2 else [1, 0]
for r in order:
    _lib.check(lib.rcull_merge_exchange(merges[r], 0, n_views, C.c_void_p(streams[r].cuda_stream)))
    print("launched exchange", r, flush=True)
for r in order:
    _lib.check(lib.rcull_merge_wait(merges[r], C.c_void_p(streams[r].cuda_stream)))
print("launched waits", flush=True)
torch.cuda.synchronize(); print("done", flush=True)
