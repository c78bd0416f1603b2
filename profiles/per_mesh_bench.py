#!/usr/bin/env python
"""BASELINE configs[3]: the Sponza mesh table instanced to 4 M entities, per-mesh indirect command
compaction (output_mode PER_MESH_INSTANCED), timed on one B200 beside the per-entity mode of the
same scene.  Prints one JSON line.  Not a bench.py line (bench.py measures the headline workload);
parity of this mode is tests/test_gpu_parity.py::test_config4_sponza_instanced_per_mesh.

    python profiles/per_mesh_bench.py [--n 4000000] [--views 1] [--steps 20]
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch

    from bench import L2_NOTE, flush_l2
    from renderer_b200 import PER_MESH_INSTANCED, CullContext, scene

    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=4_000_000)
    ap.add_argument("--views", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    args = ap.parse_args()
    table, node_scale = scene.sponza_mesh_table()
    ps, rot, mesh = scene.sponza_instances(args.n, len(table), node_scale)
    views = scene.scene_views(args.views)
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev)
    flush = torch.zeros((2, 256 << 20), dtype=torch.uint8, device=dev)
    out = {"workload": "4m_sponza_per_mesh", "entities": args.n, "meshes": len(table), "n_views": args.views, "steps": args.steps,
           "l2": L2_NOTE}
    for name, mode in (("per_entity", None), ("per_mesh", PER_MESH_INSTANCED)):
        kw = {} if mode is None else {"output_mode": mode}
        with CullContext(args.n, max_views=args.views, max_meshes=len(table), **kw) as ctx:
            ctx.set_mesh_table(table)
            ctx.upload_entities(ps, rot, mesh)
            ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(args.steps)]
            with torch.cuda.stream(stream):
                for _ in range(args.warmup):
                    ctx.frame(views, scene.CAMERA_POS, stream=stream.cuda_stream)
                    flush_l2(flush)
                torch.cuda.synchronize()
                l0 = ctx.launch_count
                for i in range(args.steps):
                    ev[i][0].record(stream)
                    ctx.frame(views, scene.CAMERA_POS, stream=stream.cuda_stream)
                    ev[i][1].record(stream)
                    flush_l2(flush)
                torch.cuda.synchronize()
            ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
            counts, totals = ctx.counts()
            out[name] = {"ms_per_frame": ms, "entities_per_s": args.n / ms * 1e3, "launches_per_frame": (ctx.launch_count - l0) / args.steps,
                         "commands": [int(c) for c in counts[: args.views]],
                         ("instances" if mode is not None else "index_totals"): [int(t) for t in totals[: args.views]]}
    out["bucketing_ms"] = out["per_mesh"]["ms_per_frame"] - out["per_entity"]["ms_per_frame"]
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
