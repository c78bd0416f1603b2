#!/usr/bin/env python
"""Peer-copy (copy engine) bandwidth GPU 0 -> GPU 1 at the sizes the record merge sends, alone and while a
memory-heavy kernel runs on the source GPU (what rcull_merge_send's copies meet under the next frame).
usage: python profiles/p2p_copy_bench.py   (2 GPUs of one box)"""
import json

import torch

assert torch.cuda.device_count() >= 2
d0, d1 = torch.device("cuda", 0), torch.device("cuda", 1)
out = {}
for mb in (1, 4, 15, 64, 256):
    n = mb * 1000 * 1000
    src = torch.empty(n, dtype=torch.uint8, device=d0)
    dst = torch.empty(n, dtype=torch.uint8, device=d1)
    big_a = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device=d0)
    big_b = torch.empty_like(big_a)
    s_copy, s_load = torch.cuda.Stream(d0), torch.cuda.Stream(d0)
    for busy in (False, True):
        with torch.cuda.device(d0):
            for _ in range(3):
                dst.copy_(src, non_blocking=True)
            torch.cuda.synchronize(d0)
            reps = 10
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if busy:
                with torch.cuda.stream(s_load):
                    for _ in range(40):
                        big_b.copy_(big_a)  # ~1 GB of HBM traffic per copy on the source GPU
            with torch.cuda.stream(s_copy):
                e0.record()
                for _ in range(reps):
                    dst.copy_(src, non_blocking=True)
                e1.record()
            torch.cuda.synchronize(d0)
            ms = e0.elapsed_time(e1) / reps
        out[f"{mb}MB_{'under_hbm_load' if busy else 'alone'}"] = {"ms": round(ms, 4), "gbs": round(n / ms / 1e6, 1)}
print(json.dumps(out))
