#!/usr/bin/env python
"""Static SASS opcode mix of one kernel of librcull.so (no GPU needed).
    python profiles/sass_static.py 'transform_cull_compact<5, false>' [lib.so]
Prints the static instruction count, the opcode histogram and the FP32 mnemonics that matter for the
contraction rule (FFMA / FFMA2 must be absent from the frame kernel)."""
import collections, re, subprocess, sys
pat = sys.argv[1]
lib = sys.argv[2] if len(sys.argv) > 2 else "renderer_b200/librcull.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
cur, funcs = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        funcs[cur] = []
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m and cur:
        funcs[cur].append(m.group(2).strip())
for name, body in funcs.items():
    if pat not in name:
        continue
    ops = collections.Counter()
    for ins in body:
        parts = ins.split()
        op = parts[1] if parts[0].startswith("@") else parts[0]
        ops[op.split(".")[0]] += 1
    print(name.split("(")[0], "static instructions", len(body))
    print("  " + ", ".join(f"{k} {v}" for k, v in ops.most_common(24)))
