"""How exposed is the visible set to the UNPINNED operation orders?  (CPU only, numpy float32.)

The oracle adopts SURVEY 8a's statement of nalgebra 0.29 / ncollide3d 0.32 (crates that are not under
/root/reference and cannot be built here) as the definition of rows A1-A3.  This study re-evaluates the
CoarseCulled flags of the benchmark scenes under other add orders / formula shapes a real crate version
could plausibly use, and counts the entity-view flags that change.  It pins nothing -- it measures how many
flags hang on each assumption, so a reader knows what an order mistake would cost.

    python profiles/order_sensitivity.py [n_entities] [half_box]

Variants (one change at a time against the adopted order):
  dot4_ltr    s = ((px*cx + py*cy) + pz*cz) + pw        instead of (px*cx + pz*cz) + (py*cy + pw)
  dot3_rtl    e = hx*|px| + (hy*|py| + hz*|pz|)         instead of (hx*|px| + hy*|py|) + hz*|pz|
  rot_1m2     R diagonal as 1 - 2*(jj + kk) etc.        instead of (ww + ii) - jj - kk
  gemv_rows   corner p = (m0*x + m2*z) + (m1*y + t)     (row-times-vector as an unrolled 4-dot) instead of the column axpy
              chain ((m0*x + m1*y) + m2*z) + t
  box_direct  planes tested on the folded min/max box   instead of the from_half_extents round trip
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from renderer_b200 import scene  # noqa: E402  (host-side input builders only)

F = np.float32


def model(ps, q, rot_1m2=False):
    i, j, k, w, s = q[:, 0], q[:, 1], q[:, 2], q[:, 3], ps[:, 3]
    ww, ii, jj, kk = w * w, i * i, j * j, k * k
    ij, wk, wj = i * j * F(2), w * k * F(2), w * j * F(2)
    ik, jk, wi = i * k * F(2), j * k * F(2), w * i * F(2)
    m = np.zeros((len(s), 16), F)
    if rot_1m2:
        d0, d1, d2 = F(1) - F(2) * (jj + kk), F(1) - F(2) * (ii + kk), F(1) - F(2) * (ii + jj)
    else:
        d0, d1, d2 = ww + ii - jj - kk, ww - ii + jj - kk, ww - ii - jj + kk
    m[:, 0] = d0 * s; m[:, 4] = (ij - wk) * s; m[:, 8] = (wj + ik) * s
    m[:, 1] = (wk + ij) * s; m[:, 5] = d1 * s; m[:, 9] = (jk - wi) * s
    m[:, 2] = (ik - wj) * s; m[:, 6] = (wi + jk) * s; m[:, 10] = d2 * s
    m[:, 12:15] = ps[:, :3]
    m[:, 15] = 1
    return m


def aabb(m, mn, mx, gemv_rows=False, box_direct=False):
    lo = np.full((len(m), 3), np.finfo(F).max, F)
    hi = np.full((len(m), 3), np.finfo(F).min, F)
    for cy in (mn[:, 1], mx[:, 1]):
        for cz in (mn[:, 2], mx[:, 2]):
            for cx in (mn[:, 0], mx[:, 0]):
                for a in range(3):
                    if gemv_rows:
                        v = (m[:, a] * cx + m[:, 8 + a] * cz) + (m[:, 4 + a] * cy + m[:, 12 + a] * F(1))
                    else:
                        v = m[:, a] * cx
                        v = m[:, 4 + a] * cy + v
                        v = m[:, 8 + a] * cz + v
                        v = m[:, 12 + a] * F(1) + v
                    lo[:, a] = np.minimum(lo[:, a], v)
                    hi[:, a] = np.maximum(hi[:, a], v)
    if box_direct:
        return lo, hi
    c, h = (hi + lo) / F(2), (hi - lo) / F(2)
    return c - h, c + h


def culled(mins, maxs, planes, dot4_ltr=False, dot3_rtl=False):
    h, c = (maxs - mins) * F(0.5), (mins + maxs) * F(0.5)
    out = np.zeros(len(h), bool)
    for p in planes:
        a = np.abs(p[:3])
        if dot3_rtl:
            e = h[:, 0] * a[0] + (h[:, 1] * a[1] + h[:, 2] * a[2])
        else:
            e = (h[:, 0] * a[0] + h[:, 1] * a[1]) + h[:, 2] * a[2]
        if dot4_ltr:
            s = ((p[0] * c[:, 0] + p[1] * c[:, 1]) + p[2] * c[:, 2]) + p[3] * F(1)
        else:
            s = (p[0] * c[:, 0] + p[2] * c[:, 2]) + (p[1] * c[:, 1] + p[3] * F(1))
        out |= (s - e) > 0
    return out


def flags(ps, rot, mesh, meshes, views, **variant):
    m = model(ps, rot, rot_1m2=variant.get("rot_1m2", False))
    mins, maxs = aabb(m, meshes["aabb_min"][mesh], meshes["aabb_max"][mesh], gemv_rows=variant.get("gemv_rows", False),
                      box_direct=variant.get("box_direct", False))
    return np.stack([culled(mins, maxs, v, dot4_ltr=variant.get("dot4_ltr", False), dot3_rtl=variant.get("dot3_rtl", False))
                     for v in views])


def study(n, half_box, n_views=5, chunk=1 << 20):
    meshes = scene.synth_mesh_table()
    views = scene.scene_views(n_views)
    names = ["dot4_ltr", "dot3_rtl", "rot_1m2", "gemv_rows", "box_direct"]
    diff = {k: 0 for k in names}
    visible = 0
    for first in range(0, n, chunk):
        c = min(chunk, n - first)
        ps, rot, mesh = scene.synth_entities(scene.SEED, first, c, half_box, len(meshes))
        base = flags(ps, rot, mesh, meshes, views)
        visible += int((~base).sum())
        for k in names:
            diff[k] += int((flags(ps, rot, mesh, meshes, views, **{k: True}) != base).sum())
    return {"n_entities": n, "half_box": half_box, "n_views": n_views, "flags": n * n_views, "visible": visible, "changed": diff}


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    hb = float(sys.argv[2]) if len(sys.argv) > 2 else 250.0
    import json

    print(json.dumps(study(n, hb)))
