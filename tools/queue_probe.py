#!/usr/bin/env python3
"""VRT_FLAG_RAY_QUEUE against the serial per-pixel loop: frames must be byte-identical; kernel times side by side.
usage (GPU box): [VRT_LIBRARY=build/x/libvrt.so] python tools/queue_probe.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, common
import voxelraytracer_b200 as vrt
from voxelraytracer_b200 import sharding
ctx = vrt.Context(0)
flush_a, flush_b = ctx.device_alloc(128 << 20), ctx.device_alloc(128 << 20)
out = ctx.device_alloc(3840 * 2160 * 4)
def timed(p, reps=20):
    t = []
    for i in range(reps + 3):
        ctx.temporal_blend(flush_a, flush_b, 32 << 20, 0.5); ctx.render_to(p, out)
        if i >= 3: t.append(ctx.last_render_ms())
    return float(np.median(t))
bad = 0
for sc, pose, fl, W, H in (("refraction128", "refraction_close", 0, 1920, 1080), ("terrain128_glass", "glass_wall", 0, 3840, 2160), ("terrain128_glass", "terrain_inside", 0, 3840, 2160),
                           ("glass_cube32", "start", vrt.FLAG_COLOR_ONLY, 400, 400), ("glass_cube32", "start", 0, 1001, 563), ("terrain32", "terrain32_above", 0, 333, 77),
                           ("refraction128", "refraction_close", vrt.FLAG_COLOR_ONLY, 1920, 1080)):
    grid, n, atlas, ts = common.scene(sc); ctx.upload_grid(grid, n); ctx.upload_atlas(atlas, ts)
    p = common.make_params(W, H, sc, pose, fl)
    q = p.copy(); q.flags |= vrt.FLAG_RAY_QUEUE
    ctx.render(p); ctx.render(p); a = ctx.read_framebuffer(p)
    ctx.render(q); b = ctx.read_framebuffer(q)
    same = bool(np.array_equal(a, b)); bad += 0 if same else 1
    # one rank's eighth in bands
    pb = sharding.band_params(p, 3, 8, 16); qb = pb.copy(); qb.flags |= vrt.FLAG_RAY_QUEUE
    ctx.render(pb); ab = ctx.read_framebuffer(pb); ctx.render(qb); bb = ctx.read_framebuffer(qb)
    sameb = bool(np.array_equal(ab, bb)); bad += 0 if sameb else 1
    print("%-17s %-16s %dx%d flags %d: %s (%d px differ), bands %s | serial %.4f ms, queue %.4f ms | bands serial %.4f, queue %.4f" %
          (sc, pose, W, H, fl, "byte-identical" if same else "DIFFERENT", int((a != b).any(axis=1).sum()), "byte-identical" if sameb else "DIFFERENT", timed(p), timed(q), timed(pb), timed(qb)), flush=True)
print("FAILED" if bad else "all byte-identical")
