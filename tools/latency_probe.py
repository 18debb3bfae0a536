#!/usr/bin/env python3
"""Kernel time of very small tiles (one CTA ... one wave): the serial latency of a pixel, per pose."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch, common
import voxelraytracer_b200 as vrt
sc = "terrain128"
grid, n, atlas, ts = common.scene(sc)
ctx = vrt.Context(0); ctx.upload_grid(grid, n); ctx.upload_atlas(atlas, ts)
buf = torch.empty(1920 * 1080 * 4, dtype=torch.uint8, device="cuda")
for pose in ("terrain_inside", "terrain_low", ((0.0, 70.0, 0.0), (1.5, 0.0, 0.0))):   # last: looking straight up = sky only
    out = []
    for (tw, th) in ((32, 4), (128, 16), (512, 64), (1920, 96), (1920, 270), (1920, 1080)):
        p = common.make_params(1920, 1080, sc, pose)
        p.tile_x0, p.tile_y0, p.tile_w, p.tile_h = (1920 - tw) // 2, (1080 - th) // 2, tw, th
        for _ in range(5): ctx.render_to(p, buf.data_ptr(), None)
        t = []
        for _ in range(40):
            ctx.render_to(p, buf.data_ptr(), None); t.append(ctx.last_render_ms())
        out.append("%dx%d %.2f us" % (tw, th, 1e3 * float(np.median(t))))
    print(pose if isinstance(pose, str) else "sky_only", " | ".join(out))
