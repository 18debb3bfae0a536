#!/usr/bin/env python3
"""Frame time vs resolution for one pose: the intercept of the line is the kernel's fixed cost (launch + tail)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch, common
import voxelraytracer_b200 as vrt
sc = sys.argv[1] if len(sys.argv) > 1 else "terrain128"
poses = sys.argv[2:] or ["terrain_inside", "terrain_low", "terrain_above"]
grid, n, atlas, ts = common.scene(sc)
ctx = vrt.Context(0); ctx.upload_grid(grid, n); ctx.upload_atlas(atlas, ts)
buf = torch.empty(7680 * 4320 * 4, dtype=torch.uint8, device="cuda")
for pose in poses:
    xs, ys = [], []
    for (W, H) in ((480, 270), (960, 540), (1920, 1080), (3840, 2160), (7680, 4320)):
        p = common.make_params(W, H, sc, pose)
        for _ in range(5): ctx.render_to(p, buf.data_ptr(), None)
        t = []
        for _ in range(30):
            ctx.render_to(p, buf.data_ptr(), None); t.append(ctx.last_render_ms())
        xs.append(W * H); ys.append(float(np.median(t)))
    k, c = np.polyfit(xs[2:], ys[2:], 1)
    print(pose, " ".join("%dx%d %.4f" % (w, h, y) for (w, h), y in zip(((480, 270), (960, 540), (1920, 1080), (3840, 2160), (7680, 4320)), ys)),
          "| slope %.4f ms/Mpx intercept %.4f ms" % (k * 1e6, c))
