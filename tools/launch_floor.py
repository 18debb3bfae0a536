#!/usr/bin/env python3
"""Event-bracketed time of a trivial kernel (temporal blend of 4 pixels) vs the smallest renders: the launch floor."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch, common
import voxelraytracer_b200 as vrt
grid, n, atlas, ts = common.scene("terrain128")
ctx = vrt.Context(0); ctx.upload_grid(grid, n); ctx.upload_atlas(atlas, ts)
a = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda"); b = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda")
def ev_time(fn, reps=50):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    t = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize(); t.append(e0.elapsed_time(e1) * 1e3)
    return float(np.median(t))
print("trivial kernel (blend 4 px): %.2f us" % ev_time(lambda: ctx.temporal_blend(a.data_ptr(), b.data_ptr(), 4, 0.5)))
buf = torch.empty(1920 * 1080 * 4, dtype=torch.uint8, device="cuda")
for name, pose in (("sky_only", ((0.0, 70.0, 0.0), (1.5, 0.0, 0.0))), ("terrain_inside", "terrain_inside")):
    for fl, fname in ((0, "textured"), (vrt.FLAG_COLOR_ONLY, "flat"), (vrt.FLAG_EXACT, "exact")):
        p = common.make_params(1920, 1080, "terrain128", pose, fl)
        p.tile_x0, p.tile_y0, p.tile_w, p.tile_h = 944, 538, 32, 4
        print("%s %s 32x4: %.2f us" % (name, fname, ev_time(lambda: ctx.render_to(p, buf.data_ptr(), None))))
