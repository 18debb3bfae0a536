#!/usr/bin/env python3
"""Minimal driver for an ncu capture of the ray-queue kernel: C4 glass_wall pose, serial then queue, a few launches each."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import common
import voxelraytracer_b200 as vrt
ctx = vrt.Context(0)
sc, pose, W, H = "terrain128_glass", "glass_wall", 1920, 1080
grid, n, atlas, ts = common.scene(sc); ctx.upload_grid(grid, n); ctx.upload_atlas(atlas, ts)
p = common.make_params(W, H, sc, pose); q = p.copy(); q.flags |= vrt.FLAG_RAY_QUEUE
for _ in range(3): ctx.render(p)
for _ in range(3): ctx.render(q)
ctx.synchronize()
print("serial %.4f" % 0, "done")
