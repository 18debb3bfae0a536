#!/usr/bin/env python3
"""Times one rank's share of a band-sharded frame on ONE GPU (no gather): separates tail / imbalance from NVLink store cost."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch, common
import voxelraytracer_b200 as vrt
from voxelraytracer_b200 import sharding
sc, pose, W, H = "terrain128_glass", sys.argv[1] if len(sys.argv) > 1 else "glass_wall", 3840, 2160
grid, n, atlas, ts = common.scene(sc)
ctx = vrt.Context(0); ctx.upload_grid(grid, n); ctx.upload_atlas(atlas, ts)
buf = torch.empty(W * H * 4, dtype=torch.uint8, device="cuda")
def timeit(p, reps=30):
    for _ in range(5): ctx.render_to(p, buf.data_ptr(), None)
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        ctx.render_to(p, buf.data_ptr(), None); ts.append(ctx.last_render_ms())
    return float(np.median(ts))
full = common.make_params(W, H, sc, pose)
t_full = timeit(full)
print("full frame %.4f ms; ideal 1/8 = %.4f" % (t_full, t_full / 8))
for br in (16, 8, 4, 32):
    t = [timeit(sharding.band_params(full, r, 8, br)) for r in range(8)]
    print("band_rows %2d: per-rank ms min %.4f max %.4f mean %.4f  (max/ideal %.2f)" % (br, min(t), max(t), np.mean(t), max(t) / (t_full / 8)))
