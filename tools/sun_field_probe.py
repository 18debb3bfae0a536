#!/usr/bin/env python3
"""One-time costs at 1024^3: upload + repack, and the sun field's build (VRT_OPT_SUN_FIELD 2), against what a frame gains.  GPU box."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, common
import voxelraytracer_b200 as vrt
ctx = vrt.Context(0)
for sc, W, H, ml, scale in (("synthetic512", 3840, 2160, 1024.0, 0.5), ("synthetic1024", 7680, 4320, 2048.0, 1.0)):
    grid, n, atlas, ts = common.scene(sc)
    t0 = time.perf_counter(); ctx.upload_grid(grid, n); t_up = time.perf_counter() - t0
    ctx.upload_atlas(atlas, ts)
    out = ctx.device_alloc(W * H * 4)
    p = common.make_params(W, H, sc, common.orbit_pose(5, 64, 600.0 * scale, 250.0 * scale), max_len=ml)
    res = {}
    for mode in (0, 2):
        ctx.set_option(vrt.OPT_SUN_FIELD, mode)
        ctx.synchronize(); t0 = time.perf_counter(); ctx.render_to(p, out); ctx.synchronize(); first = (time.perf_counter() - t0) * 1e3
        t = []
        for _ in range(6):
            ctx.render_to(p, out); t.append(ctx.last_render_ms())
        res[mode] = (first, float(np.median(t)))
    print("%s: upload + repack %.3f s | sun field off: frame %.3f ms | on: first frame incl. build %.1f ms, then %.3f ms per frame (%.1f %% faster); warning: %r" %
          (sc, t_up, res[0][1], res[2][0], res[2][1], 100 * (res[0][1] / res[2][1] - 1), ctx.last_warning()), flush=True)
    ctx.device_free(out)
    common._scene_cache.pop(sc, None)
