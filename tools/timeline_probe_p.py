#!/usr/bin/env python3
"""Per-PATCH timeline of the persistent kernel (instrumented build).  usage: VRT_LIBRARY=build/vp/libvrt.so tools/timeline_probe_p.py"""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, common
import voxelraytracer_b200 as vrt
ctx = vrt.Context(0); lib = ctx.lib
out = ctx.device_alloc(3840 * 2160 * 4)
for sc, pose, W, H in (("terrain128", "terrain_inside", 1920, 1080), ("terrain128", "terrain_low", 1920, 1080)):
    grid, n, atlas, ts = common.scene(sc); ctx.upload_grid(grid, n); ctx.upload_atlas(atlas, ts)
    p = common.make_params(W, H, sc, pose)
    for _ in range(6): ctx.render_to(p, out)
    ms = ctx.last_render_ms()
    npatch = (W // 8) * (H // 4)
    t = np.zeros((npatch, 2), np.uint64); sm = np.zeros(npatch, np.uint32)
    lib.vrt_debug_timeline(t.ctypes.data_as(C.c_void_p), sm.ctypes.data_as(C.c_void_p), npatch)
    t0 = t[:, 0].min(); s = (t[:, 0] - t0).astype(np.float64) / 1e3; e = (t[:, 1] - t0).astype(np.float64) / 1e3
    dur = e - s; span = e.max()
    grid_t = np.linspace(0, span, 41)
    active = [int(((s <= x) & (e > x)).sum()) for x in grid_t]
    print("%s event ms %.4f span %.1f us | patches %d | patch duration median %.1f p95 %.1f max %.1f | per-SM busy-until: min %.1f max %.1f" %
          (pose, ms, span, npatch, np.median(dur), np.percentile(dur, 95), dur.max(), min(e[sm == k].max() for k in np.unique(sm)), max(e[sm == k].max() for k in np.unique(sm))))
    print("   active patches over the span: " + " ".join(str(a) for a in active))
    cnt = np.bincount(sm); print("   patches per SM: min %d max %d; SMs seen %d" % (cnt[cnt > 0].min(), cnt.max(), (cnt > 0).sum()))
