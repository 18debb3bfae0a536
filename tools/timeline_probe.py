#!/usr/bin/env python3
"""Per-CTA start / end times (globaltimer) of one render, from a build instrumented for it (build/vt: tools/variant.py new vt + the
g_timeline patch).  Shows where a frame's fixed cost goes: ramp-up, steady state, tail.   usage (GPU box): VRT_LIBRARY=build/vt/libvrt.so tools/timeline_probe.py"""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, common
import voxelraytracer_b200 as vrt
from voxelraytracer_b200 import sharding
ctx = vrt.Context(0)
lib = ctx.lib
out = ctx.device_alloc(3840 * 2160 * 4)
for name, sc, pose, W, H, band in (("c2", "terrain128", "terrain_inside", 1920, 1080, False), ("c2", "terrain128", "terrain_low", 1920, 1080, False),
                                    ("c4 full", "terrain128_glass", "glass_wall", 3840, 2160, False), ("c4 1/8 bands", "terrain128_glass", "glass_wall", 3840, 2160, True),
                                    ("c4 1/8 bands", "terrain128_glass", "terrain_inside", 3840, 2160, True)):
    grid, n, atlas, ts = common.scene(sc); ctx.upload_grid(grid, n); ctx.upload_atlas(atlas, ts)
    p = common.make_params(W, H, sc, pose)
    if band:
        p = sharding.band_params(p, 3, 8, 16); p.flags |= vrt.FLAG_OUT_FRAME
    for _ in range(6): ctx.render_to(p, out)
    ms = ctx.last_render_ms()
    rows = vrt.tile_pixels(p) // W
    ncta = ((W + 15) // 16) * ((rows + 7) // 8)
    t = np.zeros((ncta, 2), np.uint64); sm = np.zeros(ncta, np.uint32)
    lib.vrt_debug_timeline(t.ctypes.data_as(C.c_void_p), sm.ctypes.data_as(C.c_void_p), ncta)
    t0 = t[:, 0].min(); s = (t[:, 0] - t0).astype(np.float64) / 1e3; e = (t[:, 1] - t0).astype(np.float64) / 1e3      # us
    span = e.max(); dur = e - s
    last_start = s.max()
    # resident CTAs over time
    grid_t = np.linspace(0, span, 41)
    active = [(int(((s <= x) & (e > x)).sum())) for x in grid_t]
    first_full = next((x for x, a in zip(grid_t, active) if a >= 0.9 * max(active)), None)
    print("%-13s %-15s event ms %.4f | CTAs %d | first start -> last end %.1f us | last CTA starts at %.1f us (tail %.1f us) | CTA duration median %.1f p95 %.1f max %.1f us"
          % (name, pose, ms, ncta, span, last_start, span - last_start, np.median(dur), np.percentile(dur, 95), dur.max()))
    print("      resident CTAs at 0..100%% of the span (41 samples): %s" % " ".join(str(a) for a in active))
    order = np.argsort(e)[-5:]
    print("      last 5 CTAs to finish: " + ", ".join("cta %d (row %d) start %.1f dur %.1f" % (i, i // ((W + 15) // 16), s[i], dur[i]) for i in order))
