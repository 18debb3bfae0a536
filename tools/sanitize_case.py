#!/usr/bin/env python3
"""Small end-to-end case for compute-sanitizer (memcheck): repack (both layouts), every kernel variant, tiles/bands,
OUT_FRAME, persistent launch, temporal filter, host round trip.  usage: compute-sanitizer --tool memcheck python tools/sanitize_case.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, common
import voxelraytracer_b200 as vrt
from voxelraytracer_b200 import sharding
ctx = vrt.Context(0)
for octants in (1, 0):
    ctx.set_option(vrt.OPT_OCTANT_FIELDS, octants)
    for sc, pose, fl in (("terrain32", "terrain32_above", 0), ("glass_cube32", "start", vrt.FLAG_COLOR_ONLY), ("refraction32", "start", 0)):
        grid, n, atlas, ts = common.scene(sc)
        ctx.upload_grid(grid, n); ctx.upload_atlas(atlas, ts)
        W, H = 203, 117
        p = common.make_params(W, H, sc, pose, fl)
        full = common.gpu_render(ctx, p, flags_extra=0).rgba8
        for extra in (vrt.FLAG_DEBUG_MAPS | vrt.FLAG_FLOAT_OUT | vrt.FLAG_COUNTERS, vrt.FLAG_PERSISTENT):
            assert np.array_equal(common.gpu_render(ctx, p, flags_extra=extra).rgba8, full)
        pe = p.copy(); pe.flags |= vrt.FLAG_EXACT
        common.gpu_render(ctx, pe)
        d = ctx.device_alloc(W * H * 4)
        for r in range(3):
            q = sharding.band_params(p, r, 3, 16); q.flags |= vrt.FLAG_OUT_FRAME
            ctx.render_to(q, d, None)
        assert np.array_equal(ctx.copy_to_host(d, W * H * 4).reshape(-1, 4), full)
        d2 = ctx.device_alloc(W * H * 4)
        ctx.temporal_filter(d, d, d2, W * H, 0.5)
        ctx.synchronize(); ctx.device_free(d); ctx.device_free(d2)
        host, hp = ctx.host_alloc(W * H * 4)
        ctx.render_host(p, hp, None)
        assert np.array_equal(host.reshape(-1, 4), full)
ctx.close()
print("sanitize case ok")
