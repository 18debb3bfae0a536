#!/usr/bin/env python3
"""Where the band-sharded (strong scaling) render loses time, on ONE GPU: kernel time of every band index of an 8-way split in isolation
(L2 flushed, library event pair), and the THROUGHPUT of the same launches with several in flight on separate streams (tails overlap the
next launch's ramp), against the full frame.  usage (GPU box): python tools/band_probe2.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch, common
import voxelraytracer_b200 as vrt
from voxelraytracer_b200 import sharding
ctx = vrt.Context(0)
sc, W, H = "terrain128_glass", 3840, 2160
grid, n, atlas, ts = common.scene(sc); ctx.upload_grid(grid, n); ctx.upload_atlas(atlas, ts)
out = torch.empty(W * H * 4, dtype=torch.uint8, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
streams = [torch.cuda.Stream() for _ in range(4)]
for pose in ("glass_wall", "terrain_inside"):
    full = common.make_params(W, H, sc, pose)
    def iso(p, reps=12):
        t = []
        for i in range(reps + 3):
            flush.fill_(i); ctx.render_to(p, out.data_ptr(), None)
            if i >= 3: t.append(ctx.last_render_ms())
        return float(np.median(t))
    def thr(ps, inflight, rounds):
        """ms per launch when `inflight` streams each run the list of launches round-robin"""
        for p in ps[:4]: ctx.render_to(p, out.data_ptr(), None)
        torch.cuda.synchronize(); t0 = time.perf_counter(); k = 0
        for r in range(rounds):
            for p in ps:
                ctx.render_to(p, out.data_ptr(), streams[k % inflight].cuda_stream); k += 1
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) * 1e3 / k
    bands = []
    for b in range(8):
        q = sharding.band_params(full, b, 8, 16); q.flags |= vrt.FLAG_OUT_FRAME; bands.append(q)
    t_full = iso(full); t_b = [iso(q) for q in bands]
    print("%s: full frame isolated %.4f ms (1/8 = %.4f) | bands isolated: %s | sum %.4f = %.2fx the full frame, max %.4f" %
          (pose, t_full, t_full / 8, " ".join("%.4f" % x for x in t_b), sum(t_b), sum(t_b) / t_full, max(t_b)))
    for inflight in (1, 2, 4):
        print("   %d in flight: full frame %.4f ms per frame | bands %.4f ms per band launch = %.4f per frame (%.2fx the full frame at the same depth)" %
              ((inflight, thr([full], inflight, 40)) + (lambda tb, tf: (tb, tb * 8, tb * 8 / tf))(thr(bands, inflight, 10), thr([full], inflight, 40))))
    # rows: the same 1/8 of the frame as ONE contiguous block of rows instead of interleaved bands
    blk = full.copy(); blk.tile_y0, blk.tile_h = 3 * (H // 8), H // 8
    print("   contiguous block of %d rows (rows %d..): isolated %.4f ms" % (H // 8, blk.tile_y0, iso(blk)))
