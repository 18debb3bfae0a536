#!/usr/bin/env python3
"""Parity report: fast traversal vs the oracle on the named configs (hit/bounce mismatches, colour RMS, step counts).
  python tools/parity_report.py [--gpu] [--small]      default: host simulation of the device headers (no GPU needed)
TEST TOOLING (uses oracle/ and tests/hostsim)."""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import common
import voxelraytracer_b200 as vrt

CASES = [("c1", "glass_cube32", "start", vrt.FLAG_COLOR_ONLY, 400, 400),
         ("c2", "terrain128", "terrain_inside", 0, 1920, 1080), ("c2", "terrain128", "terrain_low", 0, 1920, 1080),
         ("c2", "terrain128", "terrain_above", 0, 1920, 1080), ("c2", "terrain128", "terrain_outside", 0, 1920, 1080),
         ("c3", "refraction128", "refraction_close", 0, 1920, 1080), ("c4", "terrain128_glass", "glass_wall", 0, 1920, 1080),
         ("c4", "terrain128_glass", "terrain_inside", 0, 1920, 1080)]

def main():
    ap = argparse.ArgumentParser(); ap.add_argument("--gpu", action="store_true"); ap.add_argument("--small", action="store_true")
    a = ap.parse_args()
    ctx = vrt.Context(0) if a.gpu else None
    for cfg, sc, pose, fl, W, H in CASES:
        if a.small: W, H = W // 4, H // 4
        grid, n, atlas, ts = common.scene(sc)
        p = common.make_params(W, H, sc, pose, fl)
        o = common.oracle_render(p, grid, atlas)
        t0 = time.time()
        if ctx:
            ctx.upload_grid(grid, n); ctx.upload_atlas(atlas, ts)
            g = common.gpu_render(ctx, p)
        else:
            g = common.hostsim_render(p, grid, atlas, 1)
        bad_hit = int((g.hit != o.hit).sum()); bad_b = int((g.bounces != o.bounces).sum())
        bad = int(((g.hit != o.hit) | (g.bounces != o.bounces)).sum())
        oc = o.counters; ref_steps = oc["march_steps"] + oc["shadow_steps"]; rays = oc["primary_rays"] + oc["secondary_rays"] + oc["shadow_rays"]
        print(json.dumps({"cfg": cfg, "scene": sc, "pose": pose, "frame": [W, H], "mismatch_px": bad, "mismatch_frac": bad / (W * H), "hit_px": bad_hit,
                          "bounce_px": bad_b, "cap_px": int(1e-4 * W * H), "rms255": [round(float(x), 4) for x in common.rms255(g.rgba8, o.rgba8)],
                          "rays": rays, "ref_steps": ref_steps, "actual_steps": g.counters["actual_steps"], "tie3": oc["tie_index3"],
                          "impl": "gpu" if ctx else "hostsim"}))
        sys.stdout.flush()

if __name__ == "__main__":
    main()
