#!/usr/bin/env python3
"""Times builds of libvrt.so side by side on one GPU (no torch: starts in a second).
usage: tools/ab_time.py [--reps N] [--cases c2,c3,c4,c5s] LIB [LIB...]     (LIB = path of a libvrt.so build; '-' = the in-tree one)
Each build runs in its own process (VRT_LIBRARY), every frame is its own CUDA-event window (vrt_last_render_ms) and the L2 is
flushed before it by a 256 MiB temporal blend, like bench.py's kernel timing.  TEST TOOLING."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
CASES = {"c2": ("terrain128", ["terrain_inside", "terrain_low", "terrain_above"], 1920, 1080, None),
         "c3": ("refraction128", ["refraction_close"], 1920, 1080, None),
         "c4": ("terrain128_glass", ["glass_wall", "terrain_inside"], 3840, 2160, None),
         "c4b": ("terrain128_glass", ["glass_wall", "terrain_inside"], 3840, 2160, None),    # one rank's share of 8: 16-row bands, band 3
         "c5s": ("synthetic512", ["orbit0", "orbit21", "orbit42"], 3840, 2160, 1024.0)}


def worker(cases, reps):
    import numpy as np, common
    import voxelraytracer_b200 as vrt
    ctx = vrt.Context(0)
    nflush = 32 << 20                                   # pixels: 2 x 128 MiB read, 128 MiB written
    fa, fb = ctx.device_alloc(nflush * 4), ctx.device_alloc(nflush * 4)
    out = ctx.device_alloc(3840 * 2160 * 4)
    res = []
    for name in cases:
        sc, poses, W, H, max_len = CASES[name]
        grid, n, atlas, ts = common.scene(sc)
        ctx.upload_grid(grid, n); ctx.upload_atlas(atlas, ts)
        tot = 0.0
        for pose in poses:
            pz = common.orbit_pose(int(pose[5:]), 64, 300.0, 125.0) if pose.startswith("orbit") else pose
            p = common.make_params(W, H, sc, pz, 0, max_len=max_len)
            if name == "c4b":
                from voxelraytracer_b200 import sharding
                p = sharding.band_params(p, 3, 8, 16)
                p.flags |= vrt.FLAG_OUT_FRAME
            t = []
            for i in range(reps + 5):
                ctx.temporal_blend(fa, fb, nflush, 0.5)
                ctx.render_to(p, out)
                if i >= 5: t.append(ctx.last_render_ms())
            tot += float(np.mean(t))
            res.append("%s/%s %.4f" % (name, pose, float(np.median(t))))
        res.append("%s mean %.4f ms" % (name, tot / len(poses)))
    print(" | ".join(res), flush=True)


if __name__ == "__main__":
    a = sys.argv[1:]
    reps, cases = 40, ["c2", "c3", "c4"]
    while a and a[0].startswith("--"):
        if a[0] == "--reps": reps = int(a[1])
        elif a[0] == "--cases": cases = a[1].split(",")
        elif a[0] == "--worker": worker(cases, reps); sys.exit(0)
        a = a[1:] if a[0] == "--worker" else a[2:]
    for lib in a:
        env = dict(os.environ)
        if lib != "-": env["VRT_LIBRARY"] = os.path.abspath(lib)
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--reps", str(reps), "--cases", ",".join(cases), "--worker"], env=env, capture_output=True, text=True, timeout=300)
        print("%-28s %s" % (os.path.basename(lib), r.stdout.strip() if r.returncode == 0 else "FAILED: " + r.stderr.strip()[-300:]), flush=True)
