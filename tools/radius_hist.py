#!/usr/bin/env python3
"""Histogram of the empty radius at the start of every loop iteration of the fast traversal (host build of the device headers
with -DVRT_HOST_HIST): where the iterations of a pose go.  usage: tools/radius_hist.py [scene pose W H]...   TEST TOOLING."""
import ctypes as C, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import common
cs = os.path.join(ROOT, "voxelraytracer_b200", "csrc")
so = "/tmp/libhostsim_hist.so"
subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-shared", "-pthread", "-DVRT_HOST_HIST",
                       "-Wno-unknown-pragmas", os.path.join(ROOT, "tests", "hostsim", "hostsim.cpp"), os.path.join(cs, "host_scene.cpp"), "-o", so])
lib = C.CDLL(so); common._libs["h"] = lib
cases = [("terrain128", "terrain_inside"), ("terrain128", "terrain_low"), ("terrain128", "terrain_above"), ("terrain128_glass", "glass_wall")]
if len(sys.argv) > 2:
    cases = [(sys.argv[i], sys.argv[i + 1]) for i in range(1, len(sys.argv) - 1, 2)]
for sc, pose in cases:
    grid, n, atlas, ts = common.scene(sc)
    p = common.make_params(960, 540, sc, pose)
    h = (C.c_ulonglong * 256)()
    lib.hostsim_radius_hist(h, 1)
    g = common.hostsim_render(p, grid, atlas, 1)
    lib.hostsim_radius_hist(h, 1)
    a = np.array(list(h), dtype=np.float64); tot = a.sum()
    rays = g.counters["primary_rays"] + g.counters["secondary_rays"] + g.counters["shadow_rays"]
    parts = ["r=%d %.1f%%" % (r, 100 * a[r] / tot) for r in range(0, 8)] + ["r=8-15 %.1f%%" % (100 * a[8:16].sum() / tot), "r>=16 %.1f%%" % (100 * a[16:].sum() / tot)]
    print("%-16s %-15s iterations %.2f M (%.1f per ray)  " % (sc, pose, tot / 1e6, tot / rays) + "  ".join(parts))
