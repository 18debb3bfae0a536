#!/bin/bash
# compute-sanitizer is closed on this GPU pool, so the memory-safety check of the traversal code is done on the host:
# the device headers (rt_*.cuh) are compiled for the CPU with AddressSanitizer + UBSan (tests/hostsim) and every named
# scene is rendered with both tracers and both cell layouts.  GPU tests assert that the GPU computes exactly what this
# host build computes (hit map, step counts), so an out-of-bounds cell index here is one there.
set -e
cd "$(dirname "$0")/.."
g++ -std=c++17 -O1 -g -fsanitize=address,undefined -fno-sanitize-recover=undefined -ffp-contract=off -fno-fast-math -fPIC -shared -pthread \
    -Wno-unknown-pragmas tests/hostsim/hostsim.cpp voxelraytracer_b200/csrc/host_scene.cpp -o /tmp/libhostsim_asan.so
cat > /tmp/asan_run.py <<'PY'
import sys, ctypes as C
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import common
common._libs['h'] = C.CDLL('/tmp/libhostsim_asan.so')
for octants in (1, 0):
    for sc, pose, fl, W, H in (("terrain128", "terrain_inside", 0, 480, 270), ("terrain128", "terrain_outside", 0, 480, 270),
                               ("terrain128_glass", "glass_wall", 0, 480, 270), ("refraction128", "refraction_close", 0, 480, 270),
                               ("glass_cube32", "start", 1, 400, 400), ("terrain32", "terrain32_above", 0, 320, 180)):
        grid, n, atlas, ts = common.scene(sc)
        p = common.make_params(W, H, sc, pose, fl)
        for mode in (0, 1):
            g = common.hostsim_render(p, grid, atlas, mode, nthreads=8, octants=octants)
        print(sc, pose, "octants" if octants else "isotropic", g.counters['actual_steps'])
grid, n, atlas, ts = common.scene("synthetic256")
p = common.make_params(320, 180, "synthetic256", common.orbit_pose(3, 64, 150.0, 62.0), max_len=600)
print("synthetic256", common.hostsim_render(p, grid, atlas, 1, nthreads=8).counters['actual_steps'])
print("asan/ubsan clean")
PY
LD_PRELOAD=$(g++ -print-file-name=libasan.so) ASAN_OPTIONS=detect_leaks=0 python /tmp/asan_run.py
