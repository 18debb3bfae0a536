#!/usr/bin/env python3
"""Side-by-side builds of the kernels for A/B timing (authoring-side tooling).

  tools/variant.py new NAME            copy voxelraytracer_b200/csrc (as it is now) to build/NAME/csrc
  tools/variant.py build NAME [EXTRA]  make there -> build/NAME/libvrt.so      (EXTRA = extra nvcc flags, e.g. -DVRT_MIN_BLOCKS=10)
  tools/variant.py check NAME          compile the variant's device headers for the host and compare its frames, hit maps, bounce maps and
                                       loop-iteration counts with the in-tree headers on the recorded poses: a variant is kept only when
                                       every frame is bit-identical (or its differences are understood and re-checked against the oracle)
  tools/variant.py adopt NAME          copy the variant's csrc back over the in-tree one

build/ is git-ignored but travels to the GPU box, where tools/ab_time.py times build/*/libvrt.so next to the in-tree library."""
import ctypes as C
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "voxelraytracer_b200", "csrc")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))

CHECK = [("terrain128", "terrain_inside", 0, 640, 360), ("terrain128", "terrain_low", 0, 640, 360), ("terrain128", "terrain_above", 0, 640, 360),
         ("terrain128", "terrain_outside", 0, 640, 360), ("refraction128", "refraction_close", 0, 640, 360), ("terrain128_glass", "glass_wall", 0, 640, 360),
         ("terrain128_glass", "terrain_inside", 0, 640, 360), ("glass_cube32", "start", 1, 400, 400), ("terrain32", "terrain32_above", 0, 480, 270)]


def vdir(name):
    return os.path.join(ROOT, "build", name)


def hostsim_of(name):
    """libhostsim.so built against the variant's headers (None = in-tree)."""
    import common
    if name is None:
        return C.CDLL(common.build_hostsim())
    d = vdir(name)
    src = open(os.path.join(ROOT, "tests", "hostsim", "hostsim.cpp")).read().replace("../../voxelraytracer_b200/csrc/", os.path.join(d, "csrc") + "/")
    os.makedirs(os.path.join(d, "hostsim"), exist_ok=True)
    cpp = os.path.join(d, "hostsim", "hostsim.cpp")
    open(cpp, "w").write(src)
    so = os.path.join(d, "hostsim", "libhostsim.so")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-shared", "-pthread", "-Wno-unknown-pragmas",
                           "-I", os.path.join(ROOT, "tests", "hostsim"), cpp, os.path.join(d, "csrc", "host_scene.cpp"), "-o", so])
    return C.CDLL(so)


def main():
    cmd, name = sys.argv[1], sys.argv[2]
    d = vdir(name)
    if cmd == "new":
        if os.path.exists(d):
            sys.exit("exists: " + d)
        os.makedirs(d)
        shutil.copytree(CSRC, os.path.join(d, "csrc"), ignore=shutil.ignore_patterns("*.o", "*.log"))
        shutil.copytree(CSRC, os.path.join(d, "orig"), ignore=shutil.ignore_patterns("*.o", "*.log"))   # what `adopt` diffs against
        # the variant's headers include ../../include/vrt.h relative to csrc: give it one
        os.makedirs(os.path.join(ROOT, "build", "include"), exist_ok=True)
        shutil.copy(os.path.join(ROOT, "include", "vrt.h"), os.path.join(ROOT, "build", "include", "vrt.h"))
        print(d)
    elif cmd == "build":
        extra = " ".join(sys.argv[3:])
        r = subprocess.run(["make", "-C", os.path.join(d, "csrc"), "../libvrt.so", "EXTRA=" + extra], capture_output=True, text=True)
        if r.returncode:
            sys.exit(r.stdout[-3000:] + r.stderr[-3000:])
        log = open(os.path.join(d, "csrc", "kernels.ptxas.log")).read().split("\n")
        for i, l in enumerate(log):
            if "render_kernelINS_10FastTracerILb0ELb0ELb0ELb" in l and "ELi2ELi2" in l and "Compiling" in l:
                print(l.split("'")[1][40:70], "|", log[i + 1].strip(), "|", log[i + 2].strip()[13:60])
    elif cmd == "check":
        import numpy as np
        import common
        import voxelraytracer_b200 as vrt  # noqa: F401
        libs = [hostsim_of(None), hostsim_of(name)]
        bad = 0
        for sc, pose, fl, W, H in CHECK:
            grid, n, atlas, ts = common.scene(sc)
            p = common.make_params(W, H, sc, pose, fl)
            out = []
            for lib in libs:
                common._libs["h"] = lib
                common._cells_cache.clear()
                out.append(common.hostsim_render(p, grid, atlas, 1))
            a, b = out
            same = np.array_equal(a.f32, b.f32) and np.array_equal(a.hit, b.hit) and np.array_equal(a.bounces, b.bounces)
            print("%-18s %-16s %s  iterations %d -> %d" % (sc, pose, "bit-identical" if same else "DIFFERENT (%d px)" % int((a.f32 != b.f32).any(axis=1).sum()),
                                                           a.counters["actual_steps"], b.counters["actual_steps"]))
            bad += 0 if same else 1
        sys.exit(1 if bad else 0)
    elif cmd == "adopt":
        # only the files the variant CHANGED, and only if the in-tree file is still what the variant started from
        for f in sorted(os.listdir(os.path.join(d, "csrc"))):
            if not (f.endswith((".cu", ".cuh", ".h", ".cpp")) or f == "Makefile"):
                continue
            new, orig, cur = (open(os.path.join(x, f)).read() if os.path.exists(os.path.join(x, f)) else None for x in (os.path.join(d, "csrc"), os.path.join(d, "orig"), CSRC))
            if new == orig:
                continue
            if cur != orig:
                sys.exit("REFUSED: %s changed in-tree since the variant was made; merge by hand" % f)
            shutil.copy(os.path.join(d, "csrc", f), os.path.join(CSRC, f))
            print("adopted", f)
    else:
        sys.exit(__doc__)


if __name__ == "__main__":
    main()
