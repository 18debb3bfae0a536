#!/usr/bin/env python3
"""Generates tests/golden/*.npz from the REFERENCE SHADER (oracle/_ref/libvoxel_ref.so = the reference's
res/shaders/voxel.glsl compiled as C++ by oracle/build_ref.py).  Run in the authoring container, where
/root/reference exists:  python tests/golden/make_golden.py
Inputs (grid, atlas, camera) come from the deterministic host helpers; only their checksums are stored."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import common  # noqa: E402
import voxelraytracer_b200 as vrt  # noqa: E402

CASES = [
    dict(name="c1_glass_cube_color_only", scene="glass_cube32", pose="start", flags=vrt.FLAG_COLOR_ONLY, width=100, height=100, noise=[0, 0, 0], time=1.0),
    dict(name="glass_cube_textured", scene="glass_cube32", pose="start", flags=0, width=96, height=54, noise=[0, 0, 0], time=1.0),
    dict(name="terrain32_glass_walls", scene="terrain32", pose="terrain32_above", flags=0, width=96, height=54, noise=[0, 0, 0], time=1.0),
    dict(name="c2_terrain128", scene="terrain128", pose="terrain_inside", flags=0, width=128, height=72, noise=[0, 0, 0], time=1.0),
    dict(name="c2_terrain128_outside", scene="terrain128", pose="terrain_outside", flags=0, width=128, height=72, noise=[0, 0, 0], time=1.0),
    dict(name="c3_refraction128", scene="refraction128", pose="refraction_close", flags=0, width=128, height=72, noise=[0, 0, 0], time=1.0),
    dict(name="c4_terrain128_glass", scene="terrain128_glass", pose="glass_wall", flags=0, width=128, height=72, noise=[0, 0, 0], time=1.0),
    dict(name="refraction_noise", scene="refraction128", pose="refraction_close", flags=vrt.FLAG_COLOR_ONLY, width=96, height=54, noise=[0.01, 0.05, 0.05], time=5.0),
]


def main():
    assert common.ref_lib() is not None, "oracle/_ref/libvoxel_ref.so missing: needs /root/reference"
    index = []
    for c in CASES:
        grid, n, atlas, ts = common.scene(c["scene"])
        p = common.make_params(c["width"], c["height"], c["scene"], c["pose"], c["flags"], noise=tuple(c["noise"]), time=c["time"])
        r = common.ref_render(p, grid, atlas)
        fn = c["name"] + ".npz"
        np.savez_compressed(os.path.join(HERE, fn), f32=r.f32, fetch3d=r.fetch3d)
        c = dict(c, file=fn, grid_sum=int(grid.astype(np.uint64).sum()), atlas_sum=int(atlas.astype(np.uint64).sum()))
        index.append(c)
        print("wrote", fn, r.f32.shape)
    json.dump(index, open(os.path.join(HERE, "index.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
