#!/usr/bin/env python3
"""Decodes the reference's texture tiles into tests/golden/ref_tiles.npz (TEST / BENCH INPUT FIXTURE).

The reference builds its atlas from res/textures/{stone,dirt,glass,grass}.png (16^2, _HIGH_PERFORMANCE) or the *128.png set
(src/main.cpp:176-196).  The PNG decoder lives in the un-vendored engine (FreeImage), so the tiles are decoded ONCE here, in the
authoring container where /root/reference and PIL exist, and stored as raw RGBA8 arrays in image order (row 0 = top).  RGB-only
files (stone128.png) get alpha 255, which is what an RGBA texture upload of an RGB image yields.  tests/common.py and bench.py
pack them through vrt_atlas_pack in AddTexture order (stone, dirt, glass, grass).

  python tests/golden/make_tiles.py          (needs /root/reference; the GPU box only reads the committed .npz)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/res/textures"


def main():
    from PIL import Image
    out = {}
    for suffix, ts in (("", 16), ("128", 128)):
        for name in ("stone", "dirt", "glass", "grass"):
            im = Image.open(os.path.join(SRC, name + suffix + ".png"))
            a = np.array(im.convert("RGBA"), np.uint8)
            assert a.shape == (ts, ts, 4), (name, a.shape)
            out["%s%d" % (name, ts)] = a
            print("%-6s %3d  mode %-4s  mean rgba %s  alpha values %s" % (name, ts, im.mode, np.round(a.reshape(-1, 4).mean(0), 1), np.unique(a[..., 3])))
    np.savez_compressed(os.path.join(HERE, "ref_tiles.npz"), **out)
    print("wrote ref_tiles.npz, %d bytes" % os.path.getsize(os.path.join(HERE, "ref_tiles.npz")))


if __name__ == "__main__":
    sys.exit(main())
