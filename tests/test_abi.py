"""The C-ABI library loads, exports every symbol include/vrt.h declares, and fails loudly without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import voxelraytracer_b200 as vrt
from voxelraytracer_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "vrt.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vrt_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = vrt.load_library()
    syms = header_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(lib, s), "libvrt.so does not export %s" % s
    assert sorted(api.EXPORTS) == syms, "api.EXPORTS out of sync with include/vrt.h"
    assert lib.vrt_abi_version() == 1


def test_params_struct_layout_matches_header():
    # 9 ints + 16 floats + 3 ints + 3 floats + 5 floats + 2 ints + 1 uint, all 4 bytes, no padding
    assert C.sizeof(vrt.Params) == 4 * (9 + 16 + 3 + 3 + 5 + 2 + 1)
    p = vrt.default_params(1920, 1080, 128)
    assert (p.width, p.height, p.tile_w, p.tile_h, p.size) == (1920, 1080, 1920, 1080, 128)
    assert p.maxRayLength == 100.0 and p.maxReflections == 1 and p.maxTransparencies == 2  # voxel.glsl:4-5,17
    assert p.time == 1.0 and p.rayNoise == 0.0 and p.flags == 0
    assert vrt.tile_pixels(p) == 1920 * 1080


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_gpu(), reason="CPU-only check")
def test_no_cpu_fallback():
    with pytest.raises(vrt.VrtError) as e:
        vrt.Context(0)
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)


def test_product_never_references_the_oracle():
    """Nothing under voxelraytracer_b200/ or include/ may import, link or name oracle/ or hostsim."""
    bad = []
    for base in ("voxelraytracer_b200", "include"):
        for d, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp", "Makefile")):
                    t = open(os.path.join(d, f), errors="ignore").read()
                    if re.search(r"liboracle|voracle_|vref_|libhostsim|hostsim_|oracle/|import oracle|from oracle", t):
                        # comments that merely mention tests/hostsim as a checker are fine; code references are not
                        for line in t.splitlines():
                            if re.search(r"liboracle|voracle_|vref_|libhostsim|hostsim_render|hostsim_build|import oracle|from oracle", line) \
                                    and not line.strip().startswith(("//", "#", "*", "/*")):
                                bad.append((f, line.strip()))
    assert not bad, bad


def test_scene_fillers_follow_main_cpp():
    n = 32
    g = vrt.scene_fill("glass_cube", n).reshape(n, n, n)  # [z][y][x]
    assert g[n // 2, n // 2, n // 2] == 3                 # main.cpp:271
    assert (g[0] == 2).all() and (g[:, 0] == 2).all() and (g[:, :, n - 1] == 2).all()
    assert g[1:-1, 1:-1, 1:-1].sum() == 3
    r = vrt.scene_fill("refraction", n).reshape(n, n, n)
    assert r[n // 2, n // 2, n // 2] == 2 and r[n // 4, n // 4, 0] == 3 and r[0, 0, 0] == 0   # main.cpp:274-287
    assert int((r == 3).sum()) == 6 * (n // 2) ** 2
    hm = vrt.noise_heightmap(n, 1, 5, 0.5)
    assert hm.min() >= 0.0 and hm.max() <= 0.999
    t = vrt.scene_fill("terrain", n, hm).reshape(n, n, n)
    x, z = 5, 7
    top = int(hm[z, x] * n)
    assert t[z, top, x] == 3 and (t[z, :top, x] == 1).all() and (t[z, top + 1:, x] == 0).all()  # main.cpp:225-230
    assert (t[5, int(hm[5, 0] * n + 1):, 0] == 2).all()      # glass wall x = 0, size <= 64 (main.cpp:235-241)
    t128 = vrt.scene_fill("terrain", 128)
    assert not (t128 == 2).any()                              # no glass at 128 (guard main.cpp:233)
    assert (vrt.scene_fill("terrain", 128, glass_walls=True) == 2).any()
    # determinism
    assert np.array_equal(t128, vrt.scene_fill("terrain", 128))


def test_camera_and_sun_helpers():
    s = vrt.sun_dir(0.0)
    np.testing.assert_allclose(s, np.array([0, 1, 0.2]) / np.sqrt(1.04), atol=1e-7)   # main.cpp:346-348 at timeOfDay 0
    s = vrt.sun_dir(12.5, 50.0)                                                    # quarter day: (sin, cos) = (1, 0)
    np.testing.assert_allclose(s, np.array([1, 0, 0.2]) / np.sqrt(1.04), atol=1e-6)
    m = vrt.camera_invpv((1.0, 2.0, 3.0), (0.0, 0.0, 0.0), 1.0).reshape(4, 4).T    # row-major view of a column-major matrix
    near = m @ np.array([0, 0, -1, 1.0])
    far = m @ np.array([0, 0, 1, 1.0])
    near, far = near[:3] / near[3], far[:3] / far[3]
    np.testing.assert_allclose(near, [1, 2, 3 - 0.01], atol=1e-4)                   # near plane 0.01 in front, looking down -z
    d = (far - near) / np.linalg.norm(far - near)
    np.testing.assert_allclose(d, [0, 0, -1], atol=1e-5)
    # yaw by -90 degrees looks toward +x (view = RotX(-rx) * RotY(-ry) * T(-pos), main.cpp:69)
    m = vrt.camera_invpv((0, 0, 0), (0.0, -np.pi / 2, 0.0), 1.0).reshape(4, 4).T
    near = m @ np.array([0, 0, -1, 1.0]); far = m @ np.array([0, 0, 1, 1.0])
    d = far[:3] / far[3] - near[:3] / near[3]
    np.testing.assert_allclose(d / np.linalg.norm(d), [1, 0, 0], atol=1e-5)


def test_atlas_layout():
    ts = 4
    tiles = [np.full((ts, ts, 4), 10 * (i + 1), np.uint8) for i in range(4)]
    a = vrt.atlas_pack(*tiles)
    # GL memory order: row 0 is the bottom of the image.  Image layout: stone TL, dirt TR, glass BL, grass BR.
    assert a[2 * ts - 1, 0, 0] == 10 and a[2 * ts - 1, 2 * ts - 1, 0] == 20 and a[0, 0, 0] == 30 and a[0, 2 * ts - 1, 0] == 40
    s = vrt.atlas_synthetic(16)
    glass = s[:16, :16]                      # bottom-left in memory = image bottom-left = glass
    assert (glass[0, :, 3] == 255).all() and glass[8, 8, 3] == 0   # opaque frame, clear interior (like res/textures/glass.png)


def test_png_export_round_trips(tmp_path):
    """vrt_write_png (stand-in for Utils::Screenshot, main.cpp:423-429): stored-deflate PNG that a stock decoder reads back."""
    import numpy as np
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(5)
    for (w, h) in ((1, 1), (517, 300), (400, 400)):          # 517*300*3 > 65535: several stored blocks
        a = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
        path = tmp_path / ("f%dx%d.png" % (w, h))
        vrt.write_png(path, a, w, h)
        im = np.asarray(Image.open(path).convert("RGB"))
        assert np.array_equal(im, a[::-1, :, :3])            # GL row order flipped to top-down, alpha dropped
