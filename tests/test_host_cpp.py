"""The C++ host above the C ABI (voxelraytracer_b200/host: FrameBuffer / Cam / AppScene mirroring src/FrameBuffer.h and
src/main.cpp): builds with g++ alone, fails loudly without a GPU, and on a GPU renders the same bytes as the ctypes path."""
import os
import subprocess

import numpy as np
import pytest

import voxelraytracer_b200 as vrt
from common import ROOT, DEG

HOST = os.path.join(ROOT, "voxelraytracer_b200", "host")
DEMO = os.path.join(HOST, "vrt_demo")


def build_demo():
    r = subprocess.run(["make", "-C", HOST], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    return DEMO


def test_host_builds_and_has_no_cpu_path():
    demo = build_demo()
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present: the no-device behaviour is checked on the CPU box")
    except ImportError:
        pass
    r = subprocess.run([demo, "glass_cube", "1", "--hp"], capture_output=True, text=True)
    assert r.returncode == 2 and "no CPU fallback" in r.stderr, (r.returncode, r.stderr)


@pytest.mark.gpu
def test_host_appscene_renders_what_the_abi_renders(ctx, tmp_path):
    demo = build_demo()
    out = tmp_path / "frame.ppm"
    frames = 3
    r = subprocess.run([demo, "glass_cube", str(frames), str(out), "--hp", "--color-only"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "400x400" in r.stdout and "chunk 32^3" in r.stdout, r.stdout
    raw = open(out, "rb").read()
    hdr = b"P6\n400 400\n255\n"
    assert raw.startswith(hdr)
    img = np.frombuffer(raw[len(hdr):], np.uint8).reshape(400, 400, 3)[::-1]          # back to GL row order
    # the same frame through the ctypes binding: _GLASS_CUBE + _HIGH_PERFORMANCE + _COLOR_ONLY (BASELINE configs[0])
    grid = vrt.scene_fill("glass_cube", 32)
    ctx.upload_grid(grid, 32)
    ctx.upload_atlas(vrt.atlas_synthetic(16), 16)
    p = vrt.default_params(400, 400, 32)
    m = vrt.camera_invpv((-3.45, 2.17, 3.53), (-33 * DEG, -48 * DEG, 0.0), np.float32(1440) / np.float32(810))
    for i in range(16):
        p.invPV[i] = float(m[i])
    tod = np.float32(0)
    for _ in range(frames - 1):
        tod = np.float32(tod + np.float32(1.0 / 60.0))                                  # AppScene::Update before the last Render
    s = vrt.sun_dir(float(tod))
    for i in range(3):
        p.sunDir[i] = float(s[i])
    p.atlasSize, p.atlasTextureSize, p.time, p.flags = 32, 16, float(frames), vrt.FLAG_COLOR_ONLY
    ctx.render(p)
    want = ctx.read_framebuffer(p).reshape(400, 400, 4)[:, :, :3]
    assert np.array_equal(img, want)                                                    # alpha = 1: the temporal pass is a copy
