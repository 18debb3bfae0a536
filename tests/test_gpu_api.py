"""C-ABI behaviour on the GPU: error codes, external output buffers, host round trip, temporal pass, timing."""
import ctypes as C

import numpy as np
import pytest

import voxelraytracer_b200 as vrt
from common import gpu_render, make_params, scene

pytestmark = pytest.mark.gpu


def test_errors_are_codes_not_aborts():
    c = vrt.Context(0)
    p = vrt.default_params(64, 64, 32)
    with pytest.raises(vrt.VrtError) as e:
        c.render(p)
    assert e.value.code == -4                                   # VRT_E_NOT_READY: no grid
    with pytest.raises(vrt.VrtError) as e:
        c.upload_grid(np.zeros(24 ** 3, np.uint8), 24)
    assert e.value.code == -1                                   # size must be a power of two
    grid, n, atlas, ts = scene("glass_cube32")
    c.upload_grid(grid, n)
    with pytest.raises(vrt.VrtError) as e:
        c.render(p)                                             # textured without an atlas
    assert e.value.code == -4
    p.flags = vrt.FLAG_COLOR_ONLY
    c.render(p)
    p.tile_w = 65
    with pytest.raises(vrt.VrtError) as e:
        c.render(p)
    assert e.value.code == -1
    p.tile_w = 64; p.size = 64
    with pytest.raises(vrt.VrtError):
        c.render(p)
    p.size = 32; p.maxReflections = 9
    with pytest.raises(vrt.VrtError):
        c.render(p)
    with pytest.raises(vrt.VrtError):
        vrt.Context(4096)
    c.close()


def test_render_to_torch_tensor_and_render_host(ctx):
    import torch
    grid, n, atlas, ts = scene("terrain32")
    ctx.upload_grid(grid, n); ctx.upload_atlas(atlas, ts)
    W, H = 320, 200
    p = make_params(W, H, "terrain32", "terrain32_above")
    ref = gpu_render(ctx, p, flags_extra=0).rgba8
    t = torch.zeros((H, W, 4), dtype=torch.uint8, device="cuda:0")
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        ctx.render_to(p, t.data_ptr(), s.cuda_stream)
    s.synchronize()
    assert np.array_equal(t.cpu().numpy().reshape(-1, 4), ref)
    host, hp = ctx.host_alloc(W * H * 4)
    ctx.render_host(p, hp)                                     # pinned: the kernel stores straight into host memory
    assert np.array_equal(host.reshape(-1, 4), ref)
    assert ctx.last_render_ms() > 0.0
    host[:] = 0
    ctx.set_option(vrt.OPT_HOST_DIRECT, 0)                     # device frame + copy
    ctx.render_host(p, hp)
    ctx.set_option(vrt.OPT_HOST_DIRECT, 1)
    assert np.array_equal(host.reshape(-1, 4), ref)
    pageable = np.zeros(W * H * 4, np.uint8)                   # not pinned: falls back to the copy by itself
    ctx.render_host(p, pageable.ctypes.data_as(C.c_void_p))
    assert np.array_equal(pageable.reshape(-1, 4), ref)
    # ragged frame straight into pinned memory through vrt_render_to (row-store arrangement, partial patches)
    W2, H2 = 333, 77
    p2 = make_params(W2, H2, "terrain32", "terrain32_above")
    ref2 = gpu_render(ctx, p2, flags_extra=0).rgba8
    host2, hp2 = ctx.host_alloc(W2 * H2 * 4 + 64)
    host2[:] = 0xAB
    ctx.render_to(p2, hp2); ctx.synchronize()
    assert np.array_equal(host2[:W2 * H2 * 4].reshape(-1, 4), ref2)
    assert (host2[W2 * H2 * 4:] == 0xAB).all()                 # nothing written past the frame
    before = ctx.kernel_launches()
    ctx.render(p); ctx.synchronize()
    assert ctx.kernel_launches() == before + 1


def test_temporal_blend_matches_formula(ctx):
    import torch
    rng = np.random.default_rng(0)
    n = 1920 * 1080 + 3
    a = rng.integers(0, 256, (n, 4), dtype=np.uint8)
    b = rng.integers(0, 256, (n, 4), dtype=np.uint8)
    ta, tb = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    for alpha in (1.0, 0.25):                                  # alpha = 1 is the reference default (main.cpp:155)
        acc = tb.clone()
        ctx.temporal_blend(ta.data_ptr(), acc.data_ptr(), n, alpha)
        ctx.synchronize()
        got = acc.cpu().numpy()
        al = np.float32(alpha)
        v = al * (a.astype(np.float32) / np.float32(255)) + (np.float32(1) - al) * (b.astype(np.float32) / np.float32(255))  # temporal.glsl:18
        exp = np.floor(np.clip(v, 0, 1) * np.float32(255) + np.float32(0.5)).astype(np.uint8)
        assert np.array_equal(got, exp)
    assert np.array_equal(got if alpha == 1.0 else a, a)


def test_own_framebuffer_is_not_confused_with_caller_targets(ctx):
    """vrt_framebuffer_device / vrt_read_framebuffer describe the ctx-owned frame (FrameBuffer::GetTexture, FrameBuffer.cpp:27-30)
    even after a larger frame went to caller memory through vrt_render_to."""
    grid, n, atlas, ts = scene("terrain32")
    ctx.upload_grid(grid, n); ctx.upload_atlas(atlas, ts)
    small = make_params(64, 48, "terrain32", "terrain32_above")
    ctx.render(small)
    want = ctx.read_framebuffer(small)
    big = make_params(640, 400, "terrain32", "terrain32_above")
    d = ctx.device_alloc(640 * 400 * 4)
    ctx.render_to(big, d)
    ctx.synchronize()
    _, nbytes = ctx.framebuffer_device()
    assert nbytes == 64 * 48 * 4
    assert np.array_equal(ctx.read_framebuffer(small), want)
    ctx.device_free(d)


def test_two_contexts_are_independent():
    c1, c2 = vrt.Context(0), vrt.Context(0)
    g1, n1, a1, t1 = scene("glass_cube32")
    g2, n2, a2, t2 = scene("refraction32")
    c1.upload_grid(g1, n1); c2.upload_grid(g2, n2)
    p1 = make_params(128, 128, "glass_cube32", "start", vrt.FLAG_COLOR_ONLY)
    p2 = make_params(128, 128, "refraction32", "start", vrt.FLAG_COLOR_ONLY)
    f1 = gpu_render(c1, p1, flags_extra=0).rgba8
    f2 = gpu_render(c2, p2, flags_extra=0).rgba8
    assert not np.array_equal(f1, f2)
    assert np.array_equal(gpu_render(c1, p1, flags_extra=0).rgba8, f1)
    c1.close(); c2.close()


def test_non_finite_uniforms_are_refused(ctx):
    grid, n, atlas, ts = scene("glass_cube32")
    ctx.upload_grid(grid, n); ctx.upload_atlas(atlas, ts)
    for field in ("invPV", "sunDir", "time", "maxRayLength"):
        p = make_params(64, 64, "glass_cube32", "start", vrt.FLAG_COLOR_ONLY)
        if field == "invPV":
            p.invPV[5] = float("nan")
        elif field == "sunDir":
            p.sunDir[1] = float("inf")
        else:
            setattr(p, field, float("nan"))
        with pytest.raises(vrt.VrtError) as e:
            ctx.render(p)
        assert e.value.code == -1, field          # VRT_E_INVALID, and the context stays usable
    assert gpu_render(ctx, make_params(64, 64, "glass_cube32", "start", vrt.FLAG_COLOR_ONLY), flags_extra=0).rgba8.shape == (4096, 4)


def test_async_host_copy_and_radius_cap_option(ctx):
    """vrt_copy_to_host_async (double-buffered host read-back) and VRT_OPT_RADIUS_CAP (smaller cap = same pixels)."""
    grid, n, atlas, ts = scene("glass_cube32")
    ctx.upload_grid(grid, n); ctx.upload_atlas(atlas, ts)
    p = make_params(200, 120, "glass_cube32", "start", vrt.FLAG_COLOR_ONLY)
    want = gpu_render(ctx, p, flags_extra=0).rgba8
    d = ctx.device_alloc(200 * 120 * 4)
    host, hp = ctx.host_alloc(200 * 120 * 4)
    host[:] = 0
    ctx.render_to(p, d, None)
    ctx.copy_to_host_async(hp, d, 200 * 120 * 4, None)
    ctx.synchronize()
    assert np.array_equal(host.reshape(-1, 4), want)
    ctx.device_free(d)
    try:
        ctx.set_option(vrt.OPT_RADIUS_CAP, 3)
        ctx.upload_grid(grid, n)
        assert ctx.grid_info().radius_cap == 3 and int((ctx.read_cells() >> 8).max()) == 3
        got = gpu_render(ctx, p, flags_extra=0).rgba8                              # shorter jumps, same cells visited:
        assert int((got != want).any(axis=1).sum()) <= 2                           # only a grazing ray may land differently
        with pytest.raises(vrt.VrtError):
            ctx.set_option(vrt.OPT_RADIUS_CAP, 999)
    finally:
        ctx.set_option(vrt.OPT_RADIUS_CAP, 0)
        ctx.upload_grid(grid, n)


def test_read_bandwidth_probe(ctx):
    """vrt_probe_read_gbs: the L2-resident sweep is faster than the HBM-sized one, both in a sane range for a B200."""
    l2 = ctx.probe_read_gbs(32 << 20, 50)
    hbm = ctx.probe_read_gbs(1 << 30, 2)
    assert 1000.0 < hbm < 9000.0
    assert hbm < l2 < 40000.0
    with pytest.raises(vrt.VrtError):
        ctx.probe_read_gbs(1024, 1)


def test_signal_add_and_stream_wait_value(ctx):
    """The completion flag of the fused gather, on one GPU.  (1) vrt_signal_add adds at system scope; a wait for a value already reached
    lets the stream through.  (2) A wait on a counter that has NOT been reached blocks the stream's later work until the counter gets
    there - released here by a HOST write to a pinned counter, not by a kernel on another stream: two streams may share a hardware queue,
    and a kernel queued behind a blocked wait would never run (the ordering rule in include/vrt.h)."""
    import time
    import torch
    d = ctx.device_alloc(256)
    try:
        ctx.device_memset(d, 0, 256)
        ctx.synchronize()
        st = torch.cuda.Stream()
        for _ in range(3):
            ctx.signal_add(d, 1, st.cuda_stream)
        ctx.stream_wait_value(d, 3, st.cuda_stream)          # enqueued AFTER the adds it waits for
        st.synchronize()
        assert int(ctx.copy_to_host(d, 4).view(np.uint32)[0]) == 3
        with pytest.raises(vrt.VrtError):
            ctx.signal_add(d.value + 2, 1)                   # misaligned
        host, hp = ctx.host_alloc(64)
        flag = host.view(np.uint32)
        flag[:] = 0
        waiter = torch.cuda.Stream()
        ctx.stream_wait_value(hp, 5, waiter.cuda_stream)
        ctx.signal_add(d, 10, waiter.cuda_stream)            # behind the wait
        time.sleep(0.05)
        assert not waiter.query()                            # blocked: the counter is 0
        flag[0] = 5                                          # the host releases it
        waiter.synchronize()
        assert int(ctx.copy_to_host(d, 4).view(np.uint32)[0]) == 13
    finally:
        ctx.device_free(d)
