"""Shared test helpers: builds + ctypes wrappers for the checkers (oracle, oracle/_ref, hostsim) and the
named scenes/poses of SURVEY.md 8(d).  Only tests/, bench.py's CPU legs and __graft_entry__.smoke() use this."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import voxelraytracer_b200 as vrt  # noqa: E402

DEG = np.pi / 180.0
NCPU = os.cpu_count() or 1


def _run(cmd, cwd=None):
    r = subprocess.run(cmd, cwd=cwd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("command failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))
    return r.stdout


def _newer(target, deps):
    return os.path.exists(target) and all(os.path.getmtime(target) >= os.path.getmtime(d) for d in deps if os.path.exists(d))


def build_oracle():
    """oracle/liboracle.so (always) and oracle/_ref/libvoxel_ref.so (only where /root/reference exists)."""
    od = os.path.join(ROOT, "oracle")
    so = os.path.join(od, "liboracle.so")
    if not _newer(so, [os.path.join(od, "voxel_oracle.cpp"), os.path.join(ROOT, "include", "vrt.h")]):
        _run(["make", "-C", od, "liboracle.so"])
    _run([sys.executable, os.path.join(od, "build_ref.py")])
    return so


def build_hostsim():
    hd = os.path.join(ROOT, "tests", "hostsim")
    so = os.path.join(hd, "libhostsim.so")
    cs = os.path.join(ROOT, "voxelraytracer_b200", "csrc")
    deps = [os.path.join(hd, "hostsim.cpp")] + [os.path.join(cs, f) for f in
                                                ("rt_common.cuh", "rt_shade.cuh", "rt_trace_exact.cuh", "rt_trace_fast.cuh", "host_scene.cpp")]
    if not _newer(so, deps):
        _run(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-shared", "-pthread",
              "-Wno-unknown-pragmas", os.path.join(hd, "hostsim.cpp"), os.path.join(cs, "host_scene.cpp"), "-o", so])
    return so


_libs = {}


def oracle_lib():
    if "o" not in _libs:
        _libs["o"] = C.CDLL(build_oracle())
        _libs["o"].voracle_hash.restype = C.c_uint32
        _libs["o"].voracle_hash.argtypes = [C.c_uint32]
        _libs["o"].voracle_random.restype = C.c_float
        _libs["o"].voracle_random.argtypes = [C.c_float] * 4
        _libs["o"].voracle_quant.restype = C.c_uint8
        _libs["o"].voracle_quant.argtypes = [C.c_float]
    return _libs["o"]


def ref_lib():
    """The reference's own shader compiled as C++ (None when it was never built here)."""
    if "r" not in _libs:
        build_oracle()
        p = os.path.join(ROOT, "oracle", "_ref", "libvoxel_ref.so")
        _libs["r"] = C.CDLL(p) if os.path.exists(p) else None
    return _libs["r"]


def hostsim_lib():
    if "h" not in _libs:
        _libs["h"] = C.CDLL(build_hostsim())
    return _libs["h"]


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class OracleCounters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("primary_rays", "secondary_rays", "shadow_rays", "march_steps", "shadow_steps", "hits",
                                          "tie_index3", "fetch3d", "fetch2d")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


class Frame:
    """Result bundle of any renderer: f32 (n,4), rgba8 (n,4), hit (n,), bounces (n,), counters dict."""

    def __init__(self, **kw):
        self.__dict__.update(kw)


def oracle_render(p, grid, atlas, nthreads=NCPU, row_step=1, want_steps=False):
    n = vrt.tile_pixels(p)
    f32 = np.zeros((n, 4), np.float32)
    q = np.zeros((n, 4), np.uint8)
    hit = np.zeros(n, np.uint32)
    nb = np.zeros(n, np.uint8)
    steps = np.zeros(n, np.uint32) if want_steps else None
    f3 = np.zeros(n, np.uint32)
    f2 = np.zeros(n, np.uint32)
    cnt = OracleCounters()
    rc = oracle_lib().voracle_render(C.byref(p), _p(grid), _p(atlas), _p(f32), _p(q), _p(hit), _p(nb), _p(f3), _p(f2), _p(steps),
                                     C.byref(cnt), nthreads, row_step)
    assert rc == 0, "voracle_render failed"
    return Frame(f32=f32, rgba8=q, hit=hit, bounces=nb, steps=steps, fetch3d=f3, fetch2d=f2, counters=cnt.as_dict())


def ref_render(p, grid, atlas, nthreads=NCPU):
    lib = ref_lib()
    assert lib is not None
    n = p.tile_w * p.tile_h
    f32 = np.zeros((n, 4), np.float32)
    f3 = np.zeros(n, np.uint32)
    f2 = np.zeros(n, np.uint32)
    rc = lib.vref_render(C.byref(p), _p(grid), _p(atlas), _p(f32), _p(f3), _p(f2), nthreads)
    assert rc == 0, "vref_render failed"
    return Frame(f32=f32, fetch3d=f3, fetch2d=f2)


_cells_cache = {}


OCTANTS = int(os.environ.get("VRT_TEST_OCTANTS", "1"))   # what the product builds by default (vrt_set_option can change it)


def radius_cap(size):
    return 255 if size <= 256 else 96


def hostsim_cells(grid, size, octants=None):
    octants = OCTANTS if octants is None else octants
    key = (hash(grid.tobytes()), size, octants)
    if key not in _cells_cache:
        if octants:
            cells = np.zeros(8 * (size + 2) ** 3, np.uint16)
            assert hostsim_lib().hostsim_build_cells_oct(_p(grid), size, radius_cap(size), _p(cells)) == 0
        else:
            cells = np.zeros((size + 2) ** 3, np.uint16)
            assert hostsim_lib().hostsim_build_cells(_p(grid), size, _p(cells)) == 0
        _cells_cache[key] = cells
    return _cells_cache[key]


def hostsim_render(p, grid, atlas, mode, nthreads=NCPU, octants=None):
    """mode 0 = exact tracer, 1 = fast tracer (the product's device headers compiled for the host)."""
    n = vrt.tile_pixels(p)
    f32 = np.zeros((n, 4), np.float32)
    q = np.zeros((n, 4), np.uint8)
    hit = np.zeros(n, np.uint32)
    nb = np.zeros(n, np.uint8)
    cnt = (C.c_uint64 * 8)()
    octants = OCTANTS if octants is None else octants
    cells = hostsim_cells(grid, p.size, octants)
    rc = hostsim_lib().hostsim_render(C.byref(p), _p(grid), _p(cells), octants, _p(atlas), mode, _p(f32), _p(q), _p(hit), _p(nb), cnt, nthreads)
    assert rc == 0, "hostsim_render failed"
    names = ("primary_rays", "secondary_rays", "shadow_rays", "march_steps", "shadow_steps", "actual_steps", "hits", "tie_index3")
    return Frame(f32=f32, rgba8=q, hit=hit, bounces=nb, counters=dict(zip(names, [int(x) for x in cnt])))


def gpu_render(ctx, p, flags_extra=vrt.FLAG_DEBUG_MAPS | vrt.FLAG_FLOAT_OUT | vrt.FLAG_COUNTERS):
    """Through the C ABI.  Returns a Frame like the oracle's."""
    q = p.copy()
    q.flags = p.flags | flags_extra
    if flags_extra & vrt.FLAG_COUNTERS:
        ctx.reset_counters()
    ctx.render(q)
    rgba8 = ctx.read_framebuffer(q)
    hit, nb = ctx.read_debug_maps(q) if flags_extra & vrt.FLAG_DEBUG_MAPS else (None, None)
    f32 = ctx.read_float_frame(q) if flags_extra & vrt.FLAG_FLOAT_OUT else None
    cnt = ctx.counters() if flags_extra & vrt.FLAG_COUNTERS else None
    return Frame(f32=f32, rgba8=rgba8, hit=hit, bounces=nb, counters=cnt)


# ---------------------------------------------------------------- named scenes
_scene_cache = {}


_atlas_cache = {}


def ref_atlas(ts):
    """The reference's own tiles (res/textures/{stone,dirt,glass,grass}[128].png, decoded once by tests/golden/make_tiles.py into
    tests/golden/ref_tiles.npz) packed in AddTexture order (main.cpp:179-182,189-192) through vrt_atlas_pack.  ts = 16 or 128."""
    if ts not in _atlas_cache:
        t = np.load(os.path.join(ROOT, "tests", "golden", "ref_tiles.npz"))
        _atlas_cache[ts] = vrt.atlas_pack(*[t["%s%d" % (n, ts)] for n in ("stone", "dirt", "glass", "grass")])
    return _atlas_cache[ts]


def scene(name):
    """Returns (grid uint8[n^3], n, atlas uint8[as,as,4], texture_size).  C1-C4 scenes carry the reference's texture tiles; the
    synthetic C5 family keeps the procedural atlas."""
    if name in _scene_cache:
        return _scene_cache[name]
    if name == "glass_cube32":       # C1 (main.cpp:259-271), _HIGH_PERFORMANCE sizes (main.cpp:177,184)
        r = (vrt.scene_fill("glass_cube", 32), 32, ref_atlas(16), 16)
    elif name == "terrain32":        # _TERRAIN + _HIGH_PERFORMANCE: has the glass walls (size <= 64)
        r = (vrt.scene_fill("terrain", 32), 32, ref_atlas(16), 16)
    elif name == "terrain128":       # C2
        r = (vrt.scene_fill("terrain", 128), 128, ref_atlas(128), 128)
    elif name == "terrain128_glass":  # C4: guard at main.cpp:233 lifted
        r = (vrt.scene_fill("terrain", 128, glass_walls=True), 128, ref_atlas(128), 128)
    elif name == "refraction128":    # C3 (main.cpp:272-288)
        r = (vrt.scene_fill("refraction", 128), 128, ref_atlas(128), 128)
    elif name == "refraction32":
        r = (vrt.scene_fill("refraction", 32), 32, ref_atlas(16), 16)
    elif name.startswith("synthetic"):  # C5 family: synthetic256 / synthetic512 / synthetic1024
        sz = int(name[len("synthetic"):])
        r = (vrt.scene_fill("synthetic", sz, seed=7), sz, vrt.atlas_synthetic(128), 128)
    else:
        raise KeyError(name)
    _scene_cache[name] = r
    return r


# recorded poses: (position relative to the chunk centre, rotation in radians)
POSES = {
    "start": ((-3.45, 2.17, 3.53), (-33 * DEG, -48 * DEG, 0.0)),        # main.cpp:171-172 read as degrees
    "terrain_inside": ((-30.0, 30.0, 40.0), (-33 * DEG, -48 * DEG, 0.0)),
    "terrain_low": ((-50.0, -5.0, 50.0), (-10 * DEG, -45 * DEG, 0.0)),
    "terrain_above": ((0.0, 70.0, 0.0), (-89 * DEG, 0.0, 0.0)),
    "terrain_outside": ((-80.0, 50.0, 90.0), (-25 * DEG, -42 * DEG, 0.0)),
    "terrain32_above": ((-3.45, 12.17, 3.53), (-33 * DEG, -48 * DEG, 0.0)),
    "refraction_close": ((-1.5, 1.2, 3.0), (-15 * DEG, -30 * DEG, 0.0)),
    "glass_wall": ((-58.0, 20.0, 10.0), (-5 * DEG, 80 * DEG, 0.0)),
}


def orbit_pose(k, count=64, radius=600.0, height=250.0):
    """Pose k of `count` on a circle around the chunk centre, looking at the centre."""
    a = 2.0 * np.pi * k / count
    pos = np.array([radius * np.cos(a), height, radius * np.sin(a)])
    f = -pos / np.linalg.norm(pos)
    # forward = Ry(ry) Rx(rx) (0,0,-1) = (-cos rx sin ry, sin rx, -cos rx cos ry)  (camera_invpv's view = Rx(-rx) Ry(-ry) T(-pos))
    return (tuple(float(x) for x in pos), (float(np.arcsin(f[1])), float(np.arctan2(-f[0], -f[2])), 0.0))


def make_params(width, height, scene_name, pose, flags=0, aspect=None, noise=(0.0, 0.0, 0.0), time=1.0, time_of_day=0.0, max_len=None):
    grid, n, atlas, ts = scene(scene_name)
    p = vrt.default_params(width, height, n)
    pos, rot = POSES[pose] if isinstance(pose, str) else pose
    m = vrt.camera_invpv(pos, rot, aspect if aspect else width / height)
    for i in range(16):
        p.invPV[i] = float(m[i])
    s = vrt.sun_dir(time_of_day)
    for i in range(3):
        p.sunDir[i] = float(s[i])
    p.atlasSize = 2 * ts
    p.atlasTextureSize = ts
    p.flags = flags
    p.rayNoise, p.reflectionNoise, p.refractionNoise = noise
    p.time = time
    if max_len is not None:
        p.maxRayLength = float(max_len)
    return p


def rms255(a8, b8):
    """Per-channel RMS of the 8-bit frames, in 1/255 units."""
    d = a8[:, :3].astype(np.float64) - b8[:, :3].astype(np.float64)
    return np.sqrt((d ** 2).mean(axis=0))


def write_ppm(path, rgba8, width, height):
    """GL row order (row 0 = bottom) -> top-down PPM."""
    img = rgba8.reshape(height, width, 4)[::-1, :, :3]
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (width, height))
        f.write(np.ascontiguousarray(img).tobytes())
