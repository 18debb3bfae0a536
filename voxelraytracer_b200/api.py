"""ctypes binding of include/vrt.h.  Mirrors the C ABI one to one; no arithmetic happens here."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

FLAG_COLOR_ONLY = 0x1
FLAG_DEBUG_MAPS = 0x2
FLAG_EXACT = 0x4
FLAG_COUNTERS = 0x8
FLAG_FLOAT_OUT = 0x10
FLAG_OUT_FRAME = 0x20
FLAG_PERSISTENT = 0x40
FLAG_RGB8_OUT = 0x80
FLAG_CAREFUL = 0x100
FLAG_RAY_QUEUE = 0x200
OPT_OCTANT_FIELDS = 1
OPT_RADIUS_CAP = 2
OPT_HOST_DIRECT = 3
OPT_SUN_FIELD = 4
OPT_REMOTE_ROWS = 5
REPACK_STATIC_MAX_SIZE = 128
HIT_MISS = 0xFFFFFFFF

SCENE_TERRAIN, SCENE_GLASS_CUBE, SCENE_REFRACTION, SCENE_SYNTHETIC = 0, 1, 2, 3

# every symbol include/vrt.h declares (tests check that the library exports all of them)
EXPORTS = [
    "vrt_default_params", "vrt_abi_version", "vrt_create", "vrt_destroy", "vrt_last_error", "vrt_last_warning", "vrt_device_memset", "vrt_upload_grid",
    "vrt_upload_grid_device", "vrt_upload_atlas", "vrt_render", "vrt_render_to", "vrt_render_host", "vrt_tile_pixels", "vrt_probe_read_gbs",
    "vrt_framebuffer_device", "vrt_read_framebuffer", "vrt_read_debug_maps", "vrt_read_float_frame", "vrt_reset_counters",
    "vrt_get_counters", "vrt_last_render_ms", "vrt_kernel_launches", "vrt_grid_info_get", "vrt_read_cells", "vrt_read_sun_cells", "vrt_repack_static", "vrt_set_option",
    "vrt_host_alloc", "vrt_host_free", "vrt_device_alloc", "vrt_device_free", "vrt_synchronize", "vrt_copy_to_host", "vrt_copy_to_host_async", "vrt_copy_2d_async", "vrt_copy_to_device", "vrt_temporal_blend", "vrt_temporal_filter",
    "vrt_ipc_alloc_frame", "vrt_ipc_open", "vrt_ipc_close", "vrt_signal_add", "vrt_stream_wait_value", "vrt_scene_fill", "vrt_noise_heightmap", "vrt_atlas_synthetic",
    "vrt_atlas_pack", "vrt_camera_invpv", "vrt_sun_dir", "vrt_write_png",
]


class VrtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("vrt error %d: %s" % (code, msg))
        self.code = code


class Params(C.Structure):
    """struct vrt_params (include/vrt.h)."""
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32),
        ("tile_x0", C.c_int32), ("tile_y0", C.c_int32), ("tile_w", C.c_int32), ("tile_h", C.c_int32),
        ("band_rows", C.c_int32), ("band_count", C.c_int32), ("band_index", C.c_int32),
        ("invPV", C.c_float * 16),
        ("size", C.c_int32), ("atlasSize", C.c_int32), ("atlasTextureSize", C.c_int32),
        ("sunDir", C.c_float * 3),
        ("time", C.c_float), ("rayNoise", C.c_float), ("reflectionNoise", C.c_float), ("refractionNoise", C.c_float),
        ("maxRayLength", C.c_float),
        ("maxReflections", C.c_int32), ("maxTransparencies", C.c_int32),
        ("flags", C.c_uint32),
    ]

    def copy(self):
        q = Params()
        C.memmove(C.byref(q), C.byref(self), C.sizeof(Params))
        return q


class Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("primary_rays", "secondary_rays", "shadow_rays", "march_steps", "shadow_steps",
                                          "actual_steps", "hits", "tie_index3")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


class GridInfo(C.Structure):
    _fields_ = [("size", C.c_int32), ("pitch", C.c_int32), ("radius_cap", C.c_int32), ("octants", C.c_int32),
                ("cells_total", C.c_int64), ("voxels_occupied", C.c_int64), ("voxels_glass", C.c_int64)]


def lib_path():
    """In-tree libvrt.so; VRT_LIBRARY names another BUILD of it (tools/ab_time.py times kernel variants side by side)."""
    return os.environ.get("VRT_LIBRARY") or os.path.join(_HERE, "libvrt.so")


_lib = None


def load_library():
    """Loads libvrt.so (built in-tree).  Raises ImportError if it is missing: there is no Python/CPU fallback."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise ImportError("libvrt.so is not built (%s missing); run `make -C voxelraytracer_b200/csrc` or "
                          "__graft_entry__.build().  There is no CPU fallback." % path)
    lib = C.CDLL(path)
    for name, (res, args) in _signatures().items():
        try:
            fn = getattr(lib, name)  # AttributeError here == header/library mismatch
        except AttributeError:
            if os.environ.get("VRT_LIBRARY"):
                continue             # an older BUILD timed side by side (tools/ab_time.py) may lack newer entry points
            raise
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def _signatures():
    vp, i32, i64, u32, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint32, C.c_float
    PP = C.POINTER(Params)
    return {
        "vrt_default_params": (None, [PP, i32, i32, i32]),
        "vrt_abi_version": (i32, []),
        "vrt_create": (i32, [i32, C.POINTER(vp)]),
        "vrt_destroy": (i32, [vp]),
        "vrt_last_error": (C.c_char_p, [vp]),
        "vrt_last_warning": (C.c_char_p, [vp]),
        "vrt_device_memset": (i32, [vp, vp, i32, C.c_size_t, vp]),
        "vrt_upload_grid": (i32, [vp, vp, i32]),
        "vrt_upload_grid_device": (i32, [vp, vp, i32, vp]),
        "vrt_upload_atlas": (i32, [vp, vp, i32, i32]),
        "vrt_render": (i32, [vp, PP, vp]),
        "vrt_render_to": (i32, [vp, PP, vp, vp]),
        "vrt_render_host": (i32, [vp, PP, vp, vp]),
        "vrt_probe_read_gbs": (i32, [vp, C.c_size_t, i32, C.POINTER(C.c_float)]),
        "vrt_tile_pixels": (i64, [PP]),
        "vrt_framebuffer_device": (i32, [vp, C.POINTER(vp), C.POINTER(C.c_size_t)]),
        "vrt_read_framebuffer": (i32, [vp, vp, C.c_size_t]),
        "vrt_read_debug_maps": (i32, [vp, vp, vp, C.c_size_t]),
        "vrt_read_float_frame": (i32, [vp, vp, C.c_size_t]),
        "vrt_reset_counters": (i32, [vp]),
        "vrt_get_counters": (i32, [vp, C.POINTER(Counters)]),
        "vrt_last_render_ms": (i32, [vp, C.POINTER(f32)]),
        "vrt_kernel_launches": (i64, [vp]),
        "vrt_grid_info_get": (i32, [vp, C.POINTER(GridInfo)]),
        "vrt_read_cells": (i32, [vp, vp, C.c_size_t]),
        "vrt_read_sun_cells": (i32, [vp, vp, C.c_size_t]),
        "vrt_repack_static": (i32, [vp, vp, i32, vp, vp]),
        "vrt_set_option": (i32, [vp, i32, i32]),
        "vrt_host_alloc": (i32, [vp, C.c_size_t, C.POINTER(vp)]),
        "vrt_host_free": (i32, [vp, vp]),
        "vrt_device_alloc": (i32, [vp, C.c_size_t, C.POINTER(vp)]),
        "vrt_device_free": (i32, [vp, vp]),
        "vrt_synchronize": (i32, [vp, vp]),
        "vrt_copy_to_host": (i32, [vp, vp, vp, C.c_size_t, vp]),
        "vrt_copy_to_device": (i32, [vp, vp, vp, C.c_size_t, vp]),
        "vrt_copy_to_host_async": (i32, [vp, vp, vp, C.c_size_t, vp]),
        "vrt_copy_2d_async": (i32, [vp, vp, C.c_size_t, vp, C.c_size_t, C.c_size_t, C.c_size_t, vp]),
        "vrt_temporal_blend": (i32, [vp, vp, vp, i64, f32, vp]),
        "vrt_temporal_filter": (i32, [vp, vp, vp, vp, i64, f32, vp]),
        "vrt_ipc_alloc_frame": (i32, [vp, C.c_size_t, C.POINTER(vp), vp]),
        "vrt_ipc_open": (i32, [vp, vp, C.POINTER(vp)]),
        "vrt_ipc_close": (i32, [vp, vp]),
        "vrt_signal_add": (i32, [vp, vp, u32, vp]),
        "vrt_stream_wait_value": (i32, [vp, vp, u32, vp]),
        "vrt_scene_fill": (i32, [i32, i32, vp, i32, u32, vp]),
        "vrt_noise_heightmap": (i32, [i32, u32, i32, f32, vp]),
        "vrt_atlas_synthetic": (i32, [i32, u32, vp]),
        "vrt_atlas_pack": (i32, [vp, vp, vp, vp, i32, vp]),
        "vrt_camera_invpv": (i32, [C.POINTER(f32), C.POINTER(f32), f32, f32, f32, f32, C.POINTER(f32)]),
        "vrt_sun_dir": (i32, [f32, f32, C.POINTER(f32)]),
        "vrt_write_png": (i32, [C.c_char_p, vp, i32, i32, i32]),
    }


# The CUDA-free part of the ABI (host_scene.cpp), also built as libvrt_host.so: scenes, atlas, camera, sun, PNG, default params.
HOST_EXPORTS = ["vrt_default_params", "vrt_abi_version", "vrt_tile_pixels", "vrt_scene_fill", "vrt_noise_heightmap", "vrt_atlas_synthetic",
                "vrt_atlas_pack", "vrt_camera_invpv", "vrt_sun_dir", "vrt_write_png"]
_host_lib = None


def load_host_library():
    """libvrt_host.so: the host-side helpers only, no CUDA code mapped (CPU-only tooling: oracle drivers, bench.py --impl reference).
    Same functions, same object file as the ones libvrt.so exports."""
    global _host_lib
    if _host_lib is not None:
        return _host_lib
    path = os.path.join(_HERE, "libvrt_host.so")          # always the in-tree one (VRT_LIBRARY only swaps the device library)
    if not os.path.exists(path):
        raise ImportError("libvrt_host.so is not built (%s missing); run `make -C voxelraytracer_b200/csrc`" % path)
    lib = C.CDLL(path)
    full = _signatures()
    for name in HOST_EXPORTS:
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = full[name]
    _host_lib = lib
    return lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _host_check(rc, what):
    if rc != 0:
        raise VrtError(rc, "%s failed (bad argument)" % what)


# ---------------------------------------------------------------- host helpers (no GPU needed)
def default_params(width, height, size):
    p = Params()
    load_host_library().vrt_default_params(C.byref(p), width, height, size)
    return p


def tile_pixels(p):
    return int(load_host_library().vrt_tile_pixels(C.byref(p)))


def noise_heightmap(size, seed=1, octaves=5, persistence=0.125):
    out = np.zeros((size, size), np.float32)
    _host_check(load_host_library().vrt_noise_heightmap(size, seed, octaves, persistence, _ptr(out)), "vrt_noise_heightmap")
    return out


def scene_fill(kind, size, heightmap=None, glass_walls=False, seed=1):
    """Returns uint8[size**3], index x + y*size + z*size*size (src/main.cpp:219-288)."""
    kind = {"terrain": SCENE_TERRAIN, "glass_cube": SCENE_GLASS_CUBE, "refraction": SCENE_REFRACTION,
            "synthetic": SCENE_SYNTHETIC}.get(kind, kind)
    if kind == SCENE_TERRAIN and heightmap is None:
        heightmap = noise_heightmap(size, seed, 5, 0.5 if size <= 32 else 0.125)  # main.cpp:185,195
    out = np.zeros(size ** 3, np.uint8)
    hm = _ptr(np.ascontiguousarray(heightmap, np.float32)) if heightmap is not None else None
    _host_check(load_host_library().vrt_scene_fill(kind, size, hm, 1 if glass_walls else 0, seed, _ptr(out)), "vrt_scene_fill")
    return out


def atlas_synthetic(texture_size=128, seed=7):
    out = np.zeros((2 * texture_size, 2 * texture_size, 4), np.uint8)
    _host_check(load_host_library().vrt_atlas_synthetic(texture_size, seed, _ptr(out)), "vrt_atlas_synthetic")
    return out


def atlas_pack(stone, dirt, glass, grass):
    ts = stone.shape[0]
    tiles = [np.ascontiguousarray(t, np.uint8) for t in (stone, dirt, glass, grass)]
    for t in tiles:
        assert t.shape == (ts, ts, 4)
    out = np.zeros((2 * ts, 2 * ts, 4), np.uint8)
    _host_check(load_host_library().vrt_atlas_pack(_ptr(tiles[0]), _ptr(tiles[1]), _ptr(tiles[2]), _ptr(tiles[3]), ts, _ptr(out)),
                "vrt_atlas_pack")
    return out


def camera_invpv(pos, rot, aspect, fov_deg=90.0, z_near=0.01, z_far=100.0):
    """Cam (src/main.cpp:14-77): rot in radians; returns float32[16], column-major."""
    out = (C.c_float * 16)()
    _host_check(load_host_library().vrt_camera_invpv((C.c_float * 3)(*pos), (C.c_float * 3)(*rot), aspect, fov_deg, z_near, z_far, out),
                "vrt_camera_invpv")
    return np.array(out, np.float32)


def write_png(path, rgba8, width, height, gl_row_order=True):
    """RGBA8 frame (as the render calls return it) -> 8-bit RGB PNG."""
    a = np.ascontiguousarray(rgba8, np.uint8).reshape(height, width, 4)
    _host_check(load_host_library().vrt_write_png(str(path).encode(), _ptr(a), width, height, 1 if gl_row_order else 0), "vrt_write_png")


def sun_dir(time_of_day=0.0, day_time=50.0):
    out = (C.c_float * 3)()
    _host_check(load_host_library().vrt_sun_dir(time_of_day, day_time, out), "vrt_sun_dir")
    return np.array(out, np.float32)


# ---------------------------------------------------------------- device context
class Context:
    """vrt_ctx: one per GPU.  Mirrors AppScene's GL objects (chunk texture, atlas, FrameBuffer)."""

    def __init__(self, device=0):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.vrt_create(device, C.byref(h))
        if rc != 0:
            raise VrtError(rc, self.lib.vrt_last_error(None).decode())
        self.h = h
        self.device = device
        self._pinned = []

    def _check(self, rc):
        if rc != 0:
            raise VrtError(rc, self.lib.vrt_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            for p in self._pinned:
                self.lib.vrt_host_free(self.h, p)
            self._pinned = []
            self.lib.vrt_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # uploads
    def upload_grid(self, voxels, size):
        v = np.ascontiguousarray(voxels, np.uint8).ravel()
        if v.size != size ** 3:
            raise VrtError(-1, "grid has %d bytes, expected %d" % (v.size, size ** 3))
        self._check(self.lib.vrt_upload_grid(self.h, _ptr(v), size))

    def upload_grid_device(self, d_ptr, size, stream=None):
        self._check(self.lib.vrt_upload_grid_device(self.h, d_ptr, size, stream))

    def upload_atlas(self, rgba8, texture_size):
        a = np.ascontiguousarray(rgba8, np.uint8)
        self._check(self.lib.vrt_upload_atlas(self.h, _ptr(a), a.shape[0], texture_size))

    # rendering
    def render(self, p, stream=None):
        self._check(self.lib.vrt_render(self.h, C.byref(p), stream))

    def render_to(self, p, d_ptr, stream=None):
        self._check(self.lib.vrt_render_to(self.h, C.byref(p), d_ptr, stream))

    def probe_read_gbs(self, nbytes, passes):
        """Read bandwidth (GB/s) of an nbytes buffer read `passes` times by one kernel: L2 for 32 MiB, HBM for GiBs."""
        g = C.c_float()
        self._check(self.lib.vrt_probe_read_gbs(self.h, nbytes, passes, C.byref(g)))
        return float(g.value)

    def render_host(self, p, host_ptr, stream=None):
        self._check(self.lib.vrt_render_host(self.h, C.byref(p), host_ptr, stream))

    def read_framebuffer(self, p):
        n = tile_pixels(p)
        out = np.zeros((n, 3 if p.flags & FLAG_RGB8_OUT else 4), np.uint8)
        self._check(self.lib.vrt_read_framebuffer(self.h, _ptr(out), out.nbytes))
        return out

    def read_debug_maps(self, p):
        n = tile_pixels(p)
        hit = np.zeros(n, np.uint32)
        nb = np.zeros(n, np.uint8)
        self._check(self.lib.vrt_read_debug_maps(self.h, _ptr(hit), _ptr(nb), n))
        return hit, nb

    def read_float_frame(self, p):
        n = tile_pixels(p)
        out = np.zeros((n, 4), np.float32)
        self._check(self.lib.vrt_read_float_frame(self.h, _ptr(out), n))
        return out

    def framebuffer_device(self):
        d = C.c_void_p()
        n = C.c_size_t()
        self._check(self.lib.vrt_framebuffer_device(self.h, C.byref(d), C.byref(n)))
        return d.value, n.value

    def reset_counters(self):
        self._check(self.lib.vrt_reset_counters(self.h))

    def counters(self):
        c = Counters()
        self._check(self.lib.vrt_get_counters(self.h, C.byref(c)))
        return c.as_dict()

    def last_render_ms(self):
        ms = C.c_float()
        self._check(self.lib.vrt_last_render_ms(self.h, C.byref(ms)))
        return ms.value

    def kernel_launches(self):
        return int(self.lib.vrt_kernel_launches(self.h))

    def grid_info(self):
        g = GridInfo()
        self._check(self.lib.vrt_grid_info_get(self.h, C.byref(g)))
        return g

    def last_warning(self):
        return self.lib.vrt_last_warning(self.h).decode()

    def device_memset(self, d_ptr, value, nbytes, stream=None):
        self._check(self.lib.vrt_device_memset(self.h, d_ptr, value, nbytes, stream))

    def set_option(self, option, value):
        self._check(self.lib.vrt_set_option(self.h, option, value))

    def read_cells(self, size=None):
        out = np.zeros(self.grid_info().cells_total, np.uint16)
        self._check(self.lib.vrt_read_cells(self.h, _ptr(out), out.size))
        return out

    def repack_static(self, d_voxels, size, d_cells, stream=None):
        """vrt_repack_static: isotropic cells of a device-resident chunk into caller memory, no allocation (size <= 128)."""
        self._check(self.lib.vrt_repack_static(self.h, d_voxels, size, d_cells, stream))

    def read_sun_cells(self):
        out = np.zeros((self.grid_info().size + 2) ** 3, np.uint16)
        self._check(self.lib.vrt_read_sun_cells(self.h, _ptr(out), out.size))
        return out

    def synchronize(self, stream=None):
        self._check(self.lib.vrt_synchronize(self.h, stream))

    def temporal_blend(self, d_new, d_acc, npix, alpha, stream=None):
        self._check(self.lib.vrt_temporal_blend(self.h, d_new, d_acc, npix, alpha, stream))

    def temporal_filter(self, d_new, d_old, d_out, npix, alpha, stream=None):
        self._check(self.lib.vrt_temporal_filter(self.h, d_new, d_old, d_out, npix, alpha, stream))

    # memory helpers
    def host_alloc(self, nbytes):
        """Pinned host buffer as a numpy uint8 array (freed with the context)."""
        p = C.c_void_p()
        self._check(self.lib.vrt_host_alloc(self.h, nbytes, C.byref(p)))
        self._pinned.append(p)
        buf = (C.c_uint8 * nbytes).from_address(p.value)
        return np.frombuffer(buf, np.uint8), p

    def device_alloc(self, nbytes):
        p = C.c_void_p()
        self._check(self.lib.vrt_device_alloc(self.h, nbytes, C.byref(p)))
        return p

    def device_free(self, p):
        self._check(self.lib.vrt_device_free(self.h, p))

    def copy_to_host(self, d_ptr, nbytes, stream=None):
        """Device memory -> new numpy uint8 array."""
        out = np.empty(nbytes, np.uint8)
        self._check(self.lib.vrt_copy_to_host(self.h, _ptr(out), d_ptr, nbytes, stream))
        return out

    def copy_to_host_async(self, host_ptr, d_ptr, nbytes, stream=None):
        """Device -> (pinned) host, not synchronised; host_ptr from host_alloc()."""
        self._check(self.lib.vrt_copy_to_host_async(self.h, host_ptr, d_ptr, nbytes, stream))

    def copy_2d_async(self, d_dst, dst_pitch, d_src, src_pitch, width_bytes, height, stream=None):
        self._check(self.lib.vrt_copy_2d_async(self.h, d_dst, dst_pitch, d_src, src_pitch, width_bytes, height, stream))

    def copy_to_device(self, d_ptr, arr, stream=None):
        a = np.ascontiguousarray(arr)
        self._check(self.lib.vrt_copy_to_device(self.h, d_ptr, _ptr(a), a.nbytes, stream))

    def ipc_alloc_frame(self, nbytes):
        p = C.c_void_p()
        handle = (C.c_uint8 * 64)()
        self._check(self.lib.vrt_ipc_alloc_frame(self.h, nbytes, C.byref(p), handle))
        return p, bytes(handle)

    def ipc_open(self, handle):
        p = C.c_void_p()
        h = (C.c_uint8 * 64).from_buffer_copy(handle)
        self._check(self.lib.vrt_ipc_open(self.h, h, C.byref(p)))
        return p

    def signal_add(self, d_counter, value, stream=None):
        self._check(self.lib.vrt_signal_add(self.h, d_counter, value, stream))

    def stream_wait_value(self, d_counter, value, stream=None):
        self._check(self.lib.vrt_stream_wait_value(self.h, d_counter, value, stream))

    def ipc_close(self, p):
        self._check(self.lib.vrt_ipc_close(self.h, p))
