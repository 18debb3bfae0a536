"""voxelraytracer_b200 -- B200-native replacement for VoxelRayTracer's per-pixel pass.

The product is the C-ABI shared library ``libvrt.so`` (sources in ``csrc/``, declared in
``include/vrt.h``).  This package is the thin ctypes binding used by the tests and ``bench.py``; a C++
host uses ``host/`` (FrameBuffer / Cam / AppScene mirroring src/main.cpp and src/FrameBuffer.h) instead.

There is no CPU fallback: ``Context()`` raises ``VrtError`` when no CUDA device is usable, and
importing ``load_library`` raises ``ImportError`` when ``libvrt.so`` has not been built
(``python -c "import __graft_entry__ as g; g.build()"`` or ``make -C voxelraytracer_b200/csrc``).
"""
from .api import (  # noqa: F401
    FLAG_CAREFUL,
    FLAG_COLOR_ONLY,
    FLAG_COUNTERS,
    FLAG_DEBUG_MAPS,
    FLAG_EXACT,
    FLAG_FLOAT_OUT,
    FLAG_OUT_FRAME,
    FLAG_PERSISTENT,
    FLAG_RAY_QUEUE,
    FLAG_RGB8_OUT,
    HIT_MISS,
    OPT_OCTANT_FIELDS,
    OPT_RADIUS_CAP,
    OPT_HOST_DIRECT,
    OPT_REMOTE_ROWS,
    OPT_SUN_FIELD,
    Context,
    Counters,
    GridInfo,
    Params,
    VrtError,
    atlas_pack,
    atlas_synthetic,
    camera_invpv,
    default_params,
    lib_path,
    load_host_library,
    load_library,
    noise_heightmap,
    scene_fill,
    sun_dir,
    tile_pixels,
    write_png,
)
from .build import build_library  # noqa: F401
from .sharding import band_params, stitch_bands  # noqa: F401

__all__ = [n for n in dir() if not n.startswith("_")]
