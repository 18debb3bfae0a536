"""The product's device headers compiled for the host (tests/hostsim) against the oracle.
Exact tracer: bit-identical.  Fast tracer: north_star's caps (hit map / bounce count differ on at most
0.01 % of the frame; colour within 1/255 RMS per channel) - with a margin, at a size the CPU does in seconds."""
import ctypes as C

import numpy as np
import pytest

import voxelraytracer_b200 as vrt
from common import hostsim_cells, hostsim_lib, hostsim_render, make_params, oracle_render, rms255, scene

CASES = [
    ("glass_cube32", "start", vrt.FLAG_COLOR_ONLY),
    ("glass_cube32", "start", 0),
    ("terrain32", "terrain32_above", 0),
    ("terrain128", "terrain_inside", 0),
    ("terrain128", "terrain_low", 0),
    ("terrain128", "terrain_above", 0),
    ("terrain128", "terrain_outside", 0),
    ("terrain128_glass", "glass_wall", 0),
    ("refraction128", "refraction_close", 0),
    ("refraction128", "refraction_close", vrt.FLAG_COLOR_ONLY),
]
W, H = 480, 270


@pytest.mark.parametrize("sc,pose,flags", CASES)
def test_exact_tracer_is_bit_identical(sc, pose, flags):
    grid, n, atlas, ts = scene(sc)
    p = make_params(W, H, sc, pose, flags)
    o = oracle_render(p, grid, atlas)
    e = hostsim_render(p, grid, atlas, 0)
    assert np.array_equal(e.hit, o.hit) and np.array_equal(e.bounces, o.bounces)
    assert np.array_equal(e.f32.view(np.uint32), o.f32.view(np.uint32))
    for k in ("primary_rays", "secondary_rays", "shadow_rays", "march_steps", "shadow_steps", "hits", "tie_index3"):
        assert e.counters[k] == o.counters[k], k


def test_exact_tracer_with_noise_is_bit_identical():
    grid, n, atlas, ts = scene("glass_cube32")
    p = make_params(200, 200, "glass_cube32", "start", vrt.FLAG_COLOR_ONLY, noise=(0.02, 0.1, 0.1), time=7.0)
    o = oracle_render(p, grid, atlas)
    e = hostsim_render(p, grid, atlas, 0)
    assert np.array_equal(e.f32.view(np.uint32), o.f32.view(np.uint32)) and np.array_equal(e.bounces, o.bounces)


@pytest.mark.parametrize("sc,pose,flags", CASES)
def test_fast_tracer_within_caps(sc, pose, flags):
    grid, n, atlas, ts = scene(sc)
    p = make_params(W, H, sc, pose, flags)
    o = oracle_render(p, grid, atlas)
    f = hostsim_render(p, grid, atlas, 1)
    npx = len(o.hit)
    bad = int(((f.hit != o.hit) | (f.bounces != o.bounces)).sum())
    assert bad <= max(2, int(1e-4 * npx)), "hit/bounce mismatches: %d of %d" % (bad, npx)   # 0.01 % (floor 2 px at this size)
    r = rms255(f.rgba8, o.rgba8)
    assert (r <= 1.0).all(), r
    # the walk looks at far fewer cells than the shader steps through
    s_ref = o.counters["march_steps"] + o.counters["shadow_steps"]
    assert f.counters["actual_steps"] < s_ref
    assert f.counters["primary_rays"] == o.counters["primary_rays"]


def test_fast_tracer_custom_limits_and_ray_length():
    grid, n, atlas, ts = scene("glass_cube32")
    p = make_params(320, 180, "glass_cube32", "start")
    p.maxReflections, p.maxTransparencies, p.maxRayLength = 2, 3, 60.0
    o = oracle_render(p, grid, atlas)
    f = hostsim_render(p, grid, atlas, 1)
    bad = int(((f.hit != o.hit) | (f.bounces != o.bounces)).sum())
    assert bad <= 3 and o.bounces.max() > 4, o.bounces.max()  # default limits give at most 4 here
    assert (rms255(f.rgba8, o.rgba8) <= 1.0).all()


def test_cells_radius_is_exact_chebyshev_distance():
    """Both layouts against the brute-force definition: isotropic Chebyshev distance, and per octant the forward
    Chebyshev distance (obstacles = occupied voxels only - space outside the chunk is empty, the walk bounds its jumps by
    the ray's exit time - restricted to the closed forward octant)."""
    rng = np.random.default_rng(3)
    n = 16
    g = (rng.random(n ** 3) < 0.01).astype(np.uint8) * rng.integers(1, 4, n ** 3).astype(np.uint8)
    occ = np.zeros((n + 2,) * 3, bool)         # the border is not an obstacle
    occ[1:-1, 1:-1, 1:-1] = g.reshape(n, n, n) != 0
    zz, yy, xx = np.nonzero(occ)
    pts = np.stack([zz, yy, xx], 1)
    iso = hostsim_cells(g, n, 0).reshape(n + 2, n + 2, n + 2)
    octs = hostsim_cells(g, n, 1).reshape(8, n + 2, n + 2, n + 2)
    for _ in range(300):
        c = rng.integers(1, n + 1, 3)          # (z, y, x), padded coordinates
        byte = g.reshape(n, n, n)[c[0] - 1, c[1] - 1, c[2] - 1]
        d = np.abs(pts - c).max(axis=1).min() if len(pts) else 1 << 20
        v = int(iso[c[0], c[1], c[2]])
        edge = (c == 1).any() or (c == n).any()
        assert (v >> 8) == (1 if d == 2 and edge else min(d, 255)) and (v & 0xFF) == byte     # radius 2 is not given next to the border
        for o in range(8):
            sgn = np.array([-1 if o & 4 else 1, -1 if o & 2 else 1, -1 if o & 1 else 1])   # (z, y, x)
            rel = (pts - c) * sgn
            fwd = (rel >= 0).all(axis=1)
            df = rel[fwd].max(axis=1).min() if fwd.any() else 1 << 20
            w = int(octs[o, c[0], c[1], c[2]])
            fedge = ((c == np.where(sgn > 0, n, 1))).any()
            assert (w >> 8) == (1 if df == 2 and fedge else min(df, 255)) and (w & 0xFF) == byte, (o, c, w >> 8, df)
            assert (w >> 8) >= (v >> 8)        # a forward box is never smaller than the centred cube
    for cells in (iso, octs[0], octs[7]):
        assert (cells[0] == 0).all() and (cells[:, 0] == 0).all() and (cells[:, :, -1] == 0).all()


def test_fast_tracer_with_noise_uniforms():
    """Non-zero noise uniforms run the careful instantiation (rt_trace_fast.cuh): rays are built with the pinned arithmetic and, when
    reflection / refraction noise is set, every ray that ends on glass is replayed with the literal recurrence, so the hit point the
    hash sees (voxel.glsl:132-140,203-246) has the reference's bits.  Both bars of north_star hold on the default path: hit / bounce
    maps within 0.01 % and colour within 1/255 RMS - while most rays still take the skipping walk."""
    for sc, pose, fl, W, H in (("glass_cube32", "start", vrt.FLAG_COLOR_ONLY, 320, 180), ("glass_cube32", "start", 0, 320, 180),
                               ("terrain32", "terrain32_above", 0, 320, 180), ("terrain128_glass", "glass_wall", 0, 480, 270)):
        grid, n, atlas, ts = scene(sc)
        for noise in ((0.01, 0.0, 0.0), (0.01, 0.08, 0.05), (0.0, 0.1, 0.0), (0.0, 0.0, 0.1)):
            p = make_params(W, H, sc, pose, fl, noise=noise, time=9.0)
            o = oracle_render(p, grid, atlas)
            f = hostsim_render(p, grid, atlas, 1)
            bad = int(((f.hit != o.hit) | (f.bounces != o.bounces)).sum())
            assert bad <= max(3, int(1e-4 * W * H)), (sc, noise, bad)
            assert (rms255(f.rgba8, o.rgba8) <= 1.0).all(), (sc, noise, rms255(f.rgba8, o.rgba8))
            ref_steps = o.counters["march_steps"] + o.counters["shadow_steps"]
            if sc == "terrain128_glass":
                # only glass-ending rays are replayed (this pose looks AT a glass wall), and none without bounce noise
                assert f.counters["actual_steps"] < (0.7 if noise[1] or noise[2] else 0.5) * ref_steps


def test_camera_on_a_far_chunk_face_replays_the_rays_that_start_on_it():
    """Camera exactly on the +x face, looking up: the centre pixel column starts ON the plane x == size with a direction that keeps
    the float position on it for many steps; the reference reads the REPEAT-wrapped column 0 there (pin G4).  finish_frame_args
    flags the frame and the careful instantiation replays those rays (tools/fuzz_hostsim.py seed 2 case 273: 16 px before)."""
    grid, n, atlas, ts = scene("refraction32")
    p = make_params(97, 61, "refraction32", ((16.0, -7.1482414338108065, 15.1391967032303), (1.5692255304681018, -3.141592653589793, 0.0)), 0,
                    time_of_day=31.436090084183842, max_len=400.0)
    p.maxReflections, p.maxTransparencies = 0, 1
    o = oracle_render(p, grid, atlas)
    g = hostsim_render(p, grid, atlas, 1)
    assert int(((g.hit != o.hit) | (g.bounces != o.bounces)).sum()) == 0
    assert (rms255(g.rgba8, o.rgba8) <= 1.0).all()


def test_lattice_line_eye_with_an_aligned_view_is_careful_and_generic_views_stay_lean():
    """An eye on a lattice line whose view sends rays along voxel edges (axis-aligned, 45 degrees) ties on whole families of pixels:
    finish_frame_args makes the frame careful and the tie-tracking walk replays the tied rays (fuzz seed 20261105 case 4: 52 px of
    16 k before).  The same eye with a generic rotation - bench pose terrain_above sits on the line x = z = 64 - stays lean."""
    for sc, pos, rot, W, H in (("terrain32", (-5.0, 1.0, -8.0), (0.0, -np.pi / 2, 0.0), 256, 64), ("terrain128_glass", (0.0, 0.0, 0.0), (0.0, np.pi / 4, 0.0), 320, 180)):
        grid, n, atlas, ts = scene(sc)
        p = make_params(W, H, sc, (pos, rot))
        o = oracle_render(p, grid, atlas)
        g = hostsim_render(p, grid, atlas, 1)
        assert int(((g.hit != o.hit) | (g.bounces != o.bounces)).sum()) <= max(1, int(1e-4 * W * H)), (sc, rot)
        assert (rms255(g.rgba8, o.rgba8) <= 1.0).all()
    assert hostsim_lib().hostsim_frame_is_careful(C.byref(make_params(256, 64, "terrain32", ((-5.0, 1.0, -8.0), (0.0, -np.pi / 2, 0.0))))) == 1
    for pose in ("terrain_inside", "terrain_low", "terrain_above", "terrain_outside", "start", "glass_wall", "refraction_close"):
        assert hostsim_lib().hostsim_frame_is_careful(C.byref(make_params(1920, 1080, "terrain128", pose))) == 0, pose


def test_sun_field_is_conservative_and_changes_no_pixel():
    """The sun field (rt_common.cuh): (1) property check independent of its construction - from random points of random empty cells a
    ray along the sun direction touches no occupied cell for as long as the cell's radius lets a jump travel; (2) it never promises
    less than the octant field it is built from; (3) frames rendered with it equal the frames rendered with the octant field (the same
    cells are tested, fewer are looked at), with fewer loop iterations."""
    import os
    lib = hostsim_lib()
    for sc in ("terrain128", "refraction128", "glass_cube32"):
        grid, n, atlas, ts = scene(sc)
        cells = hostsim_cells(grid, n, 1)
        for sun in ((0.0, 1.0, 0.2), (0.3, 0.5, -0.8), (-1.0, 0.01, 0.0), (0.577, 0.577, 0.577), (0.0, -0.3, 1.0)):
            out = np.zeros((n + 2) ** 3, np.uint16)
            s3 = (C.c_float * 3)(*sun)
            assert lib.hostsim_build_sun_cells(cells.ctypes.data_as(C.c_void_p), n, 1, s3, out.ctypes.data_as(C.c_void_p)) == 0
            assert lib.hostsim_check_sun_cells(out.ctypes.data_as(C.c_void_p), n, s3, 60000, 4242) == 0, (sc, sun)
            o = (1 if sun[0] < 0 else 0) | (2 if sun[1] < 0 else 0) | (4 if sun[2] < 0 else 0)
            base = cells.reshape(8, -1)[o]
            assert np.array_equal(out & 0xFF, base & 0xFF) and ((out >> 8) >= (base >> 8)).all()
    for sc, pose in (("terrain128", "terrain_low"), ("refraction128", "refraction_close"), ("terrain32", "terrain32_above")):
        grid, n, atlas, ts = scene(sc)
        p = make_params(480, 270, sc, pose)
        with_field = hostsim_render(p, grid, atlas, 1)
        os.environ["HOSTSIM_NO_SUN_FIELD"] = "1"
        try:
            without = hostsim_render(p, grid, atlas, 1)
        finally:
            del os.environ["HOSTSIM_NO_SUN_FIELD"]
        assert int((with_field.rgba8 != without.rgba8).any(axis=1).sum()) <= 2 and np.array_equal(with_field.hit, without.hit)
        assert with_field.counters["actual_steps"] < without.counters["actual_steps"]


def test_randomised_poses_terminate_and_exact_replay_stays_bit_exact():
    """A slice of tools/fuzz_hostsim.py: random cameras inside / around / on the chunk, random caps and bounce limits.  Every
    case runs in its own process under a timeout (a walk that never ends is a failure, not a hang of the suite)."""
    import json, os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "fuzz_hostsim.py"), "--cases", "48", "--seed", "7", "--jobs", "4", "--timeout", "90"],
                       capture_output=True, text=True, timeout=600)
    line = json.loads(r.stdout.splitlines()[0])
    assert line["hung_or_crashed"] == 0, r.stdout
    assert line["exact_not_bit_exact"] == 0, r.stdout
    assert line["mismatch_frac"] < 5e-4, r.stdout          # lattice-aligned cameras included (whole pixel lines on voxel edges)


@pytest.mark.parametrize("world,band_rows", [(8, 16), (3, 4), (4, 7), (2, 1)])
def test_band_rows_of_the_device_headers_stitch_to_the_full_frame(world, band_rows):
    """frame_row() of rt_common.cuh (shift path for 2^k-row bands, division otherwise): every rank's compact rows are the rows
    the sharding module says they are, so the bands stitch to the frame rendered in one piece."""
    from voxelraytracer_b200 import sharding
    grid, n, atlas, ts = scene("terrain32")
    Wb, Hb = 96, 70
    p = make_params(Wb, Hb, "terrain32", "terrain32_above")
    full = hostsim_render(p, grid, atlas, 1).rgba8.reshape(Hb, Wb, 4)
    parts = [hostsim_render(sharding.band_params(p, r, world, band_rows), grid, atlas, 1).rgba8.reshape(-1, Wb, 4) for r in range(world)]
    assert np.array_equal(sharding.stitch_bands(parts, Hb, Wb, band_rows), full)


def test_rays_in_the_far_boundary_plane_sample_the_wrapped_column():
    """Pin G4 in the fast traversal: a camera placed exactly on the chunk's +x face sends a line of pixels along the plane
    x == size with direction.x == 0; GetVoxel lets that coordinate through (voxel.glsl:151) and the REPEAT wrap reads column 0,
    not the wall in column size-1 (found by tools/fuzz_hostsim.py, seed 20261105 case 46)."""
    grid, n, atlas, ts = scene("terrain128")
    p = make_params(64, 200, "terrain128", ((64.0, -8.103337157281956, 52.80251210739749), (0.8001285390172226, 2.9449588927211945, 0.0)), 0,
                    time_of_day=42.711945995857704, max_len=400.0)
    p.maxReflections, p.maxTransparencies = 2, 1
    o = oracle_render(p, grid, atlas)
    g = hostsim_render(p, grid, atlas, 1)
    assert int(((g.hit != o.hit) | (g.bounces != o.bounces)).sum()) <= 1
    assert (rms255(g.rgba8, o.rgba8) <= 1.0).all()


def test_camera_on_a_lattice_point_starts_in_the_cell_it_moves_into():
    """A camera on integer coordinates with an axis-aligned view: whole lines of pixels start exactly on a lattice plane.  The
    reference counts a ray moving in the negative direction as being in the cell below the plane already (nextPlane =
    ceil(pos - 1), voxel.glsl:306-309), so that cell is never tested (tools/fuzz_hostsim.py, seed 2 case 5: 61 px before)."""
    grid, n, atlas, ts = scene("terrain128_glass")
    p = make_params(97, 61, "terrain128_glass", ((-39.0, -42.0, -41.0), (1.5692255304681018, 3.141592653589793, 0.0)), 0,
                    time_of_day=24.424928654198467)
    p.maxReflections, p.maxTransparencies = 0, 1
    o = oracle_render(p, grid, atlas)
    g = hostsim_render(p, grid, atlas, 1)
    assert int(((g.hit != o.hit) | (g.bounces != o.bounces)).sum()) <= 1
