"""vrt_repack_static: the repack variant that allocates nothing (scratch in a fixed device array inside libvrt.so), against the existing
entry point vrt_upload_grid (+ vrt_read_cells) on the same chunks.  Integer work: the cells must be bit-identical.  Two calls in a row on
different chunks (the second must not see the first one's scratch contents), then the first chunk again.
Named test_zz_* so that pytest collects it last."""
import numpy as np
import pytest

import voxelraytracer_b200 as vrt
from common import scene

pytestmark = pytest.mark.gpu


def expected_cells(ctx, grid, n):
    """The existing path: vrt_upload_grid with the isotropic layout, read back."""
    ctx.set_option(vrt.OPT_OCTANT_FIELDS, 0)
    try:
        ctx.upload_grid(grid, n)
        assert ctx.grid_info().octants == 1
        return ctx.read_cells()
    finally:
        ctx.set_option(vrt.OPT_OCTANT_FIELDS, 1)


def test_repack_static_equals_upload_grid_on_two_calls_in_a_row(ctx):
    chunks = [scene("terrain128")[:2], scene("refraction128")[:2], scene("glass_cube32")[:2], scene("terrain128")[:2]]
    want = [expected_cells(ctx, g, n) for g, n in chunks]
    # the caller's own buffers (allocated up front, outside the call under test)
    d_grid = ctx.device_alloc(128 ** 3)
    d_cells = ctx.device_alloc(130 ** 3 * 2)
    try:
        launches = ctx.kernel_launches()
        for (grid, n), exp in zip(chunks, want):
            ctx.copy_to_device(d_grid, grid)
            ctx.device_memset(d_cells, 0xEE, 130 ** 3 * 2)
            ctx.repack_static(d_grid, n, d_cells)
            got = ctx.copy_to_host(d_cells, (n + 2) ** 3 * 2).view(np.uint16)
            assert np.array_equal(got, exp), "size %d: %d cells differ" % (n, int((got != exp).sum()))
        assert ctx.kernel_launches() - launches == 4 * len(chunks)            # 3 distance passes + 1 pack per call
        # back to back without a host round trip in between: call, call, then look at the second result
        ctx.copy_to_device(d_grid, chunks[1][0])
        ctx.repack_static(d_grid, 128, d_cells)
        ctx.repack_static(d_grid, 128, d_cells)
        assert np.array_equal(ctx.copy_to_host(d_cells, 130 ** 3 * 2).view(np.uint16), want[1])
    finally:
        ctx.device_free(d_grid)
        ctx.device_free(d_cells)


def test_repack_static_rejects_sizes_beyond_its_scratch(ctx):
    d = ctx.device_alloc(1 << 20)
    try:
        for size in (256, 1024, 100, 4, 0, -8):
            with pytest.raises(vrt.VrtError) as e:
                ctx.repack_static(d, size, d)
            assert e.value.code == -1                                          # VRT_E_INVALID
        with pytest.raises(vrt.VrtError, match="VRT_REPACK_STATIC_MAX_SIZE = 128"):
            ctx.repack_static(d, 256, d)
        with pytest.raises(vrt.VrtError):
            ctx.repack_static(None, 128, d)
    finally:
        ctx.device_free(d)
    assert vrt.api.REPACK_STATIC_MAX_SIZE == 128
