"""Screen-space sharding of one frame across ranks.

Pixels are independent (voxel.glsl:425-452 reads only uniforms and read-only textures), so a frame is
split into horizontal bands of `band_rows` scanlines dealt round-robin to the ranks: rank r renders bands
r, r+N, r+2N, ...  Interleaving balances sky rows against terrain rows.  Each rank's kernel writes its
bands compactly; `stitch_bands` (host, tests) or the all-gather + de-interleave in bench.py puts the
frame back together.  Pure index arithmetic - no device code here.
"""
import numpy as np


def band_params(p, rank, world, band_rows=16):
    """Copy of `p` restricted to the bands of `rank` (band_count = world)."""
    q = p.copy()
    if world > 1:
        q.band_rows = band_rows
        q.band_count = world
        q.band_index = rank
    return q


def band_rows_of(height, rank, world, band_rows=16):
    """Frame rows (relative to the tile) that `rank` renders, in the compact order the kernel writes them."""
    rows = []
    nb = (height + band_rows - 1) // band_rows
    for b in range(rank, nb, world):
        rows.extend(range(b * band_rows, min((b + 1) * band_rows, height)))
    return np.asarray(rows, np.int64)


def max_rows(height, world, band_rows=16):
    """Rows of the busiest rank (rank 0): the common buffer height for an equal-sized all-gather."""
    return len(band_rows_of(height, 0, world, band_rows))


def stitch_bands(parts, height, width, band_rows=16):
    """parts[r]: array (>= rows_r, width, C) from rank r -> full (height, width, C) frame."""
    world = len(parts)
    out = np.zeros((height, width) + parts[0].shape[2:], parts[0].dtype)
    for r, part in enumerate(parts):
        rows = band_rows_of(height, r, world, band_rows)
        out[rows] = part[: len(rows)]
    return out


def row_permutation(height, world, band_rows=16):
    """For the gathered buffer laid out [rank][max_rows] (row-major), the index of each frame row:
    frame[y] = gathered.reshape(world * max_rows, ...)[perm[y]]."""
    mr = max_rows(height, world, band_rows)
    perm = np.zeros(height, np.int64)
    for r in range(world):
        rows = band_rows_of(height, r, world, band_rows)
        perm[rows] = r * mr + np.arange(len(rows))
    return perm
