"""Screen-tile sharding: band arithmetic, tile renders stitched == full frame, and the
N>1 gather path over a world_size-2 gloo group on the CPU (the oracle stands in for the kernel there)."""
import os
import subprocess
import sys

import numpy as np
import pytest

import voxelraytracer_b200 as vrt
from voxelraytracer_b200 import sharding
from common import ROOT, make_params, oracle_render, scene


@pytest.mark.parametrize("height,world,br", [(1080, 1, 16), (1080, 2, 16), (1080, 8, 16), (2160, 8, 16), (270, 4, 7), (5, 8, 16)])
def test_band_partition_covers_each_row_once(height, world, br):
    seen = np.zeros(height, int)
    p = vrt.default_params(64, height, 32)
    total = 0
    for r in range(world):
        rows = sharding.band_rows_of(height, r, world, br)
        seen[rows] += 1
        q = sharding.band_params(p, r, world, br)
        assert vrt.tile_pixels(q) == len(rows) * 64          # C side agrees with the Python index arithmetic
        total += len(rows)
        assert len(rows) <= sharding.max_rows(height, world, br)
    assert (seen == 1).all() and total == height
    perm = sharding.row_permutation(height, world, br)
    assert len(set(perm.tolist())) == height


def test_tiles_and_bands_stitch_to_the_full_frame_oracle():
    grid, n, atlas, ts = scene("terrain32")
    W, H = 96, 54
    p = make_params(W, H, "terrain32", "terrain32_above")
    full = oracle_render(p, grid, atlas).rgba8.reshape(H, W, 4)
    # rectangular tiles
    out = np.zeros_like(full)
    for (x0, y0, w, h) in [(0, 0, 50, 30), (50, 0, 46, 30), (0, 30, 96, 24)]:
        q = p.copy()
        q.tile_x0, q.tile_y0, q.tile_w, q.tile_h = x0, y0, w, h
        out[y0:y0 + h, x0:x0 + w] = oracle_render(q, grid, atlas).rgba8.reshape(h, w, 4)
    assert np.array_equal(out, full)
    # interleaved bands
    parts = [oracle_render(sharding.band_params(p, r, 3, 4), grid, atlas).rgba8.reshape(-1, W, 4) for r in range(3)]
    assert np.array_equal(sharding.stitch_bands(parts, H, W, 4), full)


WORKER = r"""
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import voxelraytracer_b200 as vrt
from voxelraytracer_b200 import sharding
from common import make_params, oracle_render, scene
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
grid, n, atlas, ts = scene("terrain32")
W, H, BR = 96, 54, 4
p = make_params(W, H, "terrain32", "terrain32_above")
q = sharding.band_params(p, rank, world, BR)
mine = oracle_render(q, grid, atlas, nthreads=2).rgba8.reshape(-1, W, 4)
mr = sharding.max_rows(H, world, BR)
buf = torch.zeros((mr, W, 4), dtype=torch.uint8)
buf[: mine.shape[0]] = torch.from_numpy(mine)
gathered = torch.zeros((world * mr, W, 4), dtype=torch.uint8)
dist.all_gather_into_tensor(gathered, buf)
perm = torch.from_numpy(sharding.row_permutation(H, world, BR))
frame = gathered[perm].numpy()
if rank == 0:
    full = oracle_render(p, grid, atlas, nthreads=2).rgba8.reshape(H, W, 4)
    assert np.array_equal(frame, full), "stitched frame differs from the single-rank frame"
    print("GLOO_OK")
dist.destroy_process_group()
"""


def test_gather_path_world_size_2_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % {"root": ROOT})
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29517", str(script)], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0 and "GLOO_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
