import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Everything the tests load is built in-tree first (no-ops when up to date)."""
    import common
    from voxelraytracer_b200 import build_library, lib_path
    if not os.path.exists(lib_path()):
        build_library()
    common.build_oracle()
    common.build_hostsim()
    yield


@pytest.fixture(scope="session")
def ctx():
    import voxelraytracer_b200 as vrt
    c = vrt.Context(0)
    # Tests render ONE frame per pose and compare loop-iteration counts with the host simulation, which always walks the sun field:
    # build it at once instead of on the second frame with the same sun (the default, VRT_OPT_SUN_FIELD = 1; tested on its own).
    c.set_option(vrt.OPT_SUN_FIELD, 2)
    yield c
    c.close()
