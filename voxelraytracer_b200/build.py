"""In-tree build of libvrt.so for sm_100a (nvcc cross-compiles without a GPU)."""
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))


def build_library(force=False, verbose=False):
    """Runs `make` in csrc/ -> voxelraytracer_b200/libvrt.so.  Raises CalledProcessError on failure."""
    csrc = os.path.join(_HERE, "csrc")
    if force:
        subprocess.check_call(["make", "-C", csrc, "clean"], stdout=None if verbose else subprocess.DEVNULL)
    r = subprocess.run(["make", "-C", csrc], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout)
        print(r.stderr)
    if r.returncode != 0:
        raise subprocess.CalledProcessError(r.returncode, r.args, r.stdout, r.stderr)
    so = os.path.join(_HERE, "libvrt.so")
    assert os.path.exists(so), so
    return so
