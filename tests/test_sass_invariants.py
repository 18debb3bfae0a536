"""Properties of the GENERATED code of the timed render kernels that the measurements depend on and that an innocent edit of
the headers can silently lose.  Reads the SASS of the in-tree build with cuobjdump; no GPU needed."""
import os
import re
import shutil
import subprocess

import pytest

from voxelraytracer_b200 import build_library

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "voxelraytracer_b200", "csrc", "kernels.o")
CUOBJDUMP = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
# render_kernel<FastTracer<FLAT=0, COUNT=0, NOISE=0, GLASS=g>, FLAT=0, DEBUG=0, 2, 2, DIRECT=1, QUEUE=0>: the kernels bench.py times at C2 (g=0) and C3/C4 (g=1)
KERNEL = "_ZN3vrt13render_kernelINS_10FastTracerILb0ELb0ELb0ELb%dEEELb0ELb0ELi2ELi2ELb1ELb0EEEvNS_9FrameArgsE"

pytestmark = pytest.mark.skipif(not os.path.exists(CUOBJDUMP), reason="cuobjdump not available")


def sass_of(kernel):
    build_library()
    out = subprocess.run([CUOBJDUMP, "-sass", OBJ], capture_output=True, text=True, check=True).stdout
    keep, lines = False, []
    for ln in out.splitlines():
        if "Function :" in ln:
            keep = ln.strip().endswith(kernel)
            continue
        if keep:
            m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(.*?);", ln)
            if m:
                lines.append(m.group(1).strip())
    assert lines, "kernel not found in kernels.o: " + kernel
    return lines


@pytest.mark.parametrize("glass", [0, 1])
def test_walk_loop_joins_its_branches_before_every_load(glass):
    """walk_air is "jump, else step, then one load": ptxas must keep the BSSY/BSYNC pair that joins both branches before the
    2-byte cell load, in the march walk and in the shadow walk.  Without it the lanes of a warp drift apart from cell to cell
    (measured: C2 0.080 -> 0.102 ms, profiles/r01_ab_walk_two_phase_no_barrier.txt)."""
    s = sass_of(KERNEL % glass)
    loads = [i for i, ins in enumerate(s) if "LDG.E.U16" in ins]
    joined = [i for i in loads if any("BSYNC" in x for x in s[max(0, i - 3):i])]
    assert len(joined) >= 2, "cell loads preceded by a convergence barrier: %d of %d" % (len(joined), len(loads))


def test_kernel_for_chunks_without_glass_uses_no_local_memory():
    """GLASS = false drops the ray stack: no local loads / stores at all, and the code stays small enough for the instruction cache."""
    s = sass_of(KERNEL % 0)
    assert not [x for x in s if re.search(r"\b(LDL|STL)\b", x.split()[0] if not x.startswith("@") else x.split()[1])]
    assert len(s) < 1800, len(s)
    assert len(sass_of(KERNEL % 1)) < 3000
