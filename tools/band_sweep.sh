#!/bin/bash
# On the GPU box: CTA-size variants against the per-frame fixed cost (C2 frame time, C4 frame time, one rank's eighth of a C4 frame).
cd "$(dirname "$0")/.."
for f in "" "-DVRT_LOCAL_WX=2 -DVRT_LOCAL_WY=1" "-DVRT_LOCAL_WX=1 -DVRT_LOCAL_WY=2" "-DVRT_LOCAL_WX=1 -DVRT_LOCAL_WY=1" "-DVRT_MIN_BLOCKS=16" "-DVRT_MIN_BLOCKS=10"; do
  make -C voxelraytracer_b200/csrc clean > /dev/null; make -C voxelraytracer_b200/csrc EXTRA="$f" > /dev/null 2>&1 || { echo "BUILD FAILED: $f"; continue; }
  echo "== ${f:-default}"
  for w in c2 c4; do
    python bench.py --workload $w --steps 150 --warmup 20 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('   %-4s %9.1f Mrays/s  %.4f ms' % ('$w', d['value'], d['ms_per_step']))"
  done
  python tools/band_probe.py 2>/dev/null | sed -n '1,2p' | sed 's/^/   /'
done
make -C voxelraytracer_b200/csrc clean > /dev/null; make -C voxelraytracer_b200/csrc > /dev/null 2>&1
