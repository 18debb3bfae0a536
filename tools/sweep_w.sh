#!/bin/bash
# On the GPU box: for each -D variant rebuild libvrt.so once and time several workloads.  usage: tools/sweep_w.sh "c2 c3 c4" "<flags1>" "<flags2>" ...
cd "$(dirname "$0")/.."
WL="$1"; shift
for f in "$@"; do
  make -C voxelraytracer_b200/csrc clean > /dev/null; make -C voxelraytracer_b200/csrc EXTRA="$f" > /dev/null 2>&1 || { echo "BUILD FAILED: $f"; continue; }
  for w in $WL; do
    python bench.py --workload $w --steps 200 --warmup 20 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('%-22s %-4s %9.1f Mrays/s  %.4f ms  pipelined %.1f' % ('$f' or 'default', '$w', d['value'], d['ms_per_step'], (d.get('pipelined') or {}).get('value', 0)))"
  done
done
make -C voxelraytracer_b200/csrc clean > /dev/null; make -C voxelraytracer_b200/csrc > /dev/null 2>&1
