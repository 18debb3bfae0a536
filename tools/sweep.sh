#!/bin/bash
# On the GPU box: rebuild libvrt.so with each set of -D tunables and time the C2 bench.  usage: tools/sweep.sh "<flags1>" "<flags2>" ...
cd "$(dirname "$0")/.."
for f in "$@"; do
  make -C voxelraytracer_b200/csrc clean > /dev/null; make -C voxelraytracer_b200/csrc EXTRA="$f" > /dev/null 2>&1 || { echo "BUILD FAILED: $f"; continue; }
  regs=$(grep -A2 "FastTracerILb0ELb0" voxelraytracer_b200/csrc/kernels.ptxas.log | grep -o "Used [0-9]* registers" | head -1)
  python bench.py --steps 200 --warmup 20 --no-cpu-baseline ${SWEEP_ARGS} 2>/dev/null | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('%-60s %s  %.1f Mrays/s  %.4f ms' % ('$f', '$regs', d['value'], d['ms_per_step']))"
done
make -C voxelraytracer_b200/csrc clean > /dev/null; make -C voxelraytracer_b200/csrc > /dev/null 2>&1
