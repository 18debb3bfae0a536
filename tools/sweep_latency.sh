#!/bin/bash
cd "$(dirname "$0")/.."
for f in "$@"; do
  make -C voxelraytracer_b200/csrc clean > /dev/null; make -C voxelraytracer_b200/csrc EXTRA="$f" > /dev/null 2>&1 || { echo "BUILD FAILED: $f"; continue; }
  echo "== $f"; timeout 200 python tools/latency_probe.py 2>&1 | tail -3
done
make -C voxelraytracer_b200/csrc clean > /dev/null; make -C voxelraytracer_b200/csrc > /dev/null 2>&1
