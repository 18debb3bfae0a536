#!/bin/bash
# usage (on the GPU box, via gpurun): tools/prof.sh <tag> [bench args...]   -> gpurun_out/<tag>.ncu-rep (+ plain bench line)
# Captures 3 launches of the TIMED kernel variant only (fast tracer, no debug attachments), after the warm-up launches.
tag=$1; shift
CMD="python bench.py --steps 6 --warmup 3 --no-cpu-baseline $*"
timeout 300 $CMD > gpurun_out/${tag}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${tag}_plain.log; exit 1; }
timeout 800 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k 'regex:FastTracer<\(bool\)[01], \(bool\)0, \(bool\)[01], \(bool\)[01]>, \(bool\)[01], \(bool\)0' -s 4 -c 3 -o gpurun_out/${tag} $CMD > gpurun_out/${tag}_ncu.log 2>&1
echo ncu rc=$?
