#!/bin/bash
# usage (on the GPU box, via gpurun): tools/prof.sh <tag> [bench args...]   -> gpurun_out/<tag>.ncu-rep (+ plain bench line)
tag=$1; shift
CMD="python bench.py --steps 6 --warmup 3 --no-cpu-baseline $*"
$CMD > gpurun_out/${tag}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${tag}_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 8 -c 3 -o gpurun_out/${tag} $CMD > gpurun_out/${tag}_ncu.log 2>&1
echo ncu rc=$?
