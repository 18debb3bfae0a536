#!/bin/bash
export VRT_LIBRARY=build/vq/libvrt.so
timeout 120 python tools/queue_ncu.py > gpurun_out/r02_queue_ncu_plain.log 2>&1 && timeout 400 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:render_kernel -s 2 -c 4 -o gpurun_out/r02_queue python tools/queue_ncu.py > gpurun_out/r02_queue_ncu.log 2>&1; echo ncu rc=$?; tail -2 gpurun_out/r02_queue_ncu.log
