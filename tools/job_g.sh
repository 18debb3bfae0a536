#!/bin/bash
# ncu captures of the timed kernels at HEAD + bench lines for every workload at 1 GPU
for w in c2 c3 c4; do bash tools/prof.sh r02_g_$w --workload $w; done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_g_launches_c2.csv python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/r02_g_launches_c2.log 2>&1
python tools/size_probe.py > gpurun_out/r02_g_size_probe.txt 2>&1; cat gpurun_out/r02_g_size_probe.txt
