#!/bin/bash
timeout 250 python bench.py --workload c4 --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/r02_bench_n1_c4.json 2> gpurun_out/r02_bench_n1_c4.err; echo c4 rc=$?
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_n1_c4.json')); print('c4 n1 value %.0f ms %.4f kernel %.4f pipelined %.0f (%s)' % (d['value'], d['ms_per_step'], d['kernel_ms_per_launch'], d['pipelined']['value'], d['pipelined']['note'][:30]))"
CMD="python bench.py --steps 6 --warmup 3 --no-cpu-baseline --workload c5"
timeout 400 $CMD > gpurun_out/r02_l_c5_plain.log 2>&1 && timeout 700 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:FastTracer<\(bool\)[01], \(bool\)0, \(bool\)[01], \(bool\)[01]>, \(bool\)[01], \(bool\)0' -s 4 -c 3 -o gpurun_out/r02_l_c5 $CMD > gpurun_out/r02_l_c5_ncu.log 2>&1; echo ncu rc=$?
tail -2 gpurun_out/r02_l_c5_ncu.log
