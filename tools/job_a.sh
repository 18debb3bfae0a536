#!/bin/bash
# GPU job: tests, A/B timing, bench lines, ncu captures of the timed kernels (run on the box via gpurun)
python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/r02_pytest_a.log 2>&1; echo pytest rc=$?; tail -25 gpurun_out/r02_pytest_a.log
python tools/ab_time.py --cases c2,c3,c4,c4b - > gpurun_out/r02_ab_a.txt 2>&1; cat gpurun_out/r02_ab_a.txt
python bench.py --steps 30 --warmup 5 > gpurun_out/r02_bench_c2_a.json 2> gpurun_out/r02_bench_c2_a.err; echo bench rc=$?
python bench.py --impl reference --steps 10 --warmup 2 > gpurun_out/r02_bench_ref_a.json 2> gpurun_out/r02_bench_ref_a.err; echo ref rc=$?
bash tools/prof.sh r02_a_c2 --workload c2; bash tools/prof.sh r02_a_c4 --workload c4; bash tools/prof.sh r02_a_c3 --workload c3
python tools/size_probe.py > gpurun_out/r02_size_probe_a.txt 2>&1; cat gpurun_out/r02_size_probe_a.txt
python tools/latency_probe.py > gpurun_out/r02_latency_probe_a.txt 2>&1; cat gpurun_out/r02_latency_probe_a.txt
python tools/launch_floor.py > gpurun_out/r02_launch_floor_a.txt 2>&1; cat gpurun_out/r02_launch_floor_a.txt
