#!/bin/bash
python -m pytest tests -m gpu -x -q --durations=5 -s > gpurun_out/r02_pytest_d.log 2>&1; echo pytest rc=$?; grep -E "mismatching px|passed|failed|Error|^E " gpurun_out/r02_pytest_d.log | tail -12
python tools/ab_time.py --cases c2,c3,c4,c4b - $(ls build/*/libvrt.so 2>/dev/null) > gpurun_out/r02_ab_d.txt 2>&1; cat gpurun_out/r02_ab_d.txt
python bench.py --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/r02_bench_c2_d.json 2> gpurun_out/r02_bench_c2_d.err; echo bench rc=$?; python -c "
import json; d=json.load(open('gpurun_out/r02_bench_c2_d.json')); print(d['value'], d['ms_per_step'], d['ms_per_step_median'], d['ms_per_step_min'], d['actual_steps_per_frame'], d['e2e']['value'])"
