#!/bin/bash
# GPU job: full GPU test suite, then A/B timing of kernel variants under build/
python -m pytest tests -m gpu -x -q --durations=5 -s > gpurun_out/r02_pytest_b.log 2>&1; echo pytest rc=$?; grep -E "mismatching px|passed|failed|Error" gpurun_out/r02_pytest_b.log | tail -12
python tools/ab_time.py --cases c2,c3,c4,c4b - $(ls build/*/libvrt.so) > gpurun_out/r02_ab_b.txt 2>&1; cat gpurun_out/r02_ab_b.txt
