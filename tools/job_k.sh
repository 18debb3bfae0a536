#!/bin/bash
timeout 120 python -m pytest tests/test_gpu_api.py -m gpu -x -q > gpurun_out/r02_pytest_api.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/r02_pytest_api.log
