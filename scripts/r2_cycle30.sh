#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_cond.py tests/test_gpu_clvp.py -m gpu -x -q -s > gpurun_out/pytest_cond.log 2>&1; tail -14 gpurun_out/pytest_cond.log
