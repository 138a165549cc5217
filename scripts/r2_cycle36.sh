#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_clvp.py tests/test_gpu_cond.py -m gpu -x -q 2>&1 | tail -8
