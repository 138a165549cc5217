#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_host_cpu.py -m gpu -q -k nccl > gpurun_out/pytest_nccl.log 2>&1; tail -3 gpurun_out/pytest_nccl.log
bash scripts/r2_scale.sh 8
