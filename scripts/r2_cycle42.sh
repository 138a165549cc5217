#!/bin/bash
# attention ring depth 2 as the default: parity of the decode kernels
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_parity.log 2>&1; tail -3 gpurun_out/pytest_parity.log
bash scripts/ab_run.sh "16 300" base pre > gpurun_out/ab_depth2_confirm.log 2>&1; cat gpurun_out/ab_depth2_confirm.log
