#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -40 gpurun_out/pytest_gpu.log
bash scripts/ab_run.sh "16 300" base acqpoll > gpurun_out/ab_mode1.log 2>&1; cat gpurun_out/ab_mode1.log
timeout 900 python scripts/config_sweep.py C3_share C4_share C5_share > gpurun_out/sweep.log 2>&1; tail -5 gpurun_out/sweep.log
