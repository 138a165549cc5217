#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -40 gpurun_out/pytest_gpu.log
bash scripts/ab_run.sh "16 300" base nofence acqpoll > gpurun_out/ab_fence.log 2>&1; cat gpurun_out/ab_fence.log
