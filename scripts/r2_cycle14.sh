#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
bash scripts/ab_run.sh "16 300" base r2base > gpurun_out/ab_dataflow16.log 2>&1; cat gpurun_out/ab_dataflow16.log
bash scripts/ab_run.sh "64 300" base r2base > gpurun_out/ab_dataflow64.log 2>&1; cat gpurun_out/ab_dataflow64.log
