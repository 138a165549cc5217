#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
bash scripts/ab_run.sh "16 300" prev late base 2>&1 | grep -E "^---|^N=" > gpurun_out/ab26.log; cat gpurun_out/ab26.log
bash scripts/ab_run.sh "32 300" prev base 2>&1 | grep -E "^---|^N=" | tail -4 | tee -a gpurun_out/ab26.log
