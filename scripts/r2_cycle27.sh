#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
for n in "16 300" "32 300" "64 300"; do bash scripts/ab_run.sh "$n" prev base 2>&1 | grep -E "^---|^N=" | tail -4; done | tee gpurun_out/ab27.log
