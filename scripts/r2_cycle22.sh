#!/bin/bash
mkdir -p gpurun_out
bash scripts/ab_run.sh "16 300" base ctxasync 2>&1 | grep -E "^---|^N=" > gpurun_out/ab22.log; cat gpurun_out/ab22.log
cp ab/libtts_ar_ctxasync.so tortoise_tts_b200/lib/libtts_ar.so
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
bash scripts/ab_run.sh "32 300" base ctxasync 2>&1 | grep -E "^---|^N=" | tail -4
bash scripts/ab_run.sh "64 300" base ctxasync 2>&1 | grep -E "^---|^N=" | tail -4
