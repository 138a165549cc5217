#!/bin/bash
mkdir -p gpurun_out
for n in "16 300" "32 300" "64 300"; do bash scripts/ab_run.sh "$n" base lnasync 2>&1 | grep -E "^---|^N=" | tail -4; done | tee gpurun_out/ab35.log
cp ab/libtts_ar_lnasync.so tortoise_tts_b200/lib/libtts_ar.so
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
