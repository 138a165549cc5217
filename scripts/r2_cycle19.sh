#!/bin/bash
mkdir -p gpurun_out
bash scripts/ab_run.sh "16 300" base st16 rs1 rs2 rs4 rs8 2>&1 | grep -E "^---|^N=" > gpurun_out/ab19.log; cat gpurun_out/ab19.log
cp ab/libtts_ar_st16.so tortoise_tts_b200/lib/libtts_ar.so
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "c2_full or wide_launches or golden" 2>&1 | tail -3
