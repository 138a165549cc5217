#!/bin/bash
# decode attention: PV accumulation as paired fp32 FMAs (FFMA2, bit-identical) | + QK in two chains; same-box A/B, parity
mkdir -p gpurun_out
bash scripts/ab_run.sh "16 300" base pre qkffma2 > gpurun_out/ab_attention_ffma2.log 2>&1; cat gpurun_out/ab_attention_ffma2.log
bash scripts/ab_run.sh "64 300" base pre qkffma2 > gpurun_out/ab_attention_ffma2_64.log 2>&1; cat gpurun_out/ab_attention_ffma2_64.log
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_parity.log 2>&1; tail -3 gpurun_out/pytest_parity.log
