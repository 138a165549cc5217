#!/bin/bash
# half-tile ln_1 (MODE 1) + staged K/V granules: parity tests of the decode kernels, then a same-box A/B of the variants
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_parity.log 2>&1; tail -3 gpurun_out/pytest_parity.log
bash scripts/ab_run.sh "16 300" base old nokv lnfull > gpurun_out/ab_ln_half_kv_stage.log 2>&1; cat gpurun_out/ab_ln_half_kv_stage.log
TTS_AR_PROF_STEP=200 timeout 300 python scripts/stage_prof.py 16 300 > gpurun_out/stage_prof_16_new.log 2>&1; head -70 gpurun_out/stage_prof_16_new.log
