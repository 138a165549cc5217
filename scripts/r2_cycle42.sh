#!/bin/bash
# MODE 1 ln_1 per CTA (8 rows, no distributed-shared exchange, no cluster barrier): parity, same-box A/B
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_parity.log 2>&1; tail -3 gpurun_out/pytest_parity.log
bash scripts/ab_run.sh "16 300" base pre > gpurun_out/ab_ln1_local.log 2>&1; cat gpurun_out/ab_ln1_local.log
bash scripts/ab_run.sh "16 520" base pre > gpurun_out/ab_ln1_local_520.log 2>&1; cat gpurun_out/ab_ln1_local_520.log
TTS_AR_PROF_STEP=200 timeout 300 python scripts/stage_prof.py 16 300 > gpurun_out/stage_prof_16_ln1local.log 2>&1
grep -A14 "marks of layer 3" gpurun_out/stage_prof_16_ln1local.log
