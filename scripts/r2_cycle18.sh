#!/bin/bash
mkdir -p gpurun_out
bash scripts/ab_run.sh "16 300" base stagger relaxdiag 2>&1 | grep -E "^---|^N=" > gpurun_out/ab18.log; cat gpurun_out/ab18.log
TTS_AR_PROF_STEP=150 timeout 300 python scripts/stage_prof.py 16 300 > gpurun_out/stage_prof_16.log 2>&1; tail -60 gpurun_out/stage_prof_16.log
