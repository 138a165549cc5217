#!/bin/bash
mkdir -p gpurun_out
TTS_AR_PROF_STEP=200 timeout 300 python scripts/stage_prof.py 16 300 > gpurun_out/stage_16.log 2>&1
grep -E "^N=|barrier |SAMPLE|HEAD:" gpurun_out/stage_16.log; sed -n '/marks of layer 3/,/step span/p' gpurun_out/stage_16.log | head -60
