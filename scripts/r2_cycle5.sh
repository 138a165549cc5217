#!/bin/bash
mkdir -p gpurun_out
for n in 16 32 64; do
  TTS_AR_PROF_STEP=200 timeout 300 python scripts/stage_prof.py $n 300 > gpurun_out/stage_$n.log 2>&1
  echo "=== $n"; grep -E "^N=|barrier |SAMPLE|HEAD:" gpurun_out/stage_$n.log; sed -n '/marks of layer 3/,/step span/p' gpurun_out/stage_$n.log | head -60
done
