#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
for n in 16 32 64; do
  TTS_AR_PROF_STEP=200 timeout 300 python scripts/stage_prof.py $n 300 > gpurun_out/stage_$n.log 2>&1
  echo "=== $n"; grep -E "^N=|SAMPLE|HEAD:" gpurun_out/stage_$n.log; sed -n '/marks of layer 3/,/step span/p' gpurun_out/stage_$n.log | grep -E "attn|post-bar|pre-bar"
done
timeout 900 python scripts/config_sweep.py C3_share C4_share C5_share > gpurun_out/sweep.log 2>&1; tail -5 gpurun_out/sweep.log
