#!/bin/bash
# what bounds the decode attention: timing-only variants (results invalid): no K/V traffic | no arithmetic
mkdir -p gpurun_out
bash scripts/ab_run.sh "16 300" base attn_noload attn_nomath > gpurun_out/diag_attention_bound.log 2>&1; cat gpurun_out/diag_attention_bound.log
for v in attn_noload attn_nomath; do
  cp ab/libtts_ar_$v.so tortoise_tts_b200/lib/libtts_ar.so
  TTS_AR_PROF_STEP=200 timeout 300 python scripts/stage_prof.py 16 300 > gpurun_out/stage_prof_16_$v.log 2>&1
  echo "== $v"; grep -A32 "marks of layer 3" gpurun_out/stage_prof_16_$v.log | grep -E "attn|post-bar|pre-bar"
done
