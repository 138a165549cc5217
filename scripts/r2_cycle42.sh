#!/bin/bash
# what bounds the decode attention: timing-only variants (results invalid) and per-warp granule-loop times of every CTA
mkdir -p gpurun_out
cp tortoise_tts_b200/lib/libtts_ar.so /tmp/base.so
for v in base attn_nomath attn_l2only; do
  if [ $v = base ]; then cp /tmp/base.so tortoise_tts_b200/lib/libtts_ar.so; else cp ab/libtts_ar_$v.so tortoise_tts_b200/lib/libtts_ar.so; fi
  TTS_AR_PROF_STEP=200 timeout 300 python scripts/stage_prof.py 16 300 > gpurun_out/stage_prof_16_$v.log 2>&1
  echo "== $v"; grep -E "^N=|w0:|CTA-level|attention of" gpurun_out/stage_prof_16_$v.log; grep -A32 "marks of layer 3" gpurun_out/stage_prof_16_$v.log | grep -E "attn"
done
cp /tmp/base.so tortoise_tts_b200/lib/libtts_ar.so
