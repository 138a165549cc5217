#!/bin/bash
# diagnosis of the staged-K/V slowdown: old | half ln_1 only | + two ring slots reserved | + copies issued, granule from global | + granule from shared memory
mkdir -p gpurun_out
bash scripts/ab_run.sh "16 300" old nokv rsv dry lnfull > gpurun_out/ab_kv_stage_diag.log 2>&1; cat gpurun_out/ab_kv_stage_diag.log
for v in old nokv lnfull; do
  cp ab/libtts_ar_$v.so tortoise_tts_b200/lib/libtts_ar.so
  TTS_AR_PROF_STEP=200 timeout 300 python scripts/stage_prof.py 16 300 > gpurun_out/stage_prof_16_$v.log 2>&1
  echo "== $v"; grep -A32 "marks of layer 3" gpurun_out/stage_prof_16_$v.log | grep -E "attn|A-ready|ln-|post-bar|pre-bar"
done
