#!/bin/bash
# A/B timing on ONE box (boxes differ by a few %): the in-tree library, then ab/libtts_ar_base.so, then the in-tree one again.
# AB_ENV="VAR=val" adds a third variant: the in-tree library with that environment.
run() { env $1 TTS_AR_PROF_STEP=200 timeout 300 python scripts/stage_prof.py 16 ${STEPS:-300} 2>&1 | grep -E "^N=|SAMPLE|HEAD:|smp" ; }
echo "--- new"; run
if [ -n "$AB_ENV" ]; then echo "--- new with $AB_ENV"; run "$AB_ENV"; fi
cp tortoise_tts_b200/lib/libtts_ar.so /tmp/new.so; cp ab/libtts_ar_base.so tortoise_tts_b200/lib/libtts_ar.so
echo "--- base"; run
cp /tmp/new.so tortoise_tts_b200/lib/libtts_ar.so
echo "--- new again"; run
