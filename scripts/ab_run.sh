#!/bin/bash
# Same-box A/B of library variants (boxes differ by a few %): scripts/ab_run.sh "16 300" base varA varB ...   ("base" = the in-tree build)
ARGS=${1:-"16 300"}; shift
cp tortoise_tts_b200/lib/libtts_ar.so /tmp/base.so
run() { timeout 300 python scripts/stage_prof.py $ARGS 2>&1 | grep -E "^N=" ; }
for round in 1 2; do
  for v in "$@"; do
    if [ "$v" = "base" ]; then cp /tmp/base.so tortoise_tts_b200/lib/libtts_ar.so; else cp ab/libtts_ar_$v.so tortoise_tts_b200/lib/libtts_ar.so; fi
    touch tortoise_tts_b200/lib/libtts_ar.so
    echo "--- $v (round $round)"; run
  done
done
cp /tmp/base.so tortoise_tts_b200/lib/libtts_ar.so; touch tortoise_tts_b200/lib/libtts_ar.so
