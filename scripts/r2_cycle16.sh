#!/bin/bash
mkdir -p gpurun_out
for n in "16 500" "32 300" "64 300"; do bash scripts/ab_run.sh "$n" base d3 d5 2>&1 | grep -E "^---|^N=" ; done > gpurun_out/ab_depth.log 2>&1; cat gpurun_out/ab_depth.log
