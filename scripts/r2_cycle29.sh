#!/bin/bash
mkdir -p gpurun_out
bash scripts/ab_run.sh "16 300" base slack1 slack2 slack4 slack16 2>&1 | grep -E "^---|^N=" > gpurun_out/ab29.log; cat gpurun_out/ab29.log
