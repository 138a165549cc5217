#!/bin/bash
mkdir -p gpurun_out
bash scripts/ab_run.sh "16 300" base pfpage nopf > gpurun_out/ab_pf16.log 2>&1; cat gpurun_out/ab_pf16.log
bash scripts/ab_run.sh "64 300" base pfpage nopf > gpurun_out/ab_pf64.log 2>&1; cat gpurun_out/ab_pf64.log
