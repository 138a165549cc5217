#!/bin/bash
# refill bursts per boundary (before ln_2 / grid barrier / before ln_1): 6/6/6 (base) vs 4/6/6, 2/8/6, 6/8/2, 4/8/4
mkdir -p gpurun_out
bash scripts/ab_run.sh "16 300" base v466 v286 v682 v484 > gpurun_out/ab_refill_schedule.log 2>&1; cat gpurun_out/ab_refill_schedule.log
