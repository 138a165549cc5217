#!/bin/bash
# back-off between polls of the residual stream / the counting accumulator (ln_1 / ln_2): 0 (base) vs 50 ns vs 200 ns
mkdir -p gpurun_out
bash scripts/ab_run.sh "16 300" base sleep50 sleep200 > gpurun_out/ab_poll_backoff.log 2>&1; cat gpurun_out/ab_poll_backoff.log
