#!/bin/bash
# attn.c_proj contributions as one bulk reduction per row (cp.reduce.async.bulk .add.u64) instead of 64-bit red.global per lane: parity, same-box A/B
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_parity.log 2>&1; tail -3 gpurun_out/pytest_parity.log
bash scripts/ab_run.sh "16 300" base pre > gpurun_out/ab_proj_bulk_red.log 2>&1; cat gpurun_out/ab_proj_bulk_red.log
bash scripts/ab_run.sh "32 300" base pre > gpurun_out/ab_proj_bulk_red_32.log 2>&1; cat gpurun_out/ab_proj_bulk_red_32.log
TTS_AR_PROF_STEP=200 timeout 300 python scripts/stage_prof.py 16 300 > gpurun_out/stage_prof_16_bulkred.log 2>&1
grep -A32 "marks of layer 3" gpurun_out/stage_prof_16_bulkred.log | grep -E "attn|post-bar|pre-bar|ln-loaded"
