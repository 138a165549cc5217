#!/bin/bash
# LayerNorm gain staging overlapped with the row polls (no second CTA barrier at a soft boundary) + refills issued by warp 7: parity, same-box A/B
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_parity.log 2>&1; tail -3 gpurun_out/pytest_parity.log
bash scripts/ab_run.sh "16 300" base pre feed0 feed7nolap > gpurun_out/ab_stage_overlap.log 2>&1; cat gpurun_out/ab_stage_overlap.log
bash scripts/ab_run.sh "32 300" base pre > gpurun_out/ab_stage_overlap_32.log 2>&1; cat gpurun_out/ab_stage_overlap_32.log
bash scripts/ab_run.sh "64 300" base pre > gpurun_out/ab_stage_overlap_64.log 2>&1; cat gpurun_out/ab_stage_overlap_64.log
