#!/bin/bash
# MODE 4 attn.c_proj contributions as bulk reductions from a staging tile in the freed ring slots: parity, same-box A/B at 64 / 128 rows
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_parity.log 2>&1; tail -3 gpurun_out/pytest_parity.log
bash scripts/ab_run.sh "64 300" base pre > gpurun_out/ab_proj_bulk_red_64.log 2>&1; cat gpurun_out/ab_proj_bulk_red_64.log
bash scripts/ab_run.sh "48 300" base pre > gpurun_out/ab_proj_bulk_red_48.log 2>&1; cat gpurun_out/ab_proj_bulk_red_48.log
bash scripts/ab_run.sh "16 300" base pre > gpurun_out/ab_proj_bulk_red_16b.log 2>&1; cat gpurun_out/ab_proj_bulk_red_16b.log
