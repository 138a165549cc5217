#!/bin/bash
# round 2, GPU call 1: tcgen05 GEMM bring-up, the whole parity suite, smoke, per-config timings.
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -q -k tc_gemm > gpurun_out/pytest_tc.log 2>&1; echo "tc rc=$?" >> gpurun_out/pytest_tc.log
tail -30 gpurun_out/pytest_tc.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -60 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
timeout 900 python scripts/config_sweep.py C2 C3_share C5_share > gpurun_out/sweep.log 2>&1; tail -5 gpurun_out/sweep.log
