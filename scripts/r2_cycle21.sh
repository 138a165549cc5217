#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_cond.py tests/test_gpu_clvp.py -m gpu -x -q -s > gpurun_out/pytest_cond.log 2>&1; tail -25 gpurun_out/pytest_cond.log
timeout 900 python bench.py --steps 2 --warmup 1 > gpurun_out/bench_clvp.log 2>gpurun_out/bench_clvp.err; tail -c 3000 gpurun_out/bench_clvp.log; tail -5 gpurun_out/bench_clvp.err
