#!/bin/bash
# GPU box cycle: parity tests, short decode timing (cluster kernel and interpreter), optional full bench.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
TTS_AR_PROF_STEP=200 timeout 300 python scripts/stage_prof.py 16 300 > gpurun_out/stage_prof.log 2>&1; cat gpurun_out/stage_prof.log | tail -40
if [ "$1" = "bench" ]; then timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench1.log 2>&1; cat gpurun_out/bench1.log; fi
