#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_n1.log 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; tail -c 6000 gpurun_out/bench_n1.log; tail -5 gpurun_out/bench_n1.err
