#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/clvp_prof.py 1 > gpurun_out/clvp_plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_clvp_launches.csv python scripts/clvp_prof.py 1 > gpurun_out/clvp_ncu.log 2>&1
tail -2 gpurun_out/clvp_plain.log
