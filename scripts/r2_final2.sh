#!/bin/bash
# final cycle of the round: every GPU test, smoke, both bench arms, stage profile, then the ncu passes (launch list, --set full of the decode kernel)
bash scripts/r2_final.sh
TTS_AR_PROF_STEP=200 timeout 300 python scripts/stage_prof.py 16 300 > gpurun_out/stage_prof_16.log 2>&1; head -4 gpurun_out/stage_prof_16.log
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-sub-configs"
timeout 600 $B > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches.csv $B > gpurun_out/ncu_list.log 2>&1
tail -2 gpurun_out/ncu_list.log
export TTS_BENCH_MAX_NEW=120
timeout 600 $B > gpurun_out/plain2.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:ar_decode_kernel -c 1 -o gpurun_out/r02_decode $B > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
ls -la gpurun_out/*.ncu-rep
