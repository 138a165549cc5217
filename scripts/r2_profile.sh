#!/bin/bash
# round 2 profiles: launch list of a bench run, --set full of the decode kernel and of the tcgen05 GEMM (all ncu runs of a
# call count as one tool).
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-sub-configs"
timeout 600 $B > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches.csv $B > gpurun_out/ncu_list.log 2>&1
tail -2 gpurun_out/ncu_list.log
export TTS_BENCH_MAX_NEW=120
timeout 600 $B > gpurun_out/plain2.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:ar_decode_kernel -c 1 -o gpurun_out/r02_decode $B > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
unset TTS_BENCH_MAX_NEW
# tcgen05 GEMMs of the latent re-forward (9104 rows): 2 prompt passes x 121 GEMMs come first, then layer 10 of the re-forward
S="python scripts/config_sweep.py C2"
timeout 600 $S > gpurun_out/plain3.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:tc_gemm_kernel -s 282 -c 4 -o gpurun_out/r02_tcgemm $S > gpurun_out/ncu_tc.log 2>&1
tail -2 gpurun_out/ncu_tc.log
ls -la gpurun_out/*.ncu-rep
