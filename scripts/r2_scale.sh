#!/bin/bash
# usage: bash scripts/r2_scale.sh N   (inside gpurun --gpus N)
N=$1
mkdir -p gpurun_out
if [ "$N" = "2" ]; then timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log; fi
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
tail -c 3500 gpurun_out/bench_n$N.json; tail -3 gpurun_out/bench_n$N.err
