#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "stream or patch_reference or host_buffer" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python - <<'PY' > gpurun_out/stream.log 2>&1
import json, sys, torch
sys.path.insert(0, '.')
import bench
from tortoise_tts_b200.synthetic import ARConfig, make_state_dict
cfg = ARConfig(); sd = make_state_dict(cfg, seed=0)
print(json.dumps(bench.streaming_latency(cfg, sd, torch.device('cuda', 0))))
PY
tail -3 gpurun_out/stream.log
