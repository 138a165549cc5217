#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_clvp.py tests/test_gpu_cond.py -m gpu -x -q 2>&1 | tail -3
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "dense or reforward or golden or tc_gemm or patch" 2>&1 | tail -3
timeout 300 python scripts/clvp_prof.py 2 2>&1 | tail -1
timeout 600 python scripts/config_sweep.py C2 C5_share 2>&1 | grep -o '"case": "[A-Z0-9_a-z]*"\|"reforward_ms": [0-9.]*\|"reforward_tflops": [0-9.]*\|"prefill_ms": [0-9.]*'
