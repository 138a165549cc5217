#!/bin/bash
# Builds compile-time variants of the library into ab/ (git-ignored, shipped by gpurun) for same-box A/B timing:
#   scripts/ab_build.sh name "-DFLAG1 -DFLAG2"
set -e
cd "$(dirname "$0")/.."
mkdir -p ab
S=tortoise_tts_b200/csrc
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC $2 -o ab/libtts_ar_$1.so \
  $S/ar_kernels.cu $S/ar_decode.cu $S/dense_pass.cu $S/clvp.cu $S/cond_enc.cu $S/tts_ar.cu $S/tts_ar_run.cu
echo built ab/libtts_ar_$1.so
