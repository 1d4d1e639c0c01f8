#!/bin/bash
# Build a variant of the CUDA library for A/B runs: tools/build_variant.sh build_ab/name.so [-DSPT_...=...] ...
# (same flags as simplepathtracer_b200.binding.NVCC_FLAGS; bench.py / the tests pick it up from SPT_CUDA_LIB)
out=$1; shift
cd "$(dirname "$0")/.."
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --fmad=false \
  -Xcompiler -fPIC -Xcompiler -ffp-contract=off -Xcompiler -fvisibility=hidden -shared "$@" \
  -I include simplepathtracer_b200/csrc/spt_cuda.cu simplepathtracer_b200/csrc/spt_host_c.cpp -o "$out"
