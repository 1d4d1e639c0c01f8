#!/bin/bash
# First GPU contact of the EXPERIMENTAL tensor-core bounce kernel (spt_tensor_filter.cuh).
#   here (no GPU needed):  tools/build_variant.sh build_ab/base.so
#                          tools/build_variant.sh build_ab/tensor.so  -DSPT_TENSOR_FILTER=1
#                          tools/build_variant.sh build_ab/tensor1.so -DSPT_TENSOR_FILTER=1 -DSPT_TENSOR_TILES=1
#   then:                  gpurun --timeout 900 -- tools/tensor_filter_gpu.sh
# 1. the parity suite against the variant (bit-exact accumulators vs the oracle on every resident split render), under a
#    timeout of its own because a barrier mistake in new code shows up as a hang;  2. only if green, same-box A/B on configs 3 / 4.
set -u
mkdir -p gpurun_out
for lib in build_ab/tensor.so build_ab/tensor1.so; do
  [ -f "$lib" ] || continue
  log=gpurun_out/pytest_$(basename "$lib" .so).log
  SPT_CUDA_LIB=$PWD/$lib timeout 420 python -m pytest tests -m gpu -x -q > "$log" 2>&1
  rc=$?; echo "pytest rc=$rc ($lib)" | tee -a "$log"; tail -3 "$log"
  [ $rc -eq 0 ] || exit $rc
done
libs="build_ab/base.so build_ab/tensor.so"; [ -f build_ab/tensor1.so ] && libs="$libs build_ab/tensor1.so"
tools/ab_run.sh c3 gpurun_out/ab_tensor_c3.txt $libs $libs
tools/ab_run.sh c4 gpurun_out/ab_tensor_c4.txt $libs
