#!/bin/bash
# GPU session for the EXPERIMENTAL tensor-core bounce kernel (spt_tensor_filter.cuh).
#   here (no GPU needed):  tools/build_variant.sh build_ab/base.so
#                          tools/build_variant.sh build_ab/tensor.so  -DSPT_TENSOR_FILTER=1
#                          tools/build_variant.sh build_ab/tensor42.so -DSPT_TENSOR_FILTER=1 -DSPT_TENSOR_TILES=4 -DSPT_TENSOR_ACC=2
#                          tools/build_variant.sh build_ab/tensor21.so -DSPT_TENSOR_FILTER=1 -DSPT_TENSOR_TILES=2 -DSPT_TENSOR_ACC=1
#                          tools/build_variant.sh build_ab/tensorm.so   -DSPT_TENSOR_FILTER=1 -DSPT_TENSOR_MERGED_BARRIER=1
#                          tools/build_variant.sh build_ab/tensor42m.so -DSPT_TENSOR_FILTER=1 -DSPT_TENSOR_TILES=4 -DSPT_TENSOR_ACC=2 -DSPT_TENSOR_MERGED_BARRIER=1
#                            (time-shared accumulators, 32 / 24 warps per SM at 64 registers; one tile barrier less per iteration: NOT run yet -- the harness below is
#                             their first contact, each under its own timeout; a variant that fails is dropped, the rest go on)
#                          g++ -O2 -std=c++17 -I include -o build_ab/tensor_filter_ab tools/tensor_filter_ab.cpp -ldl
#                          python tools/export_scene.py build_ab/c3_scene.bin
#   then:                  gpurun --timeout 1200 -- tools/tensor_filter_gpu.sh
# 0. seconds: the C harness (bit-compare tensor kernel vs FFMA2 kernel from the same library, config 3's scene at 4K, timing);
# 1. the whole parity suite against the variant (accumulators bit-exact vs the oracle on every resident split render), under a
#    timeout of its own because a barrier mistake in new code shows up as a hang;
# 2. only if green: bench.py A/B on configs 3 / 4 on the same box;
# 3. one ncu --set full capture of trace_secondary_tc (and of trace_secondary for comparison) driven by the C harness, after
#    everything above has exited 0 without ncu.  Numbers printed under ncu are not measurements.
set -u
mkdir -p gpurun_out
good=
for lib in build_ab/tensor.so build_ab/tensorm.so build_ab/tensor42.so build_ab/tensor42m.so build_ab/tensor21.so; do
  [ -f "$lib" ] || continue
  out=gpurun_out/ab_$(basename "$lib" .so)_c3.txt
  timeout 60 build_ab/tensor_filter_ab "$lib" build_ab/c3_scene.bin 3840 2160 64 50 3 > "$out" 2>&1
  rc=$?; echo "rc=$rc" >> "$out"; tail -9 "$out"
  [ $rc -eq 0 ] && good="$good $lib"
done
[ -n "$good" ] || exit 1
for lib in $good; do
  log=gpurun_out/pytest_$(basename "$lib" .so).log
  SPT_CUDA_LIB=$PWD/$lib timeout 420 python -m pytest tests -m gpu -x -q > "$log" 2>&1
  rc=$?; echo "pytest rc=$rc ($lib)" | tee -a "$log"; tail -3 "$log"
  [ $rc -eq 0 ] || exit $rc
done
libs="build_ab/base.so$good"
tools/ab_run.sh c3 gpurun_out/ab_tensor_c3.txt $libs $libs
tools/ab_run.sh c4 gpurun_out/ab_tensor_c4.txt $libs
# one launch of each bounce kernel at 1920x1080, 32 spp (ncu replays a kernel ~40 times: keep it short)
for k in trace_secondary_tc trace_secondary; do
  env $([ $k = trace_secondary ] && echo SPT_TENSOR=0) timeout 600 ncu --set full --clock-control none --import-source on \
    -k regex:"^$k\$" -c 1 -o gpurun_out/ncu_$k -f \
    build_ab/tensor_filter_ab build_ab/tensor.so build_ab/c3_scene.bin 1920 1080 32 50 1 > gpurun_out/ncu_$k.log 2>&1
  echo "ncu $k rc=$?"
done
