#!/bin/bash
# round 2, GPU session 1: baseline green check, tensor-filter decision (A/B + ncu), stall/pipe metrics of trace_secondary
set -u
mkdir -p gpurun_out/s1
o=gpurun_out/s1
timeout 400 python -m pytest tests -m gpu -x -q > $o/pytest_base.log 2>&1; echo "pytest base rc=$?" | tee -a $o/pytest_base.log; tail -2 $o/pytest_base.log
good=
for lib in build_ab/tensor.so build_ab/tensorm.so build_ab/tensor42.so build_ab/tensor42m.so build_ab/tensor21.so; do
  [ -f "$lib" ] || continue
  out=$o/ab_$(basename "$lib" .so)_c3.txt
  timeout 90 build_ab/tensor_filter_ab "$lib" build_ab/c3_scene.bin 3840 2160 64 50 3 > "$out" 2>&1
  rc=$?; echo "rc=$rc" >> "$out"; echo "== $lib"; tail -9 "$out"
  [ $rc -eq 0 ] && good="$good $lib"
done
libs="build_ab/base.so$good"
tools/ab_run.sh c3 $o/ab_tensor_c3.txt $libs $libs
case "$good" in *build_ab/tensor.so*) ;; *) echo "tensor.so not good: no ncu"; exit 0;; esac
for k in trace_secondary_tc trace_secondary; do
  timeout 600 ncu --set full --clock-control none --import-source on \
    -k regex:"^$k\$" -s 4 -c 1 -o $o/ncu_$k -f \
    build_ab/tensor_filter_ab build_ab/tensor.so build_ab/c3_scene.bin 1920 1080 32 50 1 > $o/ncu_$k.log 2>&1
  echo "ncu $k rc=$?"; tail -3 $o/ncu_$k.log
done
