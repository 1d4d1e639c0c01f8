#!/bin/bash
# round 2, GPU session 10: ncu --set full on trace_secondary, samples shared inside the warp (default build of the session) against
# SPT_FLAG_OWN_PIXELS, config 3 at 32 spp (one work item of the 8 GiB cut)
set -u
o=gpurun_out/s10; mkdir -p $o
for fl in 0 256; do
  cmd="python bench.py --spp 32 --steps 1 --warmup 0 --no-cpu --no-extra --flags $fl"
  $cmd > $o/plain_f$fl.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'trace_secondary' -c 1 -f -o $o/r2_share_f$fl $cmd > $o/ncu_f$fl.log 2>&1
  echo "ncu f$fl rc=$?"; tail -2 $o/ncu_f$fl.log
done
