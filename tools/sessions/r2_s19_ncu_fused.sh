#!/bin/bash
# round 2, GPU session 19: ncu --set full on trace_fused (config 3 at 32 spp)
set -u
o=gpurun_out/s19; mkdir -p $o
cmd="python bench.py --spp 32 --steps 1 --warmup 0 --no-cpu --no-extra"
$cmd > $o/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'trace_fused' -c 1 -f -o $o/r2_fused $cmd > $o/ncu.log 2>&1
echo "ncu rc=$?"; tail -2 $o/ncu.log
