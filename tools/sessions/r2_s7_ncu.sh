#!/bin/bash
# round 2, GPU session 7: ncu on the final build -- one --set full capture of primary_hits_kernel, trace_secondary and
# resolve_kernel (config 3 at 32 spp = one work item of the default 8 GiB cut), and the launch list of the default bench command
set -u
o=gpurun_out/s7; mkdir -p $o
cmd1="python bench.py --spp 32 --steps 1 --warmup 0 --no-cpu --no-extra"
$cmd1 > $o/plain1.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'primary_hits_kernel|trace_secondary|resolve_kernel' -c 3 -f -o $o/r2_final $cmd1 > $o/ncu1.log 2>&1
echo "ncu full rc=$?"; tail -3 $o/ncu1.log
cmd2="python bench.py --steps 2 --warmup 1 --no-cpu --no-extra"
$cmd2 > $o/plain2.log 2>&1 && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/r2_launches_c3.csv $cmd2 > $o/ncu2.log 2>&1
echo "ncu launches rc=$?"; tail -1 $o/plain2.log | cut -c1-300
