#!/bin/bash
# round 2, GPU session 23: the default bench line of the final build (CPU baseline and the c4 / c5 legs included), ncu --set full on one
# launch of each kernel (config 3 at 32 spp = one work item of the 8 GiB cut) and the launch list of the bench command
set -u
o=gpurun_out/s23; mkdir -p $o
( time timeout 900 python bench.py > $o/bench_default.json 2>$o/bench_default.err ) 2>&1 | grep real
( time timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $o/bench_reference.json 2>$o/bench_reference.err ) 2>&1 | grep real
tail -c 600 $o/bench_reference.json
cmd1="python bench.py --spp 32 --steps 1 --warmup 0 --no-cpu --no-extra"
$cmd1 > $o/plain1.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'primary_hits_kernel|trace_secondary|resolve_kernel' -c 3 -f -o $o/r2_final3 $cmd1 > $o/ncu1.log 2>&1
echo "ncu full rc=$?"; tail -3 $o/ncu1.log
cmd2="python bench.py --steps 2 --warmup 1 --no-cpu --no-extra"
$cmd2 > $o/plain2.log 2>&1 && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/r2_launches_c3_final3.csv $cmd2 > $o/ncu2.log 2>&1
echo "ncu launches rc=$?"; tail -1 $o/plain2.log | cut -c1-300
python __graft_entry__.py smoke > $o/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $o/smoke.log
