#!/bin/bash
# round 2, GPU session 20: does the fused kernel's instruction footprint explain its loss?  fusedB.so = experiment build whose camera
# phase is the listed path only, one sample per trip (82 KB of SASS instead of 120 KB; valid on config 3, where no list overflows)
set -u
o=gpurun_out/s20; mkdir -p $o
for a in "fusedB:" "fusedB_c64:--split-chunk 64" "fusedB_two:--flags 512"; do
  n=${a%%:*}; f=${a#*:}
  SPT_CUDA_LIB=$PWD/build_ab/fusedB.so timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu --no-extra $f > $o/c3_$n.json 2>$o/c3_$n.err
  python - $n $o/c3_$n.json <<'P'
import json,sys
d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); print('%-14s value %8.1f ms %8.2f step frac %.4f %s'%(sys.argv[1],d['value'],d['ms_per_step'],d['roofline']['step']['frac'],d['roofline']['step']['launches'][:30]))
P
done
