#!/bin/bash
# A/B of library builds on one box: tools/ab_bench.sh <workload> <steps> lib1.so lib2.so ...
# prints value / ms_per_step / kernel ms per library (bench.py picks the library up from SPT_CUDA_LIB)
w=$1; steps=$2; shift 2
for lib in "$@"; do
  SPT_CUDA_LIB=$PWD/$lib python bench.py --workload $w --steps $steps --warmup 3 --no-cpu 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$lib', '$w', round(d['value'], 1), round(d['ms_per_step'], 2), round(d['roofline']['frac'], 4), d['clocks']['sm_mhz'], d['clocks']['reasons'])"
done
