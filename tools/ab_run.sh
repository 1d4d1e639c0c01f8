#!/bin/bash
# A/B on one box: tools/ab_run.sh <workload> <out.txt> lib.so[:cap_mib] ...   (cap_mib = bench.py --staging-mib, optional)
w=$1; out=$2; shift 2
mkdir -p "$(dirname "$out")"; : > "$out"
for spec in "$@"; do
  lib=${spec%%:*}; cap=${spec#*:}; [ "$cap" = "$spec" ] && cap=
  env SPT_CUDA_LIB=$PWD/$lib timeout 300 python bench.py --workload $w --steps 2 --warmup 2 --no-cpu ${cap:+--staging-mib $cap} 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$spec', '$w', round(d['value'], 1), round(d['ms_per_step'], 2), round(d['roofline']['frac'], 4), d['roofline']['kernel'], d['clocks']['sm_mhz'], d['clocks']['reasons'])" >> "$out" 2>&1
done
cat "$out"
