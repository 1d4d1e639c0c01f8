#!/bin/bash
# round 2, GPU session 18 (1 GPU): ONE persistent kernel (trace_fused: camera phase + bounce phase per warp tile and chunk over an
# L2-resident scratch; default) against the two-kernel path (flags 512), same library, same box; chunk lengths of the fused kernel
set -u
o=gpurun_out/s18; mkdir -p $o
( time timeout 1500 python -m pytest tests -m gpu -x -q > $o/pytest.log 2>&1 ) 2>&1 | grep real; tail -5 $o/pytest.log
line() { python - "$1" "$2" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d['roofline']; k=r['kernels'].get('trace_secondary',{})
    print('%-26s value %8.1f ms %8.2f e2e %8.1f | step frac %.4f | two-kernel path: sec %7.1f ms frac %.4f pri %6.1f ms | %s' % (sys.argv[1], d['value'], d['ms_per_step'], d['e2e']['value'], r['step']['frac'], k.get('ms_per_render',0), k.get('frac',0), r['kernels'].get('primary_hits_kernel',{}).get('ms_per_render',0), r['step']['launches'][:40]))
except Exception as e:
    print(sys.argv[1], 'FAILED', e); print(open(sys.argv[2].replace('.json','.err')).read()[-1200:])
P
}
for a in "fused_default:" "two_kernels:--flags 512" "fused_c16:--split-chunk 16" "fused_c64:--split-chunk 64" "fused_c128:--split-chunk 128" "fused_default2:" "two_kernels2:--flags 512"; do
  n=${a%%:*}; f=${a#*:}
  timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu --no-extra $f > $o/c3_$n.json 2>$o/c3_$n.err
  line c3_$n $o/c3_$n.json
done
for a in "fused_default:" "two_kernels:--flags 512"; do
  n=${a%%:*}; f=${a#*:}
  timeout 300 python bench.py --workload c4 --steps 2 --warmup 2 --no-cpu --no-extra $f > $o/c4_$n.json 2>$o/c4_$n.err
  line c4_$n $o/c4_$n.json
done
