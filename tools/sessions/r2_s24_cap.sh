#!/bin/bash
# round 2, GPU session 24 (1 GPU): capacity of the pixel-cone candidate lists (8 entries per lane; 12; 16), configs 3 and 4
set -u
o=gpurun_out/s24; mkdir -p $o
line() { python - "$1" "$2" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d['roofline']; k=r['kernels']['trace_secondary']
    print('%-22s value %8.1f ms %8.2f e2e %8.1f | sec %7.1f ms frac %.4f lanes %.4f | pri %6.1f ms | step frac %.4f' % (sys.argv[1], d['value'], d['ms_per_step'], d['e2e']['value'], k['ms_per_render'], k['frac'], k.get('lanes_with_a_ray',0), r['kernels']['primary_hits_kernel']['ms_per_render'], r['step']['frac']))
except Exception as e:
    print(sys.argv[1], 'FAILED', e)
P
}
for v in base cap12 cap16 base; do
  SPT_CUDA_LIB=$PWD/build_ab/$v.so timeout 300 python bench.py --steps 2 --warmup 2 --no-cpu --no-extra > $o/c3_$v.json 2>$o/c3_$v.err
  line c3_$v $o/c3_$v.json
done
for v in base cap12 cap16; do
  SPT_CUDA_LIB=$PWD/build_ab/$v.so timeout 300 python bench.py --workload c4 --steps 2 --warmup 2 --no-cpu --no-extra > $o/c4_$v.json 2>$o/c4_$v.err
  line c4_$v $o/c4_$v.json
done
