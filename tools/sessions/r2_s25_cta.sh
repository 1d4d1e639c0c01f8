#!/bin/bash
# round 2, GPU session 25 (1 GPU): two-warp CTAs (8x8 pixels, 16 per SM) for the resident split kernels against four-warp CTAs (base)
set -u
o=gpurun_out/s25; mkdir -p $o
line() { python - "$1" "$2" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d['roofline']; k=r['kernels']['trace_secondary']
    print('%-22s value %8.1f ms %8.2f e2e %8.1f | sec %7.1f ms frac %.4f lanes %.4f | pri %6.1f ms | step frac %.4f' % (sys.argv[1], d['value'], d['ms_per_step'], d['e2e']['value'], k['ms_per_render'], k['frac'], k.get('lanes_with_a_ray',0), r['kernels']['primary_hits_kernel']['ms_per_render'], r['step']['frac']))
except Exception as e:
    print(sys.argv[1], 'FAILED', e); print(open(sys.argv[2].replace('.json','.err')).read()[-800:])
P
}
for v in base cta2 base cta2; do
  SPT_CUDA_LIB=$PWD/build_ab/$v.so timeout 300 python bench.py --steps 2 --warmup 2 --no-cpu --no-extra > $o/c3_$v.json 2>$o/c3_$v.err
  line c3_$v $o/c3_$v.json
done
