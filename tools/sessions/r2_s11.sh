#!/bin/bash
# round 2, GPU session 11 (1 GPU): compile-time variants on config 3 with shared samples (shell sampler's local attempts, CTA shape),
# ncu on the streamed bounce kernel (config 5), then the whole GPU suite, timed
set -u
o=gpurun_out/s11; mkdir -p $o
line() { python - "$1" "$2" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d['roofline']; k=r['kernels']['trace_secondary']
    print('%-22s value %8.1f ms %8.2f e2e %8.1f | sec %7.1f ms frac %.4f lanes %.4f | pri %6.1f ms | step frac %.4f' % (sys.argv[1], d['value'], d['ms_per_step'], d['e2e']['value'], k['ms_per_render'], k['frac'], k.get('lanes_with_a_ray',0), r['kernels']['primary_hits_kernel']['ms_per_render'], r['step']['frac']))
except Exception as e:
    print(sys.argv[1], 'FAILED', e)
P
}
for v in base shell3 shell1 cta8 base; do
  SPT_CUDA_LIB=$PWD/build_ab/$v.so timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu --no-extra > $o/c3_$v.json 2>$o/c3_$v.err
  line c3_$v $o/c3_$v.json
done
for v in base cta8; do
  SPT_CUDA_LIB=$PWD/build_ab/$v.so timeout 300 python bench.py --workload c4 --steps 2 --warmup 2 --no-cpu --no-extra > $o/c4_$v.json 2>$o/c4_$v.err
  line c4_$v $o/c4_$v.json
done
cmd="python bench.py --workload c5 --spp 8 --steps 1 --warmup 0 --no-cpu --no-extra"
$cmd > $o/plain_c5.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'trace_secondary' -c 1 -f -o $o/r2_c5_secondary $cmd > $o/ncu_c5.log 2>&1
echo "ncu c5 rc=$?"; tail -2 $o/ncu_c5.log
( time timeout 1500 python -m pytest tests -m gpu -x -q > $o/pytest.log 2>&1 ) 2>&1 | grep real; tail -4 $o/pytest.log
