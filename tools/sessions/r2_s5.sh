#!/bin/bash
# round 2, GPU session 5: parity with strips / Weyl stream key / shared-memory trip counter, then same-box A/B of the work-item cut
set -u
o=gpurun_out/s5; mkdir -p $o
timeout 900 python -m pytest tests -m gpu -x -q > $o/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $o/pytest.log; tail -8 $o/pytest.log
run() { # lib staging_mib tag extra...
  lib=$1; st=$2; tag=$3; shift 3
  SPT_CUDA_LIB=$PWD/build_ab/$lib.so timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu --no-extra --staging-mib $st "$@" > $o/b_$tag.json 2>$o/b_$tag.err
  python - <<P
import json
try:
    d=json.loads(open('$o/b_$tag.json').read().strip().splitlines()[-1])
    r=d['roofline']; k=r['kernels']
    print('%-30s value %7.1f ms %8.2f e2e %7.1f | sec %6.1f ms frac %.4f lanes %.4f | pri %6.1f ms | step frac %.4f serial %.1f items %s' % ('$tag', d['value'], d['ms_per_step'], d['e2e']['value'],
      k['trace_secondary']['ms_per_render'], k['trace_secondary']['frac'], k['trace_secondary'].get('lanes_with_a_ray',0), k['primary_hits_kernel']['ms_per_render'], r['step']['frac'], r['step']['serialised_chunks_ms'], r['step']['launches'].split()[0]))
except Exception as e:
    print('$tag FAILED', e); print(open('$o/b_$tag.err').read()[-1500:])
P
}
run r4_cone_notrips 0 r4_st0
run r5_new 0 new_st0
run r5_notrips 0 notrips_st0
run r5_new 0 new_st0_s270c32 --strip-rows 270 --split-chunk 32
run r5_new 0 new_st0_s68c128 --strip-rows 68 --split-chunk 128
run r5_new 0 new_st0_s34c256 --strip-rows 34 --split-chunk 256
run r5_new 0 new_st0_s17c512 --strip-rows 17 --split-chunk 512
run r5_new 0 new_st0_s5c1024 --strip-rows 5
run r4_cone_notrips 65536 r4_st65536
run r5_new 65536 new_st65536
run r5_new 65536 new_st65536_s270c256 --strip-rows 270 --split-chunk 256
run r5_new 65536 new_st65536_s135c512 --strip-rows 135 --split-chunk 512
run r5_new 65536 new_st65536_s34c1024 --strip-rows 34
run r5_new 16384 new_st16384
run r5_new 2048 new_st2048
( time timeout 900 python bench.py > $o/bench_default.json 2>$o/bench_default.err ) 2>&1 | tail -3
python - <<P
import json
d=json.loads(open('$o/bench_default.json').read().strip().splitlines()[-1])
print('default: value', d['value'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'], 'step', d['roofline']['step']['frac'])
for k,v in d['workloads'].items():
    print(k, 'value %.1f ms %.1f step frac %.4f sec frac %.4f' % (v['value'], v['ms_per_render'], v['roofline_frac_step'], v['roofline_frac'] or 0), {n:(round(x['ms_per_render'],1), round(x.get('lanes_with_a_ray',0),4)) for n,x in v['kernels'].items()})
P
