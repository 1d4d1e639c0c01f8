#!/bin/bash
# round 2, GPU session 4: parity with the pixel-cone camera kernel, then same-box A/B: trip counter on/off, cone on/off
set -u
o=gpurun_out/s4; mkdir -p $o
timeout 900 python -m pytest tests -m gpu -x -q > $o/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $o/pytest.log; tail -8 $o/pytest.log
run() { # lib staging_mib tag
  SPT_CUDA_LIB=$PWD/build_ab/$1.so timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu --no-extra --staging-mib $2 > $o/b_$3.json 2>$o/b_$3.err
  python - <<P
import json
try:
    d=json.loads(open('$o/b_$3.json').read().strip().splitlines()[-1])
    r=d['roofline']; k=r['kernels']
    print('%-28s value %7.1f ms %8.2f e2e %7.1f | sec %6.1f ms frac %.4f lanes %.4f | pri %6.1f ms | step frac %.4f serial %.1f' % ('$3', d['value'], d['ms_per_step'], d['e2e']['value'],
      k['trace_secondary']['ms_per_render'], k['trace_secondary']['frac'], k['trace_secondary'].get('lanes_with_a_ray',0), k['primary_hits_kernel']['ms_per_render'], r['step']['frac'], r['step']['serialised_chunks_ms']))
except Exception as e:
    print('$3 FAILED', e); print(open('$o/b_$3.err').read()[-1500:])
P
}
for st in 0 65536; do
  run r2_base $st base_st$st
  run r3_trips $st trips_st$st
  run r3_notrips $st notrips_st$st
  run r4_cone $st cone_st$st
  run r4_cone_notrips $st cone_notrips_st$st
done
( time timeout 900 python bench.py > $o/bench_default.json 2>$o/bench_default.err ) 2>&1 | tail -3
python - <<P
import json
d=json.loads(open('$o/bench_default.json').read().strip().splitlines()[-1])
print('default: value', d['value'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'], 'step', d['roofline']['step']['frac'])
for k,v in d['workloads'].items():
    print(k, 'value %.1f ms %.1f step frac %.4f sec frac %.4f' % (v['value'], v['ms_per_render'], v['roofline_frac_step'], v['roofline_frac'] or 0), {n:(round(x['ms_per_render'],1), round(x.get('lanes_with_a_ray',0),4)) for n,x in v['kernels'].items()})
P
