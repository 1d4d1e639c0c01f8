#!/bin/bash
# round 2, GPU session 3: parity of the kernel changes (shell sampler local attempts, merged tail round, shared-reciprocal
# division, lanes walking several tile rows), then same-box A/B of each against the build before them
set -u
o=gpurun_out/s3; mkdir -p $o
timeout 900 python -m pytest tests -m gpu -x -q > $o/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $o/pytest.log; tail -8 $o/pytest.log
run() { # lib staging_mib lane_tiles tag
  SPT_CUDA_LIB=$PWD/build_ab/$1.so timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu --no-extra --staging-mib $2 --lane-tiles $3 > $o/b_$4.json 2>$o/b_$4.err
  python - <<P
import json
try:
    d=json.loads(open('$o/b_$4.json').read().strip().splitlines()[-1])
    r=d['roofline']; k=r['kernels']
    print('%-28s value %7.1f ms %8.2f e2e %7.1f | sec %6.1f ms frac %.4f lanes %.4f | pri %6.1f ms | step frac %.4f serial %.1f' % ('$4', d['value'], d['ms_per_step'], d['e2e']['value'],
      k['trace_secondary']['ms_per_render'], k['trace_secondary']['frac'], k['trace_secondary'].get('lanes_with_a_ray',0), k['primary_hits_kernel']['ms_per_render'], r['step']['frac'], r['step']['serialised_chunks_ms']))
except Exception as e:
    print('$4 FAILED', e); print(open('$o/b_$4.err').read()[-1500:])
P
}
for st in 0 65536; do
  run r2_base $st 0 base_st$st
  run r2_new $st 1 new_lt1_st$st
  run r2_new $st 0 new_auto_st$st
  run r2_new $st 2 new_lt2_st$st
  run r2_new $st 4 new_lt4_st$st
  run r2_new $st 8 new_lt8_st$st
  run r2_noshell $st 1 noshell_lt1_st$st
  run r2_shell1 $st 1 shell1_lt1_st$st
  run r2_nomerge $st 1 nomerge_lt1_st$st
  run r2_norcp $st 1 norcp_lt1_st$st
  run r2_base $st 0 base_again_st$st
done
