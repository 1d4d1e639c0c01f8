#!/bin/bash
# round 2, GPU session 9 (1 GPU): samples shared inside the warp (default) against one lane = one pixel (SPT_FLAG_OWN_PIXELS = 256),
# same library, same box: parity first, then config 3 at 8 GiB and 64 GiB staging, configs 4 and 5
set -u
o=gpurun_out/s9; mkdir -p $o
timeout 900 python -m pytest tests -m gpu -x -q -k "shared_inside or split_render or staging_limit or row_bands or crops or config5_lattice or pass_cap or progressive_passes" > $o/pytest_share.log 2>&1; echo "pytest rc=$?" | tee -a $o/pytest_share.log; tail -5 $o/pytest_share.log
line() { python - "$1" "$2" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d['roofline']; k=r['kernels']['trace_secondary']
    print('%-22s value %8.1f ms %8.2f e2e %8.1f | sec %7.1f ms frac %.4f lanes %.4f | pri %6.1f ms | step frac %.4f' % (sys.argv[1], d['value'], d['ms_per_step'], d['e2e']['value'], k['ms_per_render'], k['frac'], k.get('lanes_with_a_ray',0), r['kernels']['primary_hits_kernel']['ms_per_render'], r['step']['frac']))
except Exception as e:
    print(sys.argv[1], 'FAILED', e)
P
}
for st in 0 65536; do
  for fl in 0 256; do
    timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu --no-extra --staging-mib $st --flags $fl > $o/c3_st${st}_f$fl.json 2>$o/c3_st${st}_f$fl.err
    line c3_st${st}_flags$fl $o/c3_st${st}_f$fl.json
  done
done
for w in c4 c5; do
  for fl in 0 256; do
    timeout 300 python bench.py --workload $w --steps 2 --warmup 2 --no-cpu --no-extra --flags $fl > $o/${w}_f$fl.json 2>$o/${w}_f$fl.err
    line ${w}_flags$fl $o/${w}_f$fl.json
  done
done
