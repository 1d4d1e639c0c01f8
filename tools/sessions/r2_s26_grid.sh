#!/bin/bash
# round 2, GPU session 26 (1 GPU): SPT_FLAG_GRID (opt-in acceleration structure): parity (rays against the strict loop and the
# oracle, renders against the oracle), then configs 5 / 4 / 3 with and without it
set -u
o=gpurun_out/s26; mkdir -p $o
timeout 900 python -m pytest tests -m gpu -x -q -k "grid" > $o/pytest_grid.log 2>&1; echo "pytest rc=$?"; tail -15 $o/pytest_grid.log
for w in c5 c4; do for fl in 512; do
  extra=""; [ $w = c3 ] && extra="--spp 256"
  timeout 400 python bench.py --workload $w --steps 2 --warmup 1 --no-cpu --no-extra --flags $fl $extra > $o/${w}_f$fl.json 2>$o/${w}_f$fl.err
  python - $w $fl $o/${w}_f$fl.json <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[3]).read().strip().splitlines()[-1]); r=d['roofline']; k=r['kernels'].get('trace_secondary',{})
    print('%-4s flags %-4s value %9.1f Msamples/s ms %9.2f e2e %9.1f | bounce %8.1f ms camera %7.1f ms | exact evals/segment %.2f' % (sys.argv[1], sys.argv[2], d['value'], d['ms_per_step'], d['e2e']['value'], k.get('ms_per_render',0), r['kernels'].get('primary_hits_kernel',{}).get('ms_per_render',0), r['exact_evals_per_segment']))
except Exception as e:
    print(sys.argv[1], sys.argv[2], 'FAILED', e); print(open(sys.argv[3].replace('.json','.err')).read()[-1500:])
P
done; done
