#!/bin/bash
# round 2, GPU session 17 (1 GPU): the wavefront kernel with its intersect stage on the TMA chunk ring (config 5), parity, and the
# wavefront / megakernel-family comparison on configs 3, 4, 5 with the final kernels
set -u
o=gpurun_out/s17; mkdir -p $o
timeout 900 python -m pytest tests -m gpu -x -q -k "wavefront or config5 or lattice or full_size" > $o/pytest_wf.log 2>&1; echo "pytest rc=$?" | tee -a $o/pytest_wf.log; tail -3 $o/pytest_wf.log
for w in c3 c4 c5; do
  steps=2; [ $w = c3 ] && extra="--spp 128" || extra=""
  timeout 600 python bench.py --workload $w --variant wavefront --steps $steps --warmup 1 --no-cpu --no-extra $extra > $o/${w}_wavefront.json 2>$o/${w}_wavefront.err
  timeout 600 python bench.py --workload $w --steps $steps --warmup 1 --no-cpu --no-extra $extra > $o/${w}_default.json 2>$o/${w}_default.err
  python - $o $w <<'P'
import json,sys
o,w=sys.argv[1:]
for v in ('wavefront','default'):
    try:
        d=json.loads(open(f'{o}/{w}_{v}.json').read().strip().splitlines()[-1])
        print('%-4s %-10s value %9.1f Msamples/s  ms %9.2f  step frac %.4f  %s' % (w, v, d['value'], d['ms_per_step'], d['roofline']['step']['frac'], d['config']['workload'][:60]))
    except Exception as e:
        print(w, v, 'FAILED', e); print(open(f'{o}/{w}_{v}.err').read()[-1500:])
P
done
