#!/bin/bash
# round 2, GPU session 27 (1 GPU): whole GPU suite of the final tree (grid tests, full-size whole-image equality incl. the grid), default
# bench line with the c4 / c5 legs linear and on the grid, smoke
set -u
o=gpurun_out/s27; mkdir -p $o
( time timeout 1500 python -m pytest tests -m gpu -x -q > $o/pytest.log 2>&1 ) 2>&1 | grep real; tail -4 $o/pytest.log
( time timeout 900 python bench.py > $o/bench_default.json 2>$o/bench_default.err ) 2>&1 | grep real
python - <<P
import json
try:
    d=json.loads(open('$o/bench_default.json').read().strip().splitlines()[-1])
    print('value %.1f e2e %.1f frac %.4f step %.4f cpu %.2f (%d cores)' % (d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['step']['frac'], d['cpu_baseline']['value'], d['cpu_baseline']['cores']))
    for k,v in d['workloads'].items(): print(k, 'value %.1f ms %.2f' % (v['value'], v['ms_per_render']), {n:round(x['ms_per_render'],1) for n,x in v['kernels'].items()})
except Exception as e:
    print('FAILED', e); print(open('$o/bench_default.err').read()[-2000:])
P
python __graft_entry__.py smoke > $o/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $o/smoke.log | cut -c1-200
