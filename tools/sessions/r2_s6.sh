#!/bin/bash
# round 2, GPU session 6b (after the machine was replaced): the no-allocation entry point alone in a fresh process, then the
# whole GPU suite (it runs last), then the default bench line with the sample-first work-item cut
set -u
o=gpurun_out/s6; mkdir -p $o
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $o/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?" | tee -a $o/pytest_noalloc_alone.log; tail -4 $o/pytest_noalloc_alone.log
timeout 900 python -m pytest tests -m gpu -x -q > $o/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $o/pytest.log; tail -6 $o/pytest.log
( time timeout 900 python bench.py > $o/bench_default.json 2>$o/bench_default.err ) 2>&1 | tail -3
python - <<P
import json
try:
    d=json.loads(open('$o/bench_default.json').read().strip().splitlines()[-1])
    r=d['roofline']
    print('default: value %.1f ms %.2f e2e %.1f frac %.4f step %.4f %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], r['frac'], r['step']['frac'], r['step']['launches']))
    print({n:(round(x['ms_per_render'],1), round(x['frac'],4), round(x.get('lanes_with_a_ray',0),4)) for n,x in r['kernels'].items()})
    print('cpu', d['cpu_baseline']['value'], d['cpu_baseline']['cores'], d['cpu_baseline'].get('g_maxThreads_4'))
    for k,v in d['workloads'].items():
        print(k, 'value %.1f ms %.1f step frac %.4f sec frac %.4f' % (v['value'], v['ms_per_render'], v['roofline_frac_step'], v['roofline_frac'] or 0), {n:(round(x['ms_per_render'],1), round(x.get('lanes_with_a_ray',0),4)) for n,x in v['kernels'].items()}, v['scene_setup_ms'])
except Exception as e:
    print('default FAILED', e); print(open('$o/bench_default.err').read()[-2000:])
P
timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu --no-extra --staging-mib 65536 > $o/bench_st65536.json 2>$o/bench_st65536.err
python - <<P
import json
d=json.loads(open('$o/bench_st65536.json').read().strip().splitlines()[-1]); r=d['roofline']
print('64GiB: value %.1f ms %.2f e2e %.1f frac %.4f step %.4f %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], r['frac'], r['step']['frac'], r['step']['launches']))
P
