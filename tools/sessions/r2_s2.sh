#!/bin/bash
# round 2, GPU session 2: the whole GPU suite on the reworked host side (staging limit, two-stream chunks, multi-device
# context on one GPU), then the staging A/B and one complete default bench line
set -u
o=gpurun_out/s2; mkdir -p $o
timeout 900 python -m pytest tests -m gpu -x -q > $o/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $o/pytest.log; tail -15 $o/pytest.log
for mib in 0 4096 16384 65536 0 65536; do
  timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu --no-extra --staging-mib $mib > $o/bench_staging_$mib.json 2>$o/bench_staging_$mib.err
  python - <<P
import json
try:
    d=json.loads(open('$o/bench_staging_$mib.json').read().strip().splitlines()[-1])
    r=d['roofline']
    print('staging_mib=$mib value %.1f ms %.2f e2e %.1f frac_sec %.4f frac_step %.4f serial_ms %.2f kernels %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], r['frac'], r['step']['frac'], r['step']['serialised_chunks_ms'] or 0, {k:(round(v['ms_per_render'],1), round(v['frac'],4), round(v['executed_frac'],4)) for k,v in r['kernels'].items()}))
except Exception as e:
    print('staging_mib=$mib FAILED', e); print(open('$o/bench_staging_$mib.err').read()[-2000:])
P
done
( time timeout 900 python bench.py > $o/bench_default.json 2>$o/bench_default.err ) 2>&1 | tail -3
tail -c 3000 $o/bench_default.json
