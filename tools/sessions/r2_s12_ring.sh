#!/bin/bash
# round 2, GPU session 12 (1 GPU): the chunk ring refilled by the last warp to release a buffer (no "empty" barrier, no elected producer
# running ahead), ring depth 2 / 3 / 4, against the previous build (base.so = commit before): parity of the streamed kernels, config 5
set -u
o=gpurun_out/s12; mkdir -p $o
timeout 900 python -m pytest tests -m gpu -x -q -k "streamed or lattice or config5 or shared_inside or full_size or pixel_cone or row_bands" > $o/pytest_ring.log 2>&1; echo "pytest rc=$?" | tee -a $o/pytest_ring.log; tail -4 $o/pytest_ring.log
line() { python - "$1" "$2" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d['roofline']; k=r['kernels']['trace_secondary']
    print('%-22s value %8.1f ms %8.2f e2e %8.1f | sec %7.1f ms frac %.4f lanes %.4f | pri %6.1f ms | step frac %.4f' % (sys.argv[1], d['value'], d['ms_per_step'], d['e2e']['value'], k['ms_per_render'], k['frac'], k.get('lanes_with_a_ray',0), r['kernels']['primary_hits_kernel']['ms_per_render'], r['step']['frac']))
except Exception as e:
    print(sys.argv[1], 'FAILED', e)
P
}
for v in base ring3 ring3b ring4b ring2b ring3b; do
  SPT_CUDA_LIB=$PWD/build_ab/$v.so timeout 300 python bench.py --workload c5 --steps 2 --warmup 2 --no-cpu --no-extra > $o/c5_$v.json 2>$o/c5_$v.err
  line c5_$v $o/c5_$v.json
done
cmd="python bench.py --workload c5 --spp 8 --steps 1 --warmup 0 --no-cpu --no-extra"
$cmd > $o/plain_c5.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'trace_secondary' -c 1 -f -o $o/r2_c5_secondary_ring3b $cmd > $o/ncu_c5.log 2>&1
echo "ncu c5 rc=$?"; tail -2 $o/ncu_c5.log
